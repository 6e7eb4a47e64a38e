#!/bin/bash
# Quick A/B of kernel variants on the GPU box: value (ray-steps/s) per environment setting.
out=gpurun_out; mkdir -p $out
tag=${1:-sweep}
: > $out/${tag}.txt
run() {  # label, env...
  label=$1; shift
  v=$(env "$@" timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>$out/${tag}_err.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%.4g ray-steps/s  %.2f ms/step  frac %.3f' % (d['value'], d['ms_per_step'], d['roofline']['frac']))")
  echo "$label: $v" | tee -a $out/${tag}.txt
}
run occ2 PHOTON_OCC=2
run occ3 PHOTON_OCC=3
run occ4 PHOTON_OCC=4
run occ3_grid8 PHOTON_OCC=3 PHOTON_GRID_MULT=24
run texture PHOTON_SAMPLER=texture
run philox PHOTON_RNG=philox
