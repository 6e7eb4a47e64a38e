#!/bin/bash
# Quick A/B of kernel variants on the GPU box: value (ray-steps/s) per environment setting.
out=gpurun_out; mkdir -p $out
tag=${1:-sweep}; shift
: > $out/${tag}.txt
run() {  # label, env...
  label=$1; shift
  v=$(env "$@" timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>$out/${tag}_err.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%.4g ray-steps/s  %.2f ms/step  frac %.3f' % (d['value'], d['ms_per_step'], d['roofline']['frac']))")
  echo "$label: $v" | tee -a $out/${tag}.txt
}
if [ $# -eq 0 ]; then set -- PHOTON_OCC=2 PHOTON_OCC=3 PHOTON_OCC=4; fi
for cfg in "$@"; do run "$cfg" $(echo $cfg | tr ',' ' '); done
