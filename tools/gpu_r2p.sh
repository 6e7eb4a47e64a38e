#!/bin/bash
# round-2 session P (1 GPU): the two-phase kernel for scenes without a field (queue of rays that pass the first aperture),
# tree merge of the bilinear splat, float division in ray_sphere.  New tests first under their own timeout, then the suite,
# the PIV workloads with the queue on and off, the bench line with C1 / C4, ncu of the queued kernel.
tag=${1:-r4p}
out=gpurun_out; mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "queued" 2>&1 | tail -15 | tee $out/${tag}_queue_test.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee $out/${tag}_pytest.log
for q in 1 0; do echo "PHOTON_QUEUE=$q"; PHOTON_QUEUE=$q python tools/workload_probe.py piv piv_thin piv_diffuse piv_diffraction 2>&1 | tail -4; done | tee $out/${tag}_piv_workloads.txt
( time timeout 1200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workloads c1_piv,c4_1024 > $out/${tag}_bench.json 2> $out/${tag}_bench.err ) 2>&1 | tail -3
tail -2 $out/${tag}_bench.err
python - <<PY
import json
d=json.load(open("$out/${tag}_bench.json"))
print({k: d[k] for k in ("value","ms_per_step")}, d["e2e"]["value"], d["roofline"]["frac"], d.get("parity"))
for k,v in (d.get("workloads") or {}).items(): print(k, {q: v.get(q) for q in ("value","unit","ms","roofline_frac","kernel","error")})
PY
timeout 900 ncu --set full --import-source on --clock-control none -k regex:trace_nofield -s 1 -c 1 -o $out/${tag}_piv -f python tools/workload_probe.py piv > $out/${tag}_ncu_piv.log 2>&1; tail -1 $out/${tag}_ncu_piv.log
ncu -i $out/${tag}_piv.ncu-rep --page details > $out/${tag}_piv_details.txt 2>&1
ncu -i $out/${tag}_piv.ncu-rep --page raw --csv > $out/${tag}_piv_raw.csv 2>&1
ncu -i $out/${tag}_piv.ncu-rep --page source --csv 2>&1 | gzip > $out/${tag}_piv_source.csv.gz
rm -f $out/${tag}_piv.ncu-rep
ls -la $out | tail -12
