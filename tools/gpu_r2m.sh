#!/bin/bash
# round-2 session M (1 GPU): box test only off the interior cells (enter_cell), packed B-spline weights (y, z), build_field
# two voxels per thread + resident grid.  GPU tests, bench with C4 / cubic, launch list, ncu --set full of the bench kernel,
# the fan-out kernel on C4 and build_field (with source counters).
tag=${1:-r4m}
out=gpurun_out; mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee $out/${tag}_pytest.log
( time timeout 1200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workloads c4_1024,c4_512,cubic,c5 > $out/${tag}_bench.json 2> $out/${tag}_bench.err ) 2>&1 | tail -3
tail -2 $out/${tag}_bench.err
python - <<PY
import json
d=json.load(open("$out/${tag}_bench.json"))
print({k: d[k] for k in ("value","ms_per_step")}, d["e2e"]["value"], d["roofline"]["frac"])
for k,v in (d.get("workloads") or {}).items(): print(k, {q: v.get(q) for q in ("value","unit","ms","roofline_frac","kernel","error")})
PY
PCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --workloads none --no-parity"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv $PCMD > $out/${tag}_ncu.log 2>&1
grep -i "build_field\|build_cells" $out/${tag}_launches.csv | cut -c1-260
timeout 900 ncu --set full --import-source on --clock-control none -k regex:trace_rays -s 3 -c 1 -o $out/${tag}_trace -f $PCMD > $out/${tag}_ncu_trace.log 2>&1; tail -1 $out/${tag}_ncu_trace.log
PHOTON_NO_FIELD_CACHE=1 timeout 600 ncu --set full --import-source on --clock-control none -k regex:build_field -c 1 -o $out/${tag}_build_field -f $PCMD > $out/${tag}_ncu_bf.log 2>&1; tail -1 $out/${tag}_ncu_bf.log
timeout 900 ncu --set full --import-source on --clock-control none -k regex:trace_rays -s 2 -c 1 -o $out/${tag}_c4 -f python tools/workload_probe.py c4_1024 > $out/${tag}_ncu_c4.log 2>&1; tail -1 $out/${tag}_ncu_c4.log
ls -la $out/*.ncu-rep
