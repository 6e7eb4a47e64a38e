#!/bin/bash
tag=${1:-r4i}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4
PCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --workloads none"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $out/${tag}_launches.csv $PCMD > $out/${tag}_ncu_launches.log 2>&1
grep -E "build_|quad_|prefilter|extract" $out/${tag}_launches.csv | awk -F'","' '{print $5, $NF}' | head
timeout 900 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-e2e --workloads cubic > $out/${tag}_bench.json 2> $out/${tag}_bench.err
python - <<PY
import json
d=json.load(open("$out/${tag}_bench.json"))
print({k: d[k] for k in ("value","ms_per_step")}, d["roofline"]["frac"])
for k,v in (d.get("workloads") or {}).items(): print(k, {q: v.get(q) for q in ("value","unit","ms","roofline_frac","kernel","error")})
PY
PHOTON_INTERPOLATION=cubic timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $out/${tag}_launches_cubic.csv $PCMD > /dev/null 2>&1
grep -E "build_|quad_|prefilter|extract" $out/${tag}_launches_cubic.csv | awk -F'","' '{print $5, $NF}' | head
