#!/bin/bash
tag=${1:-r4h}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -6
timeout 900 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-e2e --workloads cubic,c4_1024,c1_piv > $out/${tag}_bench.json 2> $out/${tag}_bench.err
python - <<PY
import json
d=json.load(open("$out/${tag}_bench.json"))
print({k: d[k] for k in ("value","ms_per_step")}, d["roofline"]["frac"])
for k,v in (d.get("workloads") or {}).items(): print(k, {q: v.get(q) for q in ("value","unit","ms","roofline_frac","kernel","error")})
PY
tail -3 $out/${tag}_bench.err
PHOTON_INTERPOLATION=cubic timeout 900 ncu --set full --clock-control none --import-source on -k regex:trace_rays -s 3 -c 1 -o $out/${tag}_cubic -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --workloads none > $out/${tag}_ncu_cubic.log 2>&1
tail -1 $out/${tag}_ncu_cubic.log
