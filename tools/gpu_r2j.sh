#!/bin/bash
tag=${1:-r4j}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workloads c4_1024,c5 > $out/${tag}_bench.json 2> $out/${tag}_bench.err
python - <<PY
import json
d=json.load(open("$out/${tag}_bench.json"))
print({k: d[k] for k in ("value","ms_per_step")}, d["e2e"]["value"], d["roofline"]["frac"])
for k,v in (d.get("workloads") or {}).items(): print(k, {q: v.get(q) for q in ("value","unit","ms","roofline_frac","kernel","error")})
PY
timeout 600 python tools/parity_probe512.py cells "coherent:PHOTON_CELLS=0" > $out/${tag}_parity512.txt 2>&1; cat $out/${tag}_parity512.txt | cut -c1-330
PCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --workloads none"
PHOTON_NO_FIELD_CACHE=1 timeout 600 ncu --set full --clock-control none -k regex:build_field -c 1 -o $out/${tag}_build_field -f $PCMD > $out/${tag}_ncu_bf.log 2>&1
tail -1 $out/${tag}_ncu_bf.log
