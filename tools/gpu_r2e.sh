#!/bin/bash
# round-2 session E: tests, full bench (workloads), launch list + ncu of the bench kernel, ncu of C4 with the ray order
tag=${1:-r4e}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -6
( time timeout 1500 python bench.py --steps 10 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err ) 2>&1 | tail -3
python - <<PY
import json
d=json.load(open("$out/${tag}_bench.json"))
print({k: d[k] for k in ("value","ms_per_step")}, d["e2e"]["value"], d["roofline"]["frac"])
print("parity", d.get("parity"))
for k,v in (d.get("workloads") or {}).items(): print(k, {q: v.get(q) for q in ("value","unit","ms","roofline_frac","kernel","error","wall_s_including_setup")})
PY
tail -5 $out/${tag}_bench.err
PCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --workloads none"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv $PCMD > $out/${tag}_ncu_launches.log 2>&1
grep -E "build_|trace_rays" $out/${tag}_launches.csv | cut -d, -f5,12- | head -12
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:trace_rays -s 3 -c 1 -o $out/${tag}_trace -f $PCMD > $out/${tag}_ncu_full.log 2>&1
tail -1 $out/${tag}_ncu_full.log
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:trace_rays -s 2 -c 1 -o $out/${tag}_c4 -f python tools/workload_probe.py c4_1024 > $out/${tag}_ncu_c4.log 2>&1
tail -1 $out/${tag}_ncu_c4.log
