#!/bin/bash
# round-2 session U (1 GPU): evidence at the final build -- GPU tests, default bench line, reference arm, launch list of the
# bench command and ncu --set full of its trace kernel (-> profiles/ncu_current.json)
tag=${1:-r5z}
out=gpurun_out; mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee $out/${tag}_pytest.log
( time python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err ) 2>&1 | tail -3
tail -2 $out/${tag}_bench.err
( time python bench.py --impl reference > $out/${tag}_ref.json 2> $out/${tag}_ref.err ) 2>&1 | tail -3
PCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --workloads none --no-parity"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv $PCMD > $out/${tag}_ncu.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:trace_rays -s 3 -c 1 -o $out/${tag}_trace -f $PCMD > $out/${tag}_ncu_trace.log 2>&1; tail -1 $out/${tag}_ncu_trace.log
ncu -i $out/${tag}_trace.ncu-rep --page details > $out/${tag}_trace_details.txt 2>&1
ncu -i $out/${tag}_trace.ncu-rep --page raw --csv > $out/${tag}_trace_raw.csv 2>&1
ncu -i $out/${tag}_trace.ncu-rep --page source --csv 2>&1 | gzip > $out/${tag}_trace_source.csv.gz
rm -f $out/${tag}_trace.ncu-rep
python - <<PY
import json
d=json.load(open("$out/${tag}_bench.json"))
print({k: d.get(k) for k in ("value","ms_per_step","gpu_launches")}, d["e2e"]["value"], d["roofline"]["frac"], d.get("parity",{}).get("dpos_px"))
for k,v in (d.get("workloads") or {}).items(): print(k, {q: v.get(q) for q in ("value","unit","ms","roofline_frac","kernel","error")})
PY
