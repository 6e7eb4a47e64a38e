#!/bin/bash
# round-2 session R (1 GPU): what the driver runs at round end -- smoke(), the default bench line, the reference arm --
# with wall times.
tag=${1:-r4r}
out=gpurun_out; mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee $out/${tag}_pytest.log
( time python -c "import __graft_entry__ as g; g.smoke()" ) 2>&1 | tail -5 | tee $out/${tag}_smoke.log
( time python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err ) 2>&1 | tail -3 | tee $out/${tag}_bench_time.log
tail -3 $out/${tag}_bench.err
( time python bench.py --impl reference > $out/${tag}_ref.json 2> $out/${tag}_ref.err ) 2>&1 | tail -3 | tee $out/${tag}_ref_time.log
python - <<PY
import json
d=json.load(open("$out/${tag}_bench.json"))
print({k: d.get(k) for k in ("metric","value","unit","ms_per_step","steps","warmup","gpu_launches","vs_baseline")})
print("e2e", d["e2e"]); print("roofline", {k: d["roofline"][k] for k in ("bound","achieved","peak","frac","traffic")})
print("cpu_baseline", d["cpu_baseline"]); print("parity", d.get("parity")); print("clocks", d.get("clocks"))
for k,v in (d.get("workloads") or {}).items(): print(k, {q: v.get(q) for q in ("value","unit","ms","roofline_frac","kernel","error")})
r=json.load(open("$out/${tag}_ref.json")); print("ref", {k: r.get(k) for k in ("impl","value","unit","cpu_baseline")})
PY
