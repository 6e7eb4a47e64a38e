#!/bin/bash
# round-2 session B: all GPU tests (no -x), bench with the parity leg
tag=${1:-r4b}
out=gpurun_out; mkdir -p $out
timeout 2400 python -m pytest tests -m gpu -q -s 2>&1 > $out/${tag}_pytest_full.log
grep -E "passed|failed" $out/${tag}_pytest_full.log | tail -3
grep -E "^(FAILED|ERROR)" $out/${tag}_pytest_full.log | head -20
timeout 900 python bench.py --steps 10 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err || { echo "bench failed"; tail -20 $out/${tag}_bench.err; }
python - <<PY
import json
d=json.load(open("$out/${tag}_bench.json"))
print({k: d[k] for k in ("value","ms_per_step")}, d["e2e"]["value"], d["roofline"]["frac"])
print(d.get("parity"))
PY
