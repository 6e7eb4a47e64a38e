#!/bin/bash
# round-2 session L (1 GPU): the no-allocation export test alone in a fresh process, then the whole GPU suite (it runs
# last there), then the default bench line, the launch list of the same command (build_field_kernel after its interior
# fast path), the reference arm, and optionally C5 at full depth.
tag=${1:-r4l}
out=gpurun_out; mkdir -p $out
nvidia-smi --query-gpu=index,name,memory.total --format=csv > $out/${tag}_smi.csv
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -15 | tee $out/${tag}_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee $out/${tag}_pytest.log
( time timeout 1200 python bench.py --steps 10 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err ) 2>&1 | tail -3
tail -2 $out/${tag}_bench.err
python - <<PY
import json
d=json.load(open("$out/${tag}_bench.json"))
print({k: d[k] for k in ("value","ms_per_step","n_gpus")}, d["e2e"], d["roofline"], d["cpu_baseline"], d.get("parity"))
for k,v in (d.get("workloads") or {}).items(): print(k, v)
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 1 --workloads none --no-in-call --no-parity --no-cpu-baseline > $out/${tag}_ncu.log 2>&1
python - <<PY | tee $out/${tag}_launches.txt
import csv, collections
rows = [r for r in csv.reader(open("$out/${tag}_launches.csv")) if len(r) > 5]
hdr = rows[0]; k = hdr.index("Kernel Name"); v = hdr.index("Metric Value"); u = hdr.index("Metric Unit")
acc = collections.OrderedDict()
for r in rows[1:]:
    if r[k] == "Kernel Name": continue
    t = float(r[v].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(r[u], 1e-6)
    a = acc.setdefault(r[k][:110], [0, 0.0]); a[0] += 1; a[1] += t
tot = sum(a[1] for a in acc.values())
for name, (n, t) in sorted(acc.items(), key=lambda e: -e[1][1]):
    print(f"{t:10.3f} ms {100*t/tot:5.1f}% x{n:<4d} {t/n:9.3f} ms/launch  {name}")
PY
( time timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_ref.json 2> $out/${tag}_ref.err ) 2>&1 | tail -3
cat $out/${tag}_ref.json
if [ "$2" = "c5full" ]; then
  ( time timeout 900 python bench.py --steps 3 --warmup 3 --workloads c5_2048 --no-in-call --no-parity > $out/${tag}_c5full.json 2> $out/${tag}_c5full.err ) 2>&1 | tail -3
  python -c "import json; d=json.load(open('$out/${tag}_c5full.json')); print(d.get('workloads'))"; tail -3 $out/${tag}_c5full.err
fi
