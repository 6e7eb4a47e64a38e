#!/bin/bash
# round-2 session K (N GPUs): all GPU tests (multi-GPU ones included), bench under torchrun at N with in_call + sharded
# workloads, standalone whole-job call with the library's per-GPU laps
N=${1:-8}; tag=${2:-r4k}
out=gpurun_out; mkdir -p $out
nvidia-smi --query-gpu=index,name,memory.total --format=csv > $out/${tag}_smi.csv
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -5 | tee $out/${tag}_pytest.log
( time timeout 2400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > $out/${tag}_scale_$N.json 2> $out/${tag}_scale_$N.err ) 2>&1 | tail -3
python - <<PY
import json
d=json.load(open("$out/${tag}_scale_$N.json"))
print({k: d[k] for k in ("value","ms_per_step","n_gpus")}, d["e2e"]["value"])
print("in_call", d.get("in_call"))
for k,v in (d.get("workloads") or {}).items(): print(k, v)
PY
grep "photon timing" $out/${tag}_scale_$N.err | head -60
tail -3 $out/${tag}_scale_$N.err
PHOTON_TIMING=1 PHOTON_GPUS=$N timeout 600 python tools/full_run.py --gpus $N > $out/${tag}_full_$N.json 2> $out/${tag}_full_$N.err
cat $out/${tag}_full_$N.json; grep "photon timing" $out/${tag}_full_$N.err | head -40
