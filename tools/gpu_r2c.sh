#!/bin/bash
# round-2 session C: exact-normalisation A/B (parity at 512^3 + speed), ray-order A/B on C4, texture-unit probe
tag=${1:-r4c}
out=gpurun_out; mkdir -p $out
B="python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-e2e"
echo "== parity at 512^3, exact normalisation (default lib)" > $out/${tag}_parity512.txt
timeout 900 python tools/parity_probe512.py >> $out/${tag}_parity512.txt 2>&1
echo "== parity at 512^3, PH_EXACT_NORM=0 lib" >> $out/${tag}_parity512.txt
PHOTON_B200_LIB=photon_b200/lib/libprt_norm0.so timeout 900 python tools/parity_probe512.py cells "coherent:PHOTON_CELLS=0" >> $out/${tag}_parity512.txt 2>&1
cat $out/${tag}_parity512.txt
echo "== bench default lib"; timeout 600 $B 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])"
echo "== bench norm0 lib"; PHOTON_B200_LIB=photon_b200/lib/libprt_norm0.so timeout 600 $B 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])"
echo "== bench default lib, PHOTON_CELLS=0"; PHOTON_CELLS=0 timeout 600 $B 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])"
{
echo "== C4 ray order on (default)"; timeout 900 python tools/workload_probe.py c4_1024 c4_512
echo "== C4 ray order off"; PHOTON_RAY_ORDER=0 timeout 900 python tools/workload_probe.py c4_1024 c4_512
echo "== C4 ray order on, coherent gatherer (PHOTON_HOT=5, PHOTON_CELLS=0)"; PHOTON_HOT=5 PHOTON_CELLS=0 timeout 900 python tools/workload_probe.py c4_1024 c4_512
echo "== C4 ray order on, occupancy 2/4/5"; for o in 2 4 5; do PHOTON_OCC=$o timeout 900 python tools/workload_probe.py c4_1024; done
} > $out/${tag}_c4_rayorder.txt 2>&1
cat $out/${tag}_c4_rayorder.txt
timeout 300 python tools/texprobe.py $out/${tag}_texprobe.npz
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_oracle_b.py tests/test_field_ingest.py tests/test_cubic_mode.py -m gpu -q 2>&1 | tail -8
