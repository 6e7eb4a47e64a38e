#!/bin/bash
# Runs on the GPU box: the edge-case GPU tests, then one full ncu capture of the trace kernel (for the
# DRAM traffic figure of bench.py's roofline) and the launch list of the same bench command.
tag=${1:-r3b}
out=gpurun_out
mkdir -p $out
timeout 300 python -m pytest tests/test_gpu_edges.py -m gpu -q 2>&1 | tail -30 > $out/${tag}_edges.log
tail -5 $out/${tag}_edges.log
PCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $PCMD > $out/${tag}_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:trace_rays -s 3 -c 1 -o $out/${tag}_trace -f $PCMD > $out/${tag}_ncu_full.log 2>&1
tail -2 $out/${tag}_ncu_full.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv $PCMD > $out/${tag}_ncu_launches.log 2>&1
ls -la $out | tail -8
