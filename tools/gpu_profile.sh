#!/bin/bash
# Runs on the GPU box (under gpurun): GPU parity tests, the bench line, then the ncu launch list
# and one full capture of the trace kernel for the same bench command.  Outputs -> gpurun_out/<tag>_*.
tag=${1:-r1}
steps=${2:-5}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv > $out/${tag}_smi.csv 2>&1
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -60 > $out/${tag}_pytest.log
tail -3 $out/${tag}_pytest.log
CMD="python bench.py --steps $steps --warmup 3"
timeout 900 $CMD > $out/${tag}_bench.json 2> $out/${tag}_bench.err || { echo "bench failed"; tail -20 $out/${tag}_bench.err; exit 1; }
cat $out/${tag}_bench.json
if [ "${NO_NCU:-0}" = "1" ]; then exit 0; fi
PCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 $PCMD > $out/${tag}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv $PCMD > $out/${tag}_ncu_launches.log 2>&1
timeout 600 $PCMD > $out/${tag}_plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:trace_rays -s 3 -c 1 -o $out/${tag}_trace -f $PCMD > $out/${tag}_ncu_full.log 2>&1
tail -3 $out/${tag}_ncu_full.log
ls -la $out
