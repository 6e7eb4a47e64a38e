#!/bin/bash
# round-2 session A: all GPU tests, bench with the parity leg, ncu of the bench kernel, C4 fan-out probe + ncu
tag=${1:-r4a}
out=gpurun_out; mkdir -p $out
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/${tag}_smi.csv 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x -s 2>&1 | tail -150 > $out/${tag}_pytest.log
tail -5 $out/${tag}_pytest.log
timeout 900 python bench.py --steps 10 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err || { echo "bench failed"; tail -20 $out/${tag}_bench.err; }
cat $out/${tag}_bench.json
PCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv $PCMD > $out/${tag}_ncu_launches.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:trace_rays -s 3 -c 1 -o $out/${tag}_trace -f $PCMD > $out/${tag}_ncu_full.log 2>&1
tail -2 $out/${tag}_ncu_full.log
timeout 1200 python tools/workload_probe.py c4_1024 c4_512 piv piv_diffraction bos_euler > $out/${tag}_workloads.txt 2>&1
cat $out/${tag}_workloads.txt
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:trace_rays -s 2 -c 1 -o $out/${tag}_c4 -f python tools/workload_probe.py c4_1024 > $out/${tag}_ncu_c4.log 2>&1
tail -2 $out/${tag}_ncu_c4.log
ls -la $out | tail -20
