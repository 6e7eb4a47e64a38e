#!/bin/bash
# round-2 session T (1 GPU): the Euler hot loop (algorithm 1, what the reference's shipped BOS sample runs) under ncu
tag=${1:-r5b}
out=gpurun_out; mkdir -p $out
python tools/workload_probe.py bos_euler 2>&1 | tail -2 | tee $out/${tag}_euler.txt
timeout 900 ncu --set full --import-source on --clock-control none -k regex:trace_rays -s 1 -c 1 -o $out/${tag}_euler -f python tools/workload_probe.py bos_euler > $out/${tag}_ncu.log 2>&1; tail -1 $out/${tag}_ncu.log
ncu -i $out/${tag}_euler.ncu-rep --page details > $out/${tag}_euler_details.txt 2>&1
ncu -i $out/${tag}_euler.ncu-rep --page raw --csv > $out/${tag}_euler_raw.csv 2>&1
ncu -i $out/${tag}_euler.ncu-rep --page source --csv 2>&1 | gzip > $out/${tag}_euler_source.csv.gz
rm -f $out/${tag}_euler.ncu-rep
