#!/bin/bash
# round-2 session O (1 GPU): instruction-level profile of the PIV path without a field (BASELINE C1: Mie, thick lens,
# bilinear splat) -- which stage of generate / lens / sensor / splat the 1785 warp instructions per ray go to.
tag=${1:-r4o}
out=gpurun_out; mkdir -p $out
python tools/workload_probe.py piv piv_thin piv_diffuse piv_diffraction 2>&1 | tail -8 | tee $out/${tag}_piv_workloads.txt
timeout 900 ncu --set full --import-source on --clock-control none -k regex:trace_rays -s 1 -c 1 -o $out/${tag}_piv -f python tools/workload_probe.py piv > $out/${tag}_ncu_piv.log 2>&1; tail -1 $out/${tag}_ncu_piv.log
ncu -i $out/${tag}_piv.ncu-rep --page details > $out/${tag}_piv_details.txt 2>&1
ncu -i $out/${tag}_piv.ncu-rep --page raw --csv > $out/${tag}_piv_raw.csv 2>&1
ncu -i $out/${tag}_piv.ncu-rep --page source --csv 2>&1 | gzip > $out/${tag}_piv_source.csv.gz
ncu -i $out/${tag}_piv.ncu-rep --page source --csv --print-source sass,cuda 2>/dev/null | gzip > $out/${tag}_piv_source_cuda.csv.gz
rm -f $out/${tag}_piv.ncu-rep
ls -la $out | tail -8
