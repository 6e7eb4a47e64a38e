#!/bin/bash
# round-2 session D: grouped fan-out kernel on C4 (A/B), tests, ncu of the grouped kernel
tag=${1:-r4d}
out=gpurun_out; mkdir -p $out
{
echo "== C4 grouped (default)"; timeout 900 python tools/workload_probe.py c4_1024 c4_512
echo "== C4 PHOTON_GROUP=0"; PHOTON_GROUP=0 timeout 900 python tools/workload_probe.py c4_1024
echo "== C4 grouped, occupancy 2/4"; for o in 2 4; do PHOTON_OCC=$o timeout 900 python tools/workload_probe.py c4_1024; done
echo "== C4 grouped, PHOTON_RAY_ORDER=0"; PHOTON_RAY_ORDER=0 timeout 900 python tools/workload_probe.py c4_1024
echo "== C4 grouped, PHOTON_MORTON=0"; PHOTON_MORTON=0 timeout 900 python tools/workload_probe.py c4_1024
} > $out/${tag}_c4_grouped.txt 2>&1
cat $out/${tag}_c4_grouped.txt
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -6
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:trace_rays -s 2 -c 1 -o $out/${tag}_c4 -f python tools/workload_probe.py c4_1024 > $out/${tag}_ncu_c4.log 2>&1
tail -2 $out/${tag}_ncu_c4.log
