#!/bin/bash
# round-2 session S (1 GPU): build_field with the face voxels rotating over the warps; field-ingest tests, launch list, ncu.
tag=${1:-r4z}
out=gpurun_out; mkdir -p $out
timeout 600 python -m pytest tests/test_field_ingest.py -m gpu -x -q 2>&1 | tail -4 | tee $out/${tag}_ingest_test.log
PCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --workloads none --no-parity"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $out/${tag}_launches.csv $PCMD > $out/${tag}_ncu.log 2>&1
grep -i "build_field\|build_cells" $out/${tag}_launches.csv | cut -c1-260
PHOTON_NO_FIELD_CACHE=1 timeout 600 ncu --set full --import-source on --clock-control none -k regex:build_field -c 1 -o $out/${tag}_build_field -f $PCMD > $out/${tag}_ncu_bf.log 2>&1; tail -1 $out/${tag}_ncu_bf.log
ncu -i $out/${tag}_build_field.ncu-rep --page details > $out/${tag}_build_field_details.txt 2>&1
ncu -i $out/${tag}_build_field.ncu-rep --page raw --csv > $out/${tag}_build_field_raw.csv 2>&1
ncu -i $out/${tag}_build_field.ncu-rep --page source --csv 2>&1 | gzip > $out/${tag}_build_field_source.csv.gz
rm -f $out/${tag}_build_field.ncu-rep
