#!/bin/bash
# round-2 session N (1 GPU): cubic hot loop on TMA-staged bricks (PHOTON_CUBIC_BRICK=1) beside the quad-row table loop.
# The brick test first and alone under its own timeout (a wedged mbarrier wait must not take the box with it), then the
# whole GPU suite, the A/B through bench.py, and ncu --set full of both cubic kernels, of the bench kernel, of the fan-out
# kernel on C4 and of build_field.  Reports are exported to text ON the box and the .ncu-rep files are deleted: gpurun brings back at most 64 MiB (session M lost its captures to that limit).
tag=${1:-r4n}
out=gpurun_out; mkdir -p $out
export_rep() {  # <rep without extension> [keep]
    ncu -i $1.ncu-rep --page details > $1_details.txt 2>&1
    ncu -i $1.ncu-rep --page raw --csv > $1_raw.csv 2>&1
    ncu -i $1.ncu-rep --page source --csv 2>&1 | gzip > $1_source.csv.gz
    [ -n "$2" ] || rm -f $1.ncu-rep
}
timeout 300 python -m pytest tests/test_cubic_mode.py -m gpu -x -q -k brick 2>&1 | tail -15 | tee $out/${tag}_brick_test.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee $out/${tag}_pytest.log
( time timeout 1200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workloads cubic,cubic_brick,c4_1024 > $out/${tag}_bench.json 2> $out/${tag}_bench.err ) 2>&1 | tail -3
tail -2 $out/${tag}_bench.err
python - <<PY
import json
d=json.load(open("$out/${tag}_bench.json"))
print({k: d[k] for k in ("value","ms_per_step")}, d["e2e"]["value"], d["roofline"]["frac"])
for k,v in (d.get("workloads") or {}).items(): print(k, {q: v.get(q) for q in ("value","unit","ms","roofline_frac","kernel","error","brick_timeouts")})
PY
PCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --workloads none --no-parity"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv $PCMD > $out/${tag}_ncu.log 2>&1
grep -i "build_field\|build_cells" $out/${tag}_launches.csv | cut -c1-260
timeout 900 ncu --set full --import-source on --clock-control none -k regex:trace_rays -s 3 -c 1 -o $out/${tag}_trace -f $PCMD > $out/${tag}_ncu_trace.log 2>&1; tail -1 $out/${tag}_ncu_trace.log
export_rep $out/${tag}_trace
PHOTON_NO_FIELD_CACHE=1 timeout 600 ncu --set full --import-source on --clock-control none -k regex:build_field -c 1 -o $out/${tag}_build_field -f $PCMD > $out/${tag}_ncu_bf.log 2>&1; tail -1 $out/${tag}_ncu_bf.log
export_rep $out/${tag}_build_field
timeout 900 ncu --set full --import-source on --clock-control none -k regex:trace_rays -s 2 -c 1 -o $out/${tag}_c4 -f python tools/workload_probe.py c4_1024 > $out/${tag}_ncu_c4.log 2>&1; tail -1 $out/${tag}_ncu_c4.log
export_rep $out/${tag}_c4
CCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity --workloads"
timeout 900 ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k regex:BSplineFieldBrick -s 1 -c 1 -o $out/${tag}_cubic_brick -f $CCMD cubic_brick > $out/${tag}_ncu_cb.log 2>&1; tail -1 $out/${tag}_ncu_cb.log
export_rep $out/${tag}_cubic_brick
timeout 900 ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k regex:BSplineField -s 1 -c 1 -o $out/${tag}_cubic -f $CCMD cubic > $out/${tag}_ncu_c.log 2>&1; tail -1 $out/${tag}_ncu_c.log
export_rep $out/${tag}_cubic
du -sh $out; ls -la $out
