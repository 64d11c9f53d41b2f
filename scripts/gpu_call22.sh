#!/bin/bash
# GPU call 22 (1 GPU): automatic carveout; configs 3, 4, 5 and 2; ncu source-level profile of the new kernel
mkdir -p gpurun_out
run() { label=$1; shift
  out=$(timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline "$@" 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print(round(d['ms_per_step'],3), round(d['roofline']['avg_launch_us'],2), round(d['roofline']['frac'],3), round(d['e2e']['ms_per_step'],3))
")
  echo "$label $out" | tee -a gpurun_out/sweep22.txt
}
: > gpurun_out/sweep22.txt
run c3 --config 3
run c4 --config 4
run c5 --config 5 --steps 5
run c2 --config 2
timeout 900 ncu --set full --clock-control none --import-source on -k regex:vbdStream -s 77 -c 3 -f -o gpurun_out/prof_vbd22 \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full22.log 2>&1
tail -2 gpurun_out/ncu_full22.log
