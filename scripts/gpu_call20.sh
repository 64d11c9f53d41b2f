#!/bin/bash
# GPU call 20 (1 GPU): own position through the async gather tile: parity + config 3/4 A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_step.py tests/test_golden.py tests/test_gpu_multi.py -x -q -m gpu > gpurun_out/pytest20.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest20.log; tail -3 gpurun_out/pytest20.log
B="bench.py --steps 20 --warmup 3 --no-cpu-baseline"
run() { label=$1; shift
  out=$(env "$@" timeout 300 python $B 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print(round(d['ms_per_step'],3), round(d['roofline']['avg_launch_us'],2), round(d['roofline']['frac'],3), round(d['e2e']['ms_per_step'],3))
")
  echo "$label $out" | tee -a gpurun_out/sweep20.txt
}
: > gpurun_out/sweep20.txt
run "c3 lanes=2" SPGX_VBD_LANES=2
run "c3 lanes=1" SPGX_VBD_LANES=1
run "c3 lanes=4" SPGX_VBD_LANES=4
B="bench.py --steps 20 --warmup 3 --no-cpu-baseline --config 4"
run "c4" X=1
