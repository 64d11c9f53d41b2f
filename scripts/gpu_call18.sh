#!/bin/bash
# GPU call 18 (1 GPU): one-warp blocks under an explicit register cap (17 / 18 / 19 resident warps per SM) vs (128, 4)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_step.py -x -q -k variants > gpurun_out/pytest18.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest18.log; tail -3 gpurun_out/pytest18.log
B="bench.py --steps 20 --warmup 3 --no-cpu-baseline"
run() { label=$1; shift
  out=$(env "$@" timeout 300 python $B 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print(round(d['ms_per_step'],3), round(d['roofline']['avg_launch_us'],2), round(d['roofline']['frac'],3), round(d['e2e']['ms_per_step'],3))
")
  echo "$label $out" | tee -a gpurun_out/sweep18.txt
}
: > gpurun_out/sweep18.txt
for L in 2 1; do for O in 4 17 18 19; do run "lanes=$L occ=$O" SPGX_VBD_LANES=$L SPGX_VBD_OCC=$O; done; done
run "c4 occ=4" SPGX_VBD_OCC=4 BENCH_CONFIG=4
