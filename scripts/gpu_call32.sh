#!/bin/bash
# GPU call 32 (1 GPU): fused XPBD frame with a hand-rolled grid barrier: parity + config 2 A/B over block counts
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_step.py -x -q -k "fused or xpbd" > gpurun_out/pytest32.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest32.log; tail -3 gpurun_out/pytest32.log
run() { label=$1; shift
  out=$(env "$@" timeout 120 python bench.py --config 2 --steps 20 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print(round(d['ms_per_step'],3), round(d['value']/1e9,3), round(d['e2e']['ms_per_step'],3), d['gpu_launches'])
")
  echo "$label $out" | tee -a gpurun_out/sweep32.txt
}
: > gpurun_out/sweep32.txt
run "passes" SPGX_XPBD_FUSED=0
run "fused cg" SPGX_XPBD_FUSED=1
for B in 148 111 74 64 48 37; do run "fused barrier blocks=$B" SPGX_XPBD_FUSED=2 SPGX_XPBD_FUSED_BLOCKS=$B; done
