#!/bin/bash
# GPU call 21 (1 GPU): L1 prefetch of the record two steps ahead x shared-memory carveout
mkdir -p gpurun_out
B="bench.py --steps 20 --warmup 3 --no-cpu-baseline"
run() { label=$1; shift
  out=$(env "$@" timeout 300 python $B 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print(round(d['ms_per_step'],3), round(d['roofline']['avg_launch_us'],2), round(d['roofline']['frac'],3), round(d['e2e']['ms_per_step'],3))
")
  echo "$label $out" | tee -a gpurun_out/sweep21.txt
}
: > gpurun_out/sweep21.txt
run "base" X=1
for C in -1 40 50 60 75; do run "l1pf carve=$C" SPGX_VBD_L1PF=1 SPGX_VBD_CARVEOUT=$C; done
for C in 40 50 60; do run "nopf carve=$C" SPGX_VBD_CARVEOUT=$C; done
SPGX_VBD_L1PF=1 SPGX_VBD_CARVEOUT=50 timeout 600 python -m pytest tests/test_gpu_step.py -x -q -m gpu -k "cloth or beam" 2>&1 | tail -2
