#!/bin/bash
# GPU call 30 (1 GPU): register-capped variants with the owner position re-read from the tile
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_step.py -x -q -k variants > gpurun_out/pytest30.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest30.log; tail -3 gpurun_out/pytest30.log
B="bench.py --steps 20 --warmup 3 --no-cpu-baseline"
run() { label=$1; shift
  out=$(env "$@" timeout 300 python $B 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print(round(d['ms_per_step'],3), round(d['roofline']['avg_launch_us'],2), round(d['roofline']['frac'],3), round(d['e2e']['ms_per_step'],3))
")
  echo "$label $out" | tee -a gpurun_out/sweep30.txt
}
: > gpurun_out/sweep30.txt
for O in 4 17 18 19; do run "lanes=2 occ=$O" SPGX_VBD_LANES=2 SPGX_VBD_OCC=$O; done
run "lanes=1 occ=17" SPGX_VBD_LANES=1 SPGX_VBD_OCC=17
B="bench.py --steps 20 --warmup 3 --no-cpu-baseline --config 4"
for O in 4 17; do run "c4 occ=$O" SPGX_VBD_OCC=$O; done
