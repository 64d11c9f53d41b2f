#!/bin/bash
# GPU call 17 (1 GPU): knob sweep of the streaming colour pass on config 3 (lanes x occupancy, persist, lookahead)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,power.limit,memory.total --format=csv > gpurun_out/sweep17.txt
B="bench.py --steps 20 --warmup 3 --no-cpu-baseline"
run() { # label, env...
  label=$1; shift
  out=$(env "$@" timeout 300 python $B 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print(round(d['ms_per_step'],3), round(d['roofline']['avg_launch_us'],2), round(d['roofline']['frac'],3), round(d['e2e']['ms_per_step'],3))
")
  echo "$label $out" | tee -a gpurun_out/sweep17.txt
}
for L in 1 2 4; do for O in 3 4 5; do run "lanes=$L occ=$O" SPGX_VBD_LANES=$L SPGX_VBD_OCC=$O; done; done
run "persist lanes=2" SPGX_VBD_PERSIST=1
run "persist lanes=1" SPGX_VBD_PERSIST=1 SPGX_VBD_LANES=1
run "lookahead=128" SPGX_VBD_LOOKAHEAD=128
run "lookahead=256" SPGX_VBD_LOOKAHEAD=256
run "lookahead=1024" SPGX_VBD_LOOKAHEAD=1024
run "lookahead=2048" SPGX_VBD_LOOKAHEAD=2048
run "lookahead=0" SPGX_VBD_LOOKAHEAD=0
run "nopdl" SPGX_VBD_PDL=0
