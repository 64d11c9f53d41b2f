#!/bin/bash
# GPU call 47 (1 GPU): ncu --set full of the hand-derived hinge Hessian kernel and the assembly kernel (config 6)
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'hingeHessian|bsrAssemble' -s 2 -c 2 -f -o gpurun_out/prof_hinge47 \
    python bench.py --config 6 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu47.log 2>&1
tail -2 gpurun_out/ncu47.log
