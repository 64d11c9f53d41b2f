#!/bin/bash
# GPU call 42 (1 GPU): ncu --set full of the sparse product (variant 2) inside the CG loop of config 6
mkdir -p gpurun_out
SPGX_SPMV_VARIANT=2 timeout 600 ncu --set full --clock-control none --import-source on -k regex:bsrSpmv -s 40 -c 2 -f -o gpurun_out/prof_spmv42 \
    python bench.py --config 6 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu42.log 2>&1
tail -2 gpurun_out/ncu42.log
