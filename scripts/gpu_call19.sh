#!/bin/bash
# GPU call 19 (1 GPU): ncu --set full with source-level stall sampling of two colour passes (config 3)
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:vbdStream -s 77 -c 3 -f -o gpurun_out/prof_vbd19 \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full19.log 2>&1
tail -3 gpurun_out/ncu_full19.log; ls -la gpurun_out/
