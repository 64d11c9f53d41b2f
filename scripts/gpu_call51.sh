#!/bin/bash
# GPU call 51 (1 GPU): evidence for the current kernel: launch list (one frame) + ncu --set full of the 7 colour passes of one sweep
mkdir -p gpurun_out
timeout 45 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches51_c3.csv \
   python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu51a.log 2>&1
timeout 45 ncu --set full --clock-control none --import-source on -k regex:vbdStream -s 77 -c 7 -f -o gpurun_out/prof_vbd51 \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu51b.log 2>&1
tail -2 gpurun_out/ncu51b.log
