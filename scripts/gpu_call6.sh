#!/bin/bash
# GPU call 6 (1 GPU): parity tests after the partition refactor, sweep after the scoreboard-dominance fix, ncu (G=2)
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for L in 1 2; do for A in 0 512; do
SPGX_VBD_LANES=$L SPGX_VBD_LOOKAHEAD=$A python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench6_g${L}a${A}.json 2>&1
done; done
SPGX_VBD_LANES=2 SPGX_VBD_LOOKAHEAD=2048 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench6_g2a2048.json 2>&1
SPGX_VBD_LANES=2 SPGX_VBD_OCC=5 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench6_g2o5.json 2>&1
python bench.py --config 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench6_c4.json 2>&1
python bench.py --config 5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench6_c5.json 2>&1
SPGX_VBD_LANES=2 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 &&
SPGX_VBD_LANES=2 ncu --set full --clock-control none --import-source on -k regex:vbdStream -s 77 -c 7 -o gpurun_out/prof_vbd6 \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail -3
