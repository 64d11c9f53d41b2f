#!/bin/bash
# GPU call 5: parity tests, sweep after the record-prefetch reordering, ncu of the default
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for L in 1 2; do for O in 3 4; do for A in 0 512; do
SPGX_VBD_LANES=$L SPGX_VBD_OCC=$O SPGX_VBD_LOOKAHEAD=$A python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench5_g${L}o${O}a${A}.json 2>&1
done; done; done
SPGX_VBD_LANES=4 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench5_g4.json 2>&1
for L in 2 4; do for A in 0 512; do
SPGX_VBD_LANES=$L SPGX_VBD_LOOKAHEAD=$A python bench.py --config 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench5_c4_g${L}a${A}.json 2>&1
done; done
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:vbdStream -s 77 -c 7 -o gpurun_out/prof_vbd5 \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail -3
