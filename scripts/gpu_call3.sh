#!/bin/bash
# GPU call 3: parity tests, FP64 peak, streaming kernel sweep (lanes x occupancy x L2 lookahead), ncu of the default
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
./build/fp64_peak > gpurun_out/fp64_peak.json 2>&1; cat gpurun_out/fp64_peak.json
for L in 1 2; do for O in 3 4 5; do for A in 0 1024; do
SPGX_VBD_LANES=$L SPGX_VBD_OCC=$O SPGX_VBD_LOOKAHEAD=$A python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_g${L}o${O}a${A}.json 2>&1
done; done; done
for A in 128 512 4096; do
SPGX_VBD_LANES=1 SPGX_VBD_OCC=4 SPGX_VBD_LOOKAHEAD=$A python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_g1o4a${A}.json 2>&1
done
python bench.py --config 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4.json 2>&1
python bench.py --config 1 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c1.json 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:vbdStream -s 77 -c 7 -o gpurun_out/prof_vbd3 \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail -5
