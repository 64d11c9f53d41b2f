#!/bin/bash
# GPU call 2: parity tests + streaming kernel variants sweep on config 3
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for L in 1 2 4; do for O in 3 4 5; do
SPGX_VBD_LANES=$L SPGX_VBD_OCC=$O python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_g${L}o${O}.json 2>&1
done; done
SPGX_NO_STREAMS=1 SPGX_VBD_LANES=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_csr.json 2>&1
python bench.py --config 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4.json 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:vbdStream -s 77 -c 7 -o gpurun_out/prof_vbd2 \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
