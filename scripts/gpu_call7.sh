#!/bin/bash
# GPU call 7 (1 GPU): full parity suite on the restored tree, default bench + reference arm, launch list, ncu --set full
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
python bench.py > gpurun_out/bench7_default.json 2> gpurun_out/bench7_default.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench7_reference.json 2>&1
python bench.py --config 2 --steps 10 --warmup 3 > gpurun_out/bench7_c2.json 2>&1
python bench.py --config 4 --steps 10 --warmup 3 > gpurun_out/bench7_c4.json 2>&1
python bench.py --config 5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench7_c5.json 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain7.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches7.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches7.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:vbdStream -s 77 -c 7 -o gpurun_out/prof_vbd7 \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full7.log 2>&1
nproc; lscpu | grep "Model name"
ls -la gpurun_out | tail -5
