#!/bin/bash
# GPU call 1: parity tests, bench, lanes sweep, ncu launch list + full capture of the colour pass
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; tail -c 3000 gpurun_out/bench_c3.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
for L in 1 2 4; do SPGX_VBD_LANES=$L python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_lanes$L.json 2>&1; done
python bench.py --config 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2.json 2>&1
python bench.py --config 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4.json 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:vbdPass -s 77 -c 7 -o gpurun_out/prof_vbd \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
