#!/bin/bash
# GPU call 39 (1 GPU): round-end verification: full GPU suite, smoke(), both bench arms (config 3), config 6
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -8 | tee gpurun_out/c39_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3 | tee gpurun_out/c39_smoke.log
timeout 600 python bench.py > gpurun_out/c39_bench.json 2> gpurun_out/c39_bench.err; tail -c 600 gpurun_out/c39_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/c39_ref.json 2> gpurun_out/c39_ref.err; tail -c 400 gpurun_out/c39_ref.json
timeout 600 python bench.py --config 6 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/c39_c6.json 2> gpurun_out/c39_c6.err; tail -c 300 gpurun_out/c39_c6.json
