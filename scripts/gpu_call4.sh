#!/bin/bash
# GPU call 4: parity tests, sweep of the unified-pipeline streaming kernel (lanes x occupancy x PDL x lookahead), ncu
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for L in 1 2; do for O in 3 4; do for P in 0 1; do
SPGX_VBD_LANES=$L SPGX_VBD_OCC=$O SPGX_VBD_PDL=$P python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench4_g${L}o${O}p${P}.json 2>&1
done; done; done
SPGX_VBD_LANES=1 SPGX_VBD_OCC=4 SPGX_VBD_LOOKAHEAD=0 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench4_g1o4p1a0.json 2>&1
SPGX_VBD_LANES=2 SPGX_VBD_OCC=4 SPGX_VBD_LOOKAHEAD=0 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench4_g2o4p1a0.json 2>&1
SPGX_VBD_LANES=1 SPGX_VBD_OCC=5 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench4_g1o5p1.json 2>&1
for L in 1 2 4; do
SPGX_VBD_LANES=$L python bench.py --config 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench4_c4_g${L}.json 2>&1
done
python bench.py --config 1 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench4_c1.json 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:vbdStream -s 77 -c 7 -o gpurun_out/prof_vbd4 \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail -3
