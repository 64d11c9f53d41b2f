#!/bin/bash
# GPU call 12 (1 GPU): persistent-warp colour pass vs one-slice-per-warp: parity, then A/B on configs 3, 4, 5
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_step.py -x -q -k "variants" > gpurun_out/pytest12.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest12.log
tail -12 gpurun_out/pytest12.log
B="bench.py --steps 20 --warmup 3 --no-cpu-baseline"
for P in 0 1; do
SPGX_VBD_PERSIST=$P python $B > gpurun_out/b12_c3_p$P.json 2>&1
SPGX_VBD_PERSIST=$P python $B --config 4 > gpurun_out/b12_c4_p$P.json 2>&1
done
SPGX_VBD_PERSIST=1 SPGX_VBD_LANES=4 python $B > gpurun_out/b12_c3_p1_g4.json 2>&1
SPGX_VBD_PERSIST=1 SPGX_VBD_LANES=1 python $B > gpurun_out/b12_c3_p1_g1.json 2>&1
SPGX_VBD_PERSIST=1 SPGX_VBD_LOOKAHEAD=0 python $B > gpurun_out/b12_c3_p1_la0.json 2>&1
SPGX_VBD_PERSIST=1 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --config 5 > gpurun_out/b12_c5_p1.json 2>&1
python - <<'PY'
import glob, json
for f in sorted(glob.glob('gpurun_out/b12_*.json')):
    for ln in open(f):
        if ln.startswith('{'):
            d = json.loads(ln)
            print(f, d['config']['vertices'], round(d['ms_per_step'], 3), round(d['value'] / 1e9, 3), round(d['roofline']['avg_launch_us'], 2), round(d['roofline']['frac'], 3))
        elif 'rror' in ln: print(f, ln[:200])
PY
SPGX_VBD_PERSIST=1 ncu --set full --clock-control none --import-source on -k regex:vbdStream -s 77 -c 7 -o gpurun_out/prof_vbd12 \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full12.log 2>&1
