#!/bin/bash
# GPU call 13 (1 GPU): persistent-warp colour pass without spills: parity + A/B + ncu
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_step.py -x -q -k "variants" > gpurun_out/pytest13.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest13.log
tail -5 gpurun_out/pytest13.log
B="bench.py --steps 20 --warmup 3 --no-cpu-baseline"
SPGX_VBD_PERSIST=0 python $B > gpurun_out/b13_c3_p0.json 2>&1
SPGX_VBD_PERSIST=1 python $B > gpurun_out/b13_c3_p1.json 2>&1
SPGX_VBD_PERSIST=1 python $B --config 4 > gpurun_out/b13_c4_p1.json 2>&1
SPGX_VBD_PERSIST=1 SPGX_VBD_PDL=0 python $B > gpurun_out/b13_c3_p1_nopdl.json 2>&1
python - <<'PY'
import glob, json
for f in sorted(glob.glob('gpurun_out/b13_*.json')):
    for ln in open(f):
        if ln.startswith('{'):
            d = json.loads(ln)
            print(f, d['config']['vertices'], round(d['ms_per_step'], 3), round(d['value'] / 1e9, 3), round(d['roofline']['avg_launch_us'], 2), round(d['roofline']['frac'], 3))
        elif 'rror' in ln: print(f, ln[:200])
PY
SPGX_VBD_PERSIST=1 ncu --set full --clock-control none --import-source on -k regex:vbdStream -s 77 -c 7 -o gpurun_out/prof_vbd13 \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full13.log 2>&1
