#!/bin/bash
# GPU call 10 (2 GPUs): boundary-first slot order; identity test, then 2-GPU benches (weak config 3, config 4, config 5)
set -x
mkdir -p gpurun_out
export SPGX_HALO_TIMEOUT_S=5
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/pytest_multi10.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi10.log
tail -5 gpurun_out/pytest_multi10.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29601"
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/b10_solo.json 2>&1
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --config 4 > gpurun_out/b10_solo_c4.json 2>&1
SPGX_DEBUG_SKIP_HALO=1 timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 3 --halo p2p > gpurun_out/b10_w2_skip.json 2> gpurun_out/b10_w2_skip.err
timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 3 --halo p2p > gpurun_out/b10_w2_p2p.json 2> gpurun_out/b10_w2_p2p.err
timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 3 --halo nccl > gpurun_out/b10_w2_nccl.json 2> gpurun_out/b10_w2_nccl.err
timeout 300 $TR bench.py --gpus 2 --steps 10 --warmup 3 --config 4 --halo p2p > gpurun_out/b10_c4_p2p.json 2> gpurun_out/b10_c4_p2p.err
timeout 400 $TR bench.py --gpus 2 --steps 5 --warmup 3 --config 5 --halo p2p > gpurun_out/b10_c5_p2p.json 2> gpurun_out/b10_c5_p2p.err
python - <<'PY'
import glob, json
for f in sorted(glob.glob('gpurun_out/b10_*.json')):
    for ln in open(f):
        if ln.startswith('{'):
            d = json.loads(ln)
            print(f, d['n_gpus'], d['config']['vertices'], round(d['ms_per_step'], 3), round(d['value'] / 1e9, 3), round(d['roofline']['avg_launch_us'], 2))
PY
