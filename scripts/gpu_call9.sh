#!/bin/bash
# GPU call 9 (2 GPUs): where does the 2-GPU overhead come from? single-GPU controls on each device, concurrent replicas,
# partitioned sweep without any exchange, and the exchange variants
set -x
mkdir -p gpurun_out
export SPGX_HALO_TIMEOUT_S=5
B="bench.py --steps 20 --warmup 3 --no-cpu-baseline"
CUDA_VISIBLE_DEVICES=0 python $B > gpurun_out/b9_solo_gpu0.json 2>&1
CUDA_VISIBLE_DEVICES=1 python $B > gpurun_out/b9_solo_gpu1.json 2>&1
CUDA_VISIBLE_DEVICES=0 python $B > gpurun_out/b9_conc_gpu0.json 2>&1 &
CUDA_VISIBLE_DEVICES=1 python $B > gpurun_out/b9_conc_gpu1.json 2>&1
wait
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29601"
SPGX_DEBUG_SKIP_HALO=1 timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 3 --halo p2p > gpurun_out/b9_w2_skip_p2p.json 2> gpurun_out/b9_w2_skip_p2p.err
SPGX_DEBUG_SKIP_HALO=1 timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 3 --halo nccl > gpurun_out/b9_w2_skip_nccl.json 2> gpurun_out/b9_w2_skip_nccl.err
timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 3 --halo p2p > gpurun_out/b9_w2_p2p.json 2> gpurun_out/b9_w2_p2p.err
SPGX_VBD_PDL=0 timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 3 --halo p2p > gpurun_out/b9_w2_p2p_nopdl.json 2> gpurun_out/b9_w2_p2p_nopdl.err
SPGX_VBD_LOOKAHEAD=0 timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 3 --halo p2p > gpurun_out/b9_w2_p2p_nola.json 2> gpurun_out/b9_w2_p2p_nola.err
python bench.py --config 3 --steps 10 --warmup 3 --no-cpu-baseline --rows 2048 > gpurun_out/b9_solo_2048.json 2>&1
python - <<'PY'
import glob, json
for f in sorted(glob.glob('gpurun_out/b9_*.json')):
    for ln in open(f):
        if ln.startswith('{'):
            d = json.loads(ln)
            print(f, d['n_gpus'], d['config']['vertices'], round(d['ms_per_step'], 3), round(d['value'] / 1e9, 3), round(d['roofline']['avg_launch_us'], 2))
PY
