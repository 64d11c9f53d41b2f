#!/bin/bash
# GPU call 11 (8 GPUs): identity test on 4 ranks, config 5 (16M-vertex cloth) at 1/2/4/8 GPUs, weak scaling at 8
set -x
mkdir -p gpurun_out
export SPGX_HALO_TIMEOUT_S=5
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/pytest_multi11.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi11.log
tail -5 gpurun_out/pytest_multi11.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --config 5 > gpurun_out/b11_c5_n1.json 2> gpurun_out/b11_c5_n1.err &
P1=$!
CUDA_VISIBLE_DEVICES=1 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/b11_c3_n1.json 2> gpurun_out/b11_c3_n1.err
wait $P1
for N in 8 4 2; do
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29601"
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --config 5 --halo p2p > gpurun_out/b11_c5_n${N}_p2p.json 2> gpurun_out/b11_c5_n${N}_p2p.err
done
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29601"
timeout 600 $TR bench.py --gpus 8 --steps 5 --warmup 3 --config 5 --halo nccl > gpurun_out/b11_c5_n8_nccl.json 2> gpurun_out/b11_c5_n8_nccl.err
timeout 600 $TR bench.py --gpus 8 --steps 20 --warmup 3 --halo p2p > gpurun_out/b11_w8_p2p.json 2> gpurun_out/b11_w8_p2p.err
python - <<'PY'
import glob, json
for f in sorted(glob.glob('gpurun_out/b11_*.json')):
    for ln in open(f):
        if ln.startswith('{'):
            d = json.loads(ln)
            print(f, d['n_gpus'], d['config']['vertices'], round(d['ms_per_step'], 3), round(d['value'] / 1e9, 3), round(d['roofline']['avg_launch_us'], 2), round(d['e2e']['value']/1e9, 3))
PY
tail -n 3 gpurun_out/b11_*.err | cut -c1-240 | tail -40
