#!/bin/bash
# GPU call 8 (2 GPUs): k-GPU == 1-GPU identity for the NCCL and the fused peer-memory halo exchange, then 2-GPU benches
set -x
mkdir -p gpurun_out
export SPGX_HALO_TIMEOUT_S=5
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/pytest_multi8.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi8.log
tail -30 gpurun_out/pytest_multi8.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29601"
for H in p2p nccl; do
timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 3 --halo $H > gpurun_out/bench8_w2_$H.json 2> gpurun_out/bench8_w2_$H.err; echo "exit $?"
timeout 300 $TR bench.py --gpus 2 --steps 10 --warmup 3 --config 4 --halo $H > gpurun_out/bench8_c4_2_$H.json 2> gpurun_out/bench8_c4_2_$H.err; echo "exit $?"
done
timeout 400 $TR bench.py --gpus 2 --steps 5 --warmup 3 --config 5 --halo p2p > gpurun_out/bench8_c5_2_p2p.json 2> gpurun_out/bench8_c5_2_p2p.err; echo "exit $?"
tail -3 gpurun_out/*8*.err
nvidia-smi topo -m | head -12
