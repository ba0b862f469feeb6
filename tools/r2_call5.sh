#!/bin/bash
# round-2 call 5 (2 GPUs): peer-memory tests on real peers, then bench at N=2 (parity check, F2 roofline), sharded variants
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/r2c5_mgpu.log 2>&1; echo "mgpu rc=$?"; tail -25 gpurun_out/r2c5_mgpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2c5_bench2.log 2>&1; echo "bench2 rc=$?"; tail -1 gpurun_out/r2c5_bench2.log | cut -c1-400
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 --shard-optimizer 1 --no-f2-large > gpurun_out/r2c5_bench2_shard1.log 2>&1; echo "shard1 rc=$?"; tail -1 gpurun_out/r2c5_bench2_shard1.log | cut -c1-300
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 --shard-optimizer 2 --no-f2-large > gpurun_out/r2c5_bench2_shard2.log 2>&1; echo "shard2 rc=$?"; tail -1 gpurun_out/r2c5_bench2_shard2.log | cut -c1-300
B2_DDP_OVERLAP=0 timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 --no-f2-large > gpurun_out/r2c5_bench2_nooverlap.log 2>&1; echo "nooverlap rc=$?"; tail -1 gpurun_out/r2c5_bench2_nooverlap.log | cut -c1-300
