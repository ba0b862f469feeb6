#!/bin/bash
# round-2 call 8 (8 GPUs): W=8 tests on real peers (NVLS kernel), bench at N=8: GPT-2, Llama-1B and Llama-8B shapes, replicated vs sharded
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 420 python -m pytest tests/test_multigpu.py -m gpu -x -q -k "4" > gpurun_out/r2c8w4_mgpu.log 2>&1; echo "mgpu rc=$?"; tail -4 gpurun_out/r2c8w4_mgpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29514"
B2_DDP_TRACE=1 timeout 240 $TR bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/r2c8w4_gpt2_n4.log 2>&1; echo "gpt2 rc=$?"; tail -1 gpurun_out/r2c8w4_gpt2_n4.log | cut -c1-160
for v in 0 2; do
  B2_DDP_TRACE=1 timeout 300 $TR bench.py --gpus 4 --steps 20 --warmup 5 --workload llama-1b --shard-optimizer $v --no-f2-large > gpurun_out/r2c8w4_l1b_n4_shard$v.log 2>&1; echo "l1b shard$v rc=$?"; tail -1 gpurun_out/r2c8w4_l1b_n4_shard$v.log | cut -c1-160
done
for v in 2 0; do
  timeout 420 $TR bench.py --gpus 4 --steps 20 --warmup 5 --workload llama-8b --shard-optimizer $v --no-f2-large > gpurun_out/r2c8w4_l8b_n4_shard$v.log 2>&1; echo "l8b shard$v rc=$?"; tail -1 gpurun_out/r2c8w4_l8b_n4_shard$v.log | cut -c1-160
done
