#!/bin/bash
# round-2 call 7 (2 GPUs): Llama-3.2-1B shape (configs[2], GA=4) replicated vs sharded optimizer, + the examples on 2 GPUs
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512"
for v in 0 1 2; do
  timeout 900 $TR bench.py --gpus 2 --steps 20 --warmup 5 --workload llama-1b --shard-optimizer $v --no-f2-large > gpurun_out/r2c7_l1b_n2_shard$v.log 2>&1; echo "shard$v rc=$?"; tail -1 gpurun_out/r2c7_l1b_n2_shard$v.log | cut -c1-200
done
B2_DDP_BUCKET_MB=400 timeout 900 $TR bench.py --gpus 2 --steps 20 --warmup 5 --workload llama-1b --shard-optimizer 2 --no-f2-large > gpurun_out/r2c7_l1b_n2_shard2_400mb.log 2>&1; tail -1 gpurun_out/r2c7_l1b_n2_shard2_400mb.log | cut -c1-200
timeout 600 python -m pytest tests/test_example_gpu.py -m gpu -x -q > gpurun_out/r2c7_examples.log 2>&1; echo "examples rc=$?"; tail -3 gpurun_out/r2c7_examples.log
