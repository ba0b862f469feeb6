#!/bin/bash
# round-2 call 15 (8 GPUs), final build: W=8 tests on real peers, GPT-2 / Llama-1B / Llama-8B (sharded) at N=8
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_multigpu.py -m gpu -x -q -k "8" > gpurun_out/r2c15_mgpu.log 2>&1; echo "mgpu rc=$?"; tail -4 gpurun_out/r2c15_mgpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518"
B2_DDP_TRACE=1 timeout 240 $TR bench.py --gpus 8 --steps 30 --warmup 5 > gpurun_out/r2c15_gpt2_n8.log 2>&1; echo "gpt2 rc=$?"
B2_DDP_TRACE=1 timeout 300 $TR bench.py --gpus 8 --steps 20 --warmup 5 --workload llama-1b --no-f2-large > gpurun_out/r2c15_l1b_n8_shard0.log 2>&1; echo "l1b rc=$?"
timeout 420 $TR bench.py --gpus 8 --steps 20 --warmup 5 --workload llama-8b --shard-optimizer 2 --no-f2-large > gpurun_out/r2c15_l8b_n8_shard2.log 2>&1; echo "l8b rc=$?"
python - <<PY
import json
for f in ("gpt2_n8","l1b_n8_shard0","l8b_n8_shard2"):
    try:
        d=json.loads(open(f"gpurun_out/r2c15_{f}.log").read().strip().splitlines()[-1])
        print(f, "%.3f ms/step e2e %.3f value %.0f" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["value"]), d.get("parity_check",{}).get("f2"), (d.get("ddp_trace") or {}).get("exposed_ms"))
    except Exception as e:
        print(f, "failed", e)
PY
