#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu.py tests/test_example_gpu.py -m gpu -x -q -k "2 or two" > gpurun_out/r2c14_mgpu.log 2>&1; echo "mgpu rc=$?"; tail -4 gpurun_out/r2c14_mgpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
B2_DDP_TRACE=1 timeout 600 $TR bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/r2c14_bench2.log 2>&1; echo "bench2 rc=$?"
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 --workload llama-1b --shard-optimizer 2 --no-f2-large > gpurun_out/r2c14_l1b_shard2.log 2>&1; echo "l1b rc=$?"
timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-stock > gpurun_out/r2c14_n1.log 2>&1
python - <<PY
import json
for f in ("bench2","l1b_shard2","n1"):
    d=json.loads(open(f"gpurun_out/r2c14_{f}.log").read().strip().splitlines()[-1])
    print(f, "%.3f ms/step e2e %.3f" % (d["ms_per_step"], d["e2e"]["ms_per_step"]), d.get("parity_check"), (d.get("ddp_trace") or {}).get("exposed_ms"))
PY
