#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_multigpu.py -m gpu -x -q -k "2" > gpurun_out/r2c19_mgpu.log 2>&1; echo "mgpu rc=$?"; tail -2 gpurun_out/r2c19_mgpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/r2c19_bench2.log 2>&1; echo "bench2 rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/r2c19_bench2.log").read().strip().splitlines()[-1])
print("%.3f ms/step e2e %.3f" % (d["ms_per_step"], d["e2e"]["ms_per_step"]), d["parity_check"])
for r in d["rooflines_other"][1:]: print({k:(round(v,3) if isinstance(v,float) else v) for k,v in r.items() if k in ("achieved","frac","ms","nccl_ms","nccl_busbw_gbs")})
PY
