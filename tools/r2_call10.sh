#!/bin/bash
# round-2 call 10 (2 GPUs): what does overlapping the gradient exchange buy on GPT-2? stream priority, overlap on/off, CTA caps; 2 runs each
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513"
run() { tag=$1; shift; for i in 1 2; do env "$@" timeout 300 $TR bench.py --gpus 2 --steps 30 --warmup 5 --no-f2-large > gpurun_out/r2c10_${tag}_$i.log 2>&1; python - <<PY
import json
d=json.loads(open("gpurun_out/r2c10_${tag}_$i.log").read().strip().splitlines()[-1])
t=d.get("ddp_trace") or {}
print("${tag} run $i: %.3f ms/step, e2e %.3f, exposed %s" % (d["ms_per_step"], d["e2e"]["ms_per_step"], t.get("exposed_ms")))
PY
done; }
run prio_high B2_DDP_TRACE=1 B2_DDP_STREAM_PRIORITY=-1
run prio_norm B2_DDP_TRACE=1 B2_DDP_STREAM_PRIORITY=0
run no_overlap B2_DDP_TRACE=1 B2_DDP_OVERLAP=0
run prio_high_blocks32 B2_DDP_TRACE=1 B2_DDP_STREAM_PRIORITY=-1 B2_DDP_MAX_BLOCKS=32
run prio_high_blocks296 B2_DDP_TRACE=1 B2_DDP_STREAM_PRIORITY=-1 B2_DDP_MAX_BLOCKS=296
run prio_high_bucket25 B2_DDP_TRACE=1 B2_DDP_STREAM_PRIORITY=-1 B2_DDP_BUCKET_MB=25
timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-stock > gpurun_out/r2c10_n1.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/r2c10_n1.log').read().strip().splitlines()[-1]); print('N=1 on this box: %.3f ms/step, e2e %.3f' % (d['ms_per_step'], d['e2e']['ms_per_step']))"
