#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2c13_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/r2c13_tests.log
timeout 200 python tools/host_bound_probe.py 2>&1 | tail -1 | tee gpurun_out/r2c13_probe.log
B2_LAND_FLUSH_MB=0 timeout 200 python tools/host_bound_probe.py 2>&1 | tail -1 | tee -a gpurun_out/r2c13_probe.log
B2_LAND_FLUSH_MB=1024 timeout 200 python tools/host_bound_probe.py 2>&1 | tail -1 | tee -a gpurun_out/r2c13_probe.log
for mb in 64 0 256; do B2_LAND_FLUSH_MB=$mb timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-stock > gpurun_out/r2c13_bench_$mb.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/r2c13_bench_$mb.log').read().strip().splitlines()[-1]); print('flush $mb MB: %.3f ms/step, e2e %.3f, launches %d' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['gpu_launches']))"; done
