#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 200 python tools/host_bound_probe.py 2>&1 | tail -1 | tee gpurun_out/r2c11_probe.log
B2_LAND_LATE=1 timeout 200 python tools/host_bound_probe.py 2>&1 | tail -1 | tee -a gpurun_out/r2c11_probe.log
timeout 200 python tools/host_bound_probe.py --stock 2>&1 | tail -1 | tee -a gpurun_out/r2c11_probe.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-stock > gpurun_out/r2c11_bench.log 2>&1; python -c "
import json; d=json.loads(open('gpurun_out/r2c11_bench.log').read().strip().splitlines()[-1]); r=d['roofline']; print('bench %.3f ms/step; F1 family %.3f ms frac %.3f; launches %s; F3 %.3f' % (d['ms_per_step'], r['ms'], r['frac'], {k:round(v['ms'],3) for k,v in r['launches'].items()}, d['rooflines_other'][0]['frac']))"
