#!/bin/bash
# round-2 call 18 (1 GPU): does L2 residency between the head's kernels pay? transform row order x store eviction hint (family time from bench.py's roofline)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for rev in 1 0; do for pol in first normal; do
  for i in 1 2; do
  B2_TRANSFORM_REVERSE=$rev B2_STORE_POLICY=$pol timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-stock > gpurun_out/r2c18_rev${rev}_${pol}_$i.log 2>&1
  python -c "
import json; d=json.loads(open('gpurun_out/r2c18_rev${rev}_${pol}_$i.log').read().strip().splitlines()[-1]); r=d['roofline']
print('reverse=$rev policy=$pol run $i: family %.4f ms (frac %.3f), step %.3f ms, launches %s' % (r['ms'], r['frac'], d['ms_per_step'], {k: round(v['ms'],3) for k,v in r['launches'].items()}))"
  done
done; done
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_e2e_gpu.py -m gpu -x -q > gpurun_out/r2c18_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/r2c18_tests.log
