#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2c4_tests.log 2>&1; echo "tests rc=$?"; tail -15 gpurun_out/r2c4_tests.log
timeout 300 python tools/f1_probe.py --iters 7 > gpurun_out/r2c4_gpt2.log 2>&1; grep -E "^(fwd|bwd|dlogits|gemm)" gpurun_out/r2c4_gpt2.log | cut -c1-64
timeout 300 python tools/f1_probe.py --iters 5 --N 4096 --chunk 4096 --D 4096 --V 128256 --T 2048 > gpurun_out/r2c4_l8b.log 2>&1; grep -E "^(fwd|bwd|dlogits|gemm)" gpurun_out/r2c4_l8b.log | cut -c1-64
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2c4_bench.log 2>&1; echo "bench rc=$?"; tail -c 1500 gpurun_out/r2c4_bench.log
