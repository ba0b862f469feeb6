#!/bin/bash
# round-2 call 3: double-buffered staging (3/5-stage ring) + pair heuristic vs the previous build
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
BASE=$PWD/gke-ray-train_b200/build/ab/libb2train_base.so
P="timeout 300 python tools/f1_probe.py --iters 7"
L1="--N 4096 --chunk 4096 --D 2048 --V 128256 --T 2048"
L8="--N 4096 --chunk 4096 --D 4096 --V 128256 --T 2048"
$P > gpurun_out/r2c3_gpt2_new.log 2>&1
B2_PAIR=0 $P > gpurun_out/r2c3_gpt2_new_single.log 2>&1
B2_PAIR=1 $P > gpurun_out/r2c3_gpt2_new_pair.log 2>&1
B2_LIB=$BASE $P > gpurun_out/r2c3_gpt2_base.log 2>&1
$P $L1 > gpurun_out/r2c3_l1b_new.log 2>&1
B2_LIB=$BASE B2_PAIR=1 $P $L1 > gpurun_out/r2c3_l1b_base_pair.log 2>&1
$P $L8 > gpurun_out/r2c3_l8b_new.log 2>&1
B2_PAIR=0 $P $L8 > gpurun_out/r2c3_l8b_new_single.log 2>&1
for f in gpurun_out/r2c3_*.log; do echo "== $f"; grep -E "^(fwd|bwd|dlogits|gemm)" $f | cut -c1-64; done
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "lmhead or gemm" > gpurun_out/r2c3_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/r2c3_tests.log
