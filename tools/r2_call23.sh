#!/bin/bash
# round-2 call 23 (1 GPU): earlier accumulator release in the epilogue vs the previous build
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
BASE=$PWD/gke-ray-train_b200/build/ab/libb2train_base.so
P="timeout 300 python tools/f1_probe.py --iters 9"
L1="--N 4096 --chunk 4096 --D 2048 --V 128256 --T 2048"
for tag in new base new2 base2; do
  if [ ${tag:0:4} = base ]; then export B2_LIB=$BASE; else unset B2_LIB; fi
  $P > gpurun_out/r2c23_gpt2_$tag.log 2>&1
done
for tag in new base; do
  if [ $tag = base ]; then export B2_LIB=$BASE; else unset B2_LIB; fi
  $P $L1 > gpurun_out/r2c23_l1b_$tag.log 2>&1
done
unset B2_LIB
for f in gpurun_out/r2c23_*.log; do echo "== $f"; grep -E "^(fwd|bwd|dlogits|gemm)" $f | cut -c1-64; done
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_e2e_gpu.py -m gpu -x -q > gpurun_out/r2c23_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/r2c23_tests.log
