#!/bin/bash
# round-2 call 9 (1 GPU): A/B of the GEMM epilogue streaming stores + wider split-K window, full tests, smoke, F2 single-GPU ncu, bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
BASE=$PWD/gke-ray-train_b200/build/ab/libb2train_base.so
P="timeout 300 python tools/f1_probe.py --iters 7"
L1="--N 4096 --chunk 4096 --D 2048 --V 128256 --T 2048"
L8="--N 4096 --chunk 4096 --D 4096 --V 128256 --T 2048"
for tag in new base; do
  if [ $tag = base ]; then export B2_LIB=$BASE; else unset B2_LIB; fi
  $P > gpurun_out/r2c9_gpt2_$tag.log 2>&1
  $P $L1 > gpurun_out/r2c9_l1b_$tag.log 2>&1
  $P $L8 > gpurun_out/r2c9_l8b_$tag.log 2>&1
done
unset B2_LIB
for f in gpurun_out/r2c9_*_new.log gpurun_out/r2c9_*_base.log; do echo "== $f"; grep -E "^(fwd|bwd|dlogits|gemm)" $f | cut -c1-64; done
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2c9_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/r2c9_tests.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r2c9_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2c9_smoke.log
timeout 120 python tools/f2_emul_probe.py --world 8 > gpurun_out/r2c9_f2emul.log 2>&1 && \
timeout 600 ncu --set full --clock-control none -k regex:"peer_allreduce" -c 2 -o gpurun_out/r2c9_f2emul -f python tools/f2_emul_probe.py --world 8 --iters 2 > gpurun_out/r2c9_f2emul_ncu.log 2>&1
cat gpurun_out/r2c9_f2emul.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2c9_bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/r2c9_bench.log | cut -c1-250
