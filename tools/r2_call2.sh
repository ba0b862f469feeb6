#!/bin/bash
# round-2 call 2: where do the storing epilogues lose their time?  (merged uniform-issue build)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
P="timeout 300 python tools/f1_probe.py --iters 7"
$P > gpurun_out/r2c2_gpt2_default.log 2>&1
B2_STORE_DIRECT=1 $P > gpurun_out/r2c2_gpt2_direct.log 2>&1
B2_DEBUG_STORE=1 $P > gpurun_out/r2c2_gpt2_dbg1_stage_only.log 2>&1
B2_DEBUG_STORE=2 $P > gpurun_out/r2c2_gpt2_dbg2_drop.log 2>&1
B2_PAIR=1 $P > gpurun_out/r2c2_gpt2_pair.log 2>&1
B2_PAIR=1 B2_STORE_DIRECT=1 $P > gpurun_out/r2c2_gpt2_pair_direct.log 2>&1
B2_STORE_BOX=32 $P > gpurun_out/r2c2_gpt2_box32.log 2>&1
L1="--N 4096 --chunk 4096 --D 2048 --V 128256 --T 2048"
B2_STORE_DIRECT=1 $P $L1 > gpurun_out/r2c2_l1b_direct.log 2>&1
B2_PAIR=1 $P $L1 > gpurun_out/r2c2_l1b_pair.log 2>&1
for f in gpurun_out/r2c2_*.log; do echo "== $f"; grep -E "^(fwd|bwd|dlogits|gemm)" $f | cut -c1-64; done
B2_STORE_DIRECT=1 timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_e2e_gpu.py -m gpu -x -q > gpurun_out/r2c2_tests_direct.log 2>&1; echo "direct tests rc=$?"; tail -2 gpurun_out/r2c2_tests_direct.log
B2_PAIR=1 timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_e2e_gpu.py -m gpu -x -q > gpurun_out/r2c2_tests_pair.log 2>&1; echo "pair tests rc=$?"; tail -2 gpurun_out/r2c2_tests_pair.log
# ncu --set full of the shipped (kept-logits) route: fwd_keep, inplace transform, dH (+split-K reduce), dW
timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:"umma_kernel|ce_dlogits_inplace|splitk" -c 14 -o gpurun_out/r2c2_f1 -f \
    python tools/f1_probe.py --iters 1 > gpurun_out/r2c2_ncu.log 2>&1
tail -2 gpurun_out/r2c2_ncu.log; ls -la gpurun_out/*.ncu-rep
