#!/bin/bash
# round-2 call 1: full GPU tests (incl. the new Llama-head shapes), then A/B of the uniform-issue build (B2_LIB) on the
# three head shapes, and a default bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
UI=$PWD/gke-ray-train_b200/build/ab/libb2train_ui.so
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2c1_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2c1_tests.log
B2_LIB=$UI timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_e2e_gpu.py -m gpu -x -q > gpurun_out/r2c1_tests_ui.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2c1_tests_ui.log
for lib in main ui; do
  if [ $lib = ui ]; then export B2_LIB=$UI; else unset B2_LIB; fi
  timeout 300 python tools/f1_probe.py --iters 7 --plain > gpurun_out/r2c1_probe_gpt2_$lib.log 2>&1
  timeout 300 python tools/f1_probe.py --iters 5 --N 4096 --chunk 4096 --D 2048 --V 128256 --T 2048 > gpurun_out/r2c1_probe_l1b_$lib.log 2>&1
  timeout 300 python tools/f1_probe.py --iters 5 --N 4096 --chunk 4096 --D 4096 --V 128256 --T 2048 > gpurun_out/r2c1_probe_l8b_$lib.log 2>&1
done
unset B2_LIB
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2c1_bench.log 2>&1
B2_LIB=$UI timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2c1_bench_ui.log 2>&1
tail -3 gpurun_out/r2c1_tests.log gpurun_out/r2c1_tests_ui.log
cat gpurun_out/r2c1_probe_*.log
