#!/bin/bash
# round-2 call 6 (1 GPU): full GPU tests, default bench line (stock + CPU baselines), launch list of the step, ncu --set full of F1
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2c6_tests.log 2>&1; echo "tests rc=$?"; tail -12 gpurun_out/r2c6_tests.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2c6_bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/r2c6_bench.log | cut -c1-300
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2c6_bench_ref.log 2>&1; echo "ref rc=$?"; tail -1 gpurun_out/r2c6_bench_ref.log | cut -c1-200
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-stock > gpurun_out/r2c6_bench_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 3300 -c 1400 --csv \
    --log-file gpurun_out/r2c6_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-stock > gpurun_out/r2c6_bench_ncu.log 2>&1
gzip -f gpurun_out/r2c6_launches.csv; ls -la gpurun_out/r2c6_launches.csv.gz
timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:"umma_kernel|ce_dlogits_inplace|splitk" -c 14 -o gpurun_out/r2c6_f1 -f \
    python tools/f1_probe.py --iters 1 > gpurun_out/r2c6_ncu.log 2>&1; tail -2 gpurun_out/r2c6_ncu.log
