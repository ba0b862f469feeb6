#!/bin/bash
# round-2 call 16 (1 GPU), final build: tests, smoke, default bench + reference arm, launch list, ncu of the landing / AdamW / sumsq kernels
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2c16_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2c16_tests.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r2c16_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2c16_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2c16_bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/r2c16_bench.log | cut -c1-200
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2c16_bench_ref.log 2>&1; echo "ref rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1300 --csv \
    --log-file gpurun_out/r2c16_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-stock > gpurun_out/r2c16_bench_ncu.log 2>&1
gzip -f gpurun_out/r2c16_launches.csv
timeout 900 ncu --set full --clock-control none -k regex:"land_copy|adamw_grouped|sumsq_kernel" -c 8 -o gpurun_out/r2c16_tail -f \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-stock > gpurun_out/r2c16_tail_ncu.log 2>&1; tail -1 gpurun_out/r2c16_tail_ncu.log | cut -c1-100
ls -la gpurun_out/r2c16*
