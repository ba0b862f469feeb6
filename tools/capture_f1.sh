#!/bin/bash
# One gpurun call's worth of F1 evidence (run from the repo root ON THE GPU BOX, i.e. as the gpurun command):
#   1. the probe without a profiler (CUDA-event timings; also proves the command exits 0 before ncu touches it),
#   2. an `ncu --set full` capture of the F1 kernels of the same command (one GPU, source-level samples included),
#   3. the launch list of bench.py (gpu__time_duration only) for the share-of-step table.
# Reports land in gpurun_out/; read them back in the build container with
#   ncu -i gpurun_out/f1.ncu-rep --page raw --csv > ...          (per-launch metrics)
#   ncu -i gpurun_out/f1.ncu-rep --page source --csv --print-source sass > /tmp/sass.csv
#   python tools/ncu_source_stalls.py /tmp/sass.csv --kernel 'umma_kernel<(int)3'
# and copy the summaries you quote into profiles/.
set -u
mkdir -p gpurun_out
TAG=${1:-f1}
timeout 120 python tools/f1_probe.py --iters 5 > gpurun_out/${TAG}_probe.log 2>&1 || { echo "probe failed"; tail -5 gpurun_out/${TAG}_probe.log; exit 1; }
cat gpurun_out/${TAG}_probe.log
timeout 600 ncu --set full --clock-control none --import-source on \
    -k regex:"umma_kernel|ce_dlogits_inplace|splitk" -c 14 -o gpurun_out/${TAG} -f \
    python tools/f1_probe.py --iters 1 > gpurun_out/${TAG}_ncu.log 2>&1
tail -3 gpurun_out/${TAG}_ncu.log
if [ "${2:-}" = "--launches" ]; then
  timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_plain.log 2>&1 && \
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 3300 -c 1400 --csv \
      --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline \
      > gpurun_out/${TAG}_bench_ncu.log 2>&1
  gzip -f gpurun_out/${TAG}_launches.csv
fi
