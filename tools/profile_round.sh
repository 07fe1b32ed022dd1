#!/bin/bash
# One gpurun call's worth of evidence for profiles/: GPU tests, the bench line (own arm and the CPU arm), the ncu launch
# list of one bench step and an `ncu --set full` capture of one launch of every kernel of the invert.
#   gpurun --timeout 1800 -- 'bash tools/profile_round.sh r2z'
# Every profiler pass runs after the same command has exited 0 without the profiler.  Output: gpurun_out/<tag>_*.
set -u
TAG=${1:-rX}
OUT=gpurun_out
mkdir -p $OUT
KERNELS='regex:^(cols[12]_kernel|correct_kernel|finv_kernel|grid_|hist_kernel|uext_kernel|items_kernel|records_kernel|rows_kernel|scan_|whist|wscreen|transpose_crop|vector_fft|regular_fft)'
(time timeout 900 python -m pytest tests -m gpu -x -q) > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest rc=$?"
timeout 400 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_reference.json 2>> $OUT/${TAG}_bench.err
echo "reference arm rc=$?"
CMD="python bench.py --steps 1 --warmup 1 --no-cpu --no-parity"
if $CMD > $OUT/plain.log 2>&1; then
  # launches of the last timed step and the e2e step included; warm-up launches skipped by name filter + count
  ncu --metrics gpu__time_duration.sum --clock-control none -k "$KERNELS" -c 400 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/ncu1.log 2>&1
  # one launch of each kernel of one step (the first step is the warm-up, buffers grow there); the report of ~40
  # kernels exceeds what gpurun brings back, so it is summarised on the box (tools/ncu_kernels.py) and dropped
  ncu --set full --clock-control none -k "$KERNELS" -s 36 -c 40 -o $OUT/${TAG}_full $CMD > $OUT/ncu2.log 2>&1
  python tools/ncu_kernels.py $OUT/${TAG}_full.ncu-rep --traffic $OUT/grid_kernel_traffic.json grid_warp_kernel > $OUT/${TAG}_kernels_ncu_summary.txt 2>&1
  rm -f $OUT/${TAG}_full.ncu-rep
  # the gridding kernel alone, with source-level sampling
  ncu --set full --clock-control none --import-source on -k regex:grid_warp_kernel -s 1 -c 1 -o $OUT/${TAG}_grid $CMD > $OUT/ncu3.log 2>&1
  echo "ncu done"
else
  echo "plain run failed; not profiled"
fi
ls -la $OUT | tail -12
