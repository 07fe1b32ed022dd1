TAG=${1:-r1c}
timeout 600 python tools/gpu_check.py > gpurun_out/check_$TAG.log 2>&1; echo "check rc=$?"
grep -c "^\[ok" gpurun_out/check_$TAG.log; grep "ERR" gpurun_out/check_$TAG.log | head
grep -E "rel_rms" gpurun_out/check_$TAG.log | cut -c1-100 | head -8
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"
cat gpurun_out/bench_$TAG.json
python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:'cip::|cub::|fft' -c 300 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu1.log 2>&1
echo "ncu1 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:grid_kernel -s 1 -c 1 -o gpurun_out/prof_grid_$TAG python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu2.log 2>&1
echo "ncu2 rc=$?"
