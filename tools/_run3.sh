timeout 600 python tools/gpu_check.py > gpurun_out/check2.log 2>&1; echo "check rc=$?"
grep -c "^\[ok" gpurun_out/check2.log; grep "ERR" gpurun_out/check2.log | head
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r1b.json 2> gpurun_out/bench_r1b.err; echo "bench rc=$?"
cat gpurun_out/bench_r1b.json
python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'cip|cub|fft|Fft|FFT' -c 300 --csv --log-file gpurun_out/launches_r1b.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu1.log 2>&1
echo "ncu1 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:grid_kernel -s 1 -c 1 -o gpurun_out/prof_grid_r1b python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu2.log 2>&1
echo "ncu2 rc=$?"
