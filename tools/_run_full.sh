# usage (on the GPU box, through gpurun): bash tools/_run_full.sh TAG
# GPU tests, the default bench (with the CPU baseline), the reference arm, the ncu launch list of one step and
# one `ncu --set full` capture of the gridding kernel -- each profiler pass only after the plain run passed.
TAG=${1:-x}
python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/${TAG}_pytest_gpu.log; cat gpurun_out/${TAG}_pytest_gpu.log
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; cat gpurun_out/${TAG}_bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2>> gpurun_out/${TAG}_bench.err; echo "ref rc=$?"; cat gpurun_out/${TAG}_bench_reference.json
python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:'cip::|cub::|fft' -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu1.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:grid_warp_kernel -s 1 -c 1 -o gpurun_out/${TAG}_grid_kernel python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"
