# usage: tools/_run_quick.sh TAG [kernel-regex]   (run on the GPU box through gpurun)
TAG=${1:-x}; K=${2:-grid_warp_kernel}
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/pytest_$TAG.log; cat gpurun_out/pytest_$TAG.log
python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench_$TAG.json'));print(d['value'],d['e2e']['value'],d['stage_ms'])"
ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 1 -o gpurun_out/prof_$TAG python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_$TAG.log 2>&1
echo "ncu rc=$?"
