# session 12, call A: state of HEAD on a fresh box — GPU tests, smoke, default bench line, ncu --set full of tiles_kernel on cfg4
set -x
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_cfg5_n1.json 2> gpurun_out/bench_cfg5_n1.err
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
$B --workload cfg4 --steps 50 --warmup 5 > gpurun_out/plain_cfg4.json 2>gpurun_out/plain_cfg4.err &&
ncu --set full --clock-control none --import-source on -k regex:tiles_kernel -s 4 -c 1 -o gpurun_out/tiles_cfg4 $B --workload cfg4 --steps 4 --warmup 3 > gpurun_out/ncu_cfg4.log 2>&1
ls -la gpurun_out/
