# session 12, call B: tile renderer with the packed-fp32 arithmetic sweep — its tests, cfg4 bench, ncu --set full
set -x
python -m pytest tests/test_tiles_path.py tests/test_gpu_parity.py -m gpu -q > gpurun_out/pytest_tiles.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_tiles.log; tail -15 gpurun_out/pytest_tiles.log
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
$B --workload cfg4 --steps 200 --warmup 10 > gpurun_out/plain_cfg4.json 2>gpurun_out/plain_cfg4.err &&
ncu --set full --clock-control none --import-source on -k regex:tiles_kernel -s 4 -c 1 -o gpurun_out/tiles_cfg4_v2 $B --workload cfg4 --steps 4 --warmup 3 > gpurun_out/ncu_cfg4.log 2>&1
python profiles/scripts/accum_one.py > gpurun_out/plain_accum.log 2>&1
