# session 12, call K: dominance pruning in the tile queue — tests, cfg4 bench + ncu
set -x
python -m pytest tests/test_tiles_path.py tests/test_gpu_parity.py tests/test_gpu_properties.py -m gpu -q -x 2>&1 | tail -4 | tee gpurun_out/pytest_dense.log
B="python bench.py --no-cpu-baseline --no-side-legs"
$B --workload cfg4 --steps 200 --warmup 10 > gpurun_out/plain_cfg4.json 2>/dev/null
python profiles/scripts/padded_t_probe.py 2>&1 | grep -E "cfg4|\"T\": 256" | tee gpurun_out/padded_t_dense.txt
ncu --set full --clock-control none --import-source on -k regex:groups_kernel -s 4 -c 1 -o gpurun_out/dense_cfg4 $B --no-e2e --workload cfg4 --steps 4 --warmup 3 > gpurun_out/ncu_cfg4.log 2>&1
