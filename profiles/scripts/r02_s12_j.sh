# session 12, call J: dense branch out of line (noinline) — tests, padded-T probe, cfg4 / cfg3 / cfg5 bench
set -x
python -m pytest tests/test_tiles_path.py tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_round2_surface.py -m gpu -q -x 2>&1 | tail -4 | tee gpurun_out/pytest_dense.log
python profiles/scripts/padded_t_probe.py 2>&1 | tee gpurun_out/padded_t_dense.txt
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
$B --workload cfg4 --steps 200 --warmup 10 > gpurun_out/plain_cfg4.json 2>/dev/null
$B --workload cfg3 --steps 400 --warmup 20 > gpurun_out/plain_cfg3.json 2>/dev/null
$B --steps 50 --warmup 5 > gpurun_out/plain_cfg5.json 2>/dev/null
