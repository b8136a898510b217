# session 12, call N: cost-based per-unit rule as shipped (FRESH 3 x, ACCUMULATE 7 x the unit's pixels): tests, probes, cfg4
set -x
python -m pytest tests/test_tiles_path.py tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_round2_surface.py tests/test_gpu_python_api.py -m gpu -q -x 2>&1 | tail -4 | tee gpurun_out/pytest_dense.log
echo "== rule as shipped" | tee -a gpurun_out/dense_rule.txt
python profiles/scripts/dense_rule_probe.py 2>&1 | tee -a gpurun_out/dense_rule.txt
python profiles/scripts/padded_t_probe.py 2>&1 | tee gpurun_out/padded_t_dense.txt
B="python bench.py --no-cpu-baseline --no-side-legs"
$B --workload cfg4 --steps 200 --warmup 10 > gpurun_out/plain_cfg4.json 2>/dev/null
