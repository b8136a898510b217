# session 12, call D: accumulate sparse path with both rows' loads issued up front
set -x
python -m pytest tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_gpu_python_api.py tests/test_guard_bands.py tests/test_handout_order.py -m gpu -q -x 2>&1 | tail -3
python profiles/scripts/accum_bench.py 2>&1 | tee gpurun_out/accum_d.txt
