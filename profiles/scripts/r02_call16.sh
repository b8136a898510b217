# round 2, call 16: object-heavy FRESH renders (cfg4) with the lean layout: one staging buffer per warp, 4 CTAs per SM
set -x
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
T=profiles/trace_build
DH_B200_LIB=$T/libdh_lean.so python -m pytest tests/test_gpu_parity.py tests/test_two_render_paths.py tests/test_gpu_properties.py tests/test_gpu_python_api.py tests/test_guard_bands.py -m gpu -x -q 2>&1 | tail -3
for v in 0 1; do
  DH_G_LEAN=$v DH_B200_LIB=$T/libdh_lean.so $B --workload cfg4 --steps 100 --warmup 10 > gpurun_out/cfg4_lean$v.json 2>/dev/null
done
python profiles/show_bench.py gpurun_out/cfg4_lean?.json | awk '{print $1, $4, $5}'
