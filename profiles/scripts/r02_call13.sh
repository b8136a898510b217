# round 2, call 13: ACCUMULATE with 4 resident CTAs per SM (no zero tile, one staging buffer per warp) vs 3
set -x
T=profiles/trace_build
for v in acc5 acc6; do
  DH_B200_LIB=$T/libdh_$v.so python -m pytest tests/test_gpu_parity.py tests/test_two_render_paths.py tests/test_guard_bands.py tests/test_gpu_properties.py -m gpu -x -q 2>&1 | tail -2
  DH_B200_LIB=$T/libdh_$v.so python profiles/scripts/accum_bench.py > gpurun_out/accum_$v.txt 2>&1
  cat gpurun_out/accum_$v.txt
done
