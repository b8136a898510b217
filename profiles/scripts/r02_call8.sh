# round 2, call 8: two block barriers per unit instead of three (double-buffered list counters); 4-warp CTAs (6 per SM)
set -x
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
T=profiles/trace_build
for v in base w4; do
  DH_B200_LIB=$T/libdh_$v.so python -m pytest tests/test_gpu_parity.py tests/test_two_render_paths.py tests/test_guard_bands.py tests/test_lowp_planes.py -m gpu -x -q 2>&1 | tail -3
  for cfg in cfg1 cfg2 cfg3 cfg4; do
    DH_B200_LIB=$T/libdh_$v.so $B --workload $cfg --steps 200 --warmup 20 > gpurun_out/${cfg}_$v.json 2>/dev/null
  done
  DH_B200_LIB=$T/libdh_$v.so $B --workload cfg5 --batch 384 --steps 100 --warmup 10 > gpurun_out/b384_$v.json 2>/dev/null
  DH_B200_LIB=$T/libdh_$v.so $B --workload cfg5 --steps 30 --warmup 5 > gpurun_out/cfg5_$v.json 2>/dev/null
done
python profiles/show_bench.py gpurun_out/cfg?_base.json gpurun_out/b384_base.json gpurun_out/cfg?_w4.json gpurun_out/b384_w4.json
