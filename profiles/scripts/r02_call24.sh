# round 2, call 24: sparse path two rows at a time + single-object instantiation (sp2) against the round-1 form (sp1)
set -x
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
T=profiles/trace_build
DH_B200_LIB=$T/libdh_sp2.so python -m pytest tests/test_gpu_parity.py tests/test_two_render_paths.py tests/test_gpu_properties.py tests/test_gpu_python_api.py tests/test_guard_bands.py tests/test_lowp_planes.py tests/test_handout_order.py tests/test_round2_surface.py -m gpu -x -q 2>&1 | tail -3
for v in sp1 sp2; do
  for cfg in cfg1 cfg2 cfg3 cfg4; do
    DH_B200_LIB=$T/libdh_$v.so $B --workload $cfg --steps 300 --warmup 20 > gpurun_out/${cfg}_$v.json 2>/dev/null
  done
  DH_B200_LIB=$T/libdh_$v.so $B --workload cfg5 --batch 384 --steps 150 --warmup 10 > gpurun_out/b384_$v.json 2>/dev/null
  DH_B200_LIB=$T/libdh_$v.so $B --workload cfg5 --steps 40 --warmup 5 > gpurun_out/cfg5_$v.json 2>/dev/null
  DH_B200_LIB=$T/libdh_$v.so python profiles/scripts/accum_bench.py 2>&1 | head -3 > gpurun_out/accum_$v.txt
done
python profiles/show_bench.py gpurun_out/cfg?_sp?.json gpurun_out/b384_sp?.json | awk '{print $1, $4, $5}'
cat gpurun_out/accum_sp1.txt gpurun_out/accum_sp2.txt
