# round 2, call 27: 32-bit unit-index divisions in unit_header
set -x
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
T=profiles/trace_build
DH_B200_LIB=$T/libdh_div32.so python -m pytest tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_handout_order.py tests/test_guard_bands.py -m gpu -x -q 2>&1 | tail -1
for v in base div32 base div32; do
  for cfg in cfg1 cfg2 cfg3; do
    DH_B200_LIB=$T/libdh_$v.so $B --workload $cfg --steps 300 --warmup 20 > gpurun_out/${cfg}_$v.json 2>/dev/null
  done
  DH_B200_LIB=$T/libdh_$v.so $B --workload cfg5 --batch 384 --steps 150 --warmup 10 > gpurun_out/b384_$v.json 2>/dev/null
  DH_B200_LIB=$T/libdh_$v.so $B --workload cfg5 --steps 40 --warmup 5 > gpurun_out/cfg5_$v.json 2>/dev/null
  python profiles/show_bench.py gpurun_out/cfg?_$v.json gpurun_out/b384_$v.json | awk '{print $1, $4, $5}'
done
