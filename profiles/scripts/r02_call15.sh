# round 2, call 15: dense path (cfg4): FFMA / k>=0 fast sweep + direct list walk
set -x
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
T=profiles/trace_build
DH_B200_LIB=$T/libdh_dense.so python -m pytest tests/test_gpu_parity.py tests/test_two_render_paths.py tests/test_gpu_properties.py tests/test_gpu_python_api.py tests/test_lowp_planes.py -m gpu -x -q 2>&1 | tail -3
for dd in 0 1; do
  DH_G_DENSE_DIRECT=$dd DH_B200_LIB=$T/libdh_dense.so $B --workload cfg4 --steps 100 --warmup 10 > gpurun_out/cfg4_dd$dd.json 2>/dev/null
done
DH_B200_LIB=$T/libdh_dense.so $B --workload cfg3 --steps 200 --warmup 20 > gpurun_out/cfg3_dense.json 2>/dev/null
DH_B200_LIB=$T/libdh_dense.so $B --workload cfg5 --batch 384 --steps 100 --warmup 10 > gpurun_out/b384_dense.json 2>/dev/null
DH_B200_LIB=$T/libdh_dense.so $B --workload cfg2 --steps 200 --warmup 20 > gpurun_out/cfg2_dense.json 2>/dev/null
python profiles/show_bench.py gpurun_out/cfg4_dd?.json gpurun_out/cfg3_dense.json gpurun_out/b384_dense.json gpurun_out/cfg2_dense.json | awk '{print $1, $4, $5}'
