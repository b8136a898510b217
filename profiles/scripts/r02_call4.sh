# round 2, call 4: who issues the untouched runs in groups2 (first-arriver claim / rotating warp / warp 7) + unit timelines
set -x
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
T=profiles/trace_build
DH_B200_LIB=$T/libdh_z0.so python -m pytest tests/test_single_launch_path.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -5
for v in z0 z1 z2; do
  DH_B200_LIB=$T/libdh_$v.so $B --workload cfg3 --steps 200 --warmup 20 > gpurun_out/cfg3_$v.json 2>/dev/null
  DH_B200_LIB=$T/libdh_$v.so $B --workload cfg5 --batch 384 --steps 100 --warmup 10 > gpurun_out/b384_$v.json 2>/dev/null
done
DH_B200_LIB=$T/libdh_z0.so $B --workload cfg4 --steps 100 --warmup 10 > gpurun_out/cfg4_z0.json 2>/dev/null
DH_G_NOBIN=1 DH_B200_LIB=$T/libdh_z0.so $B --workload cfg4 --steps 100 --warmup 10 > gpurun_out/cfg4_z0_nobin.json 2>/dev/null
python profiles/show_bench.py gpurun_out/cfg3_z?.json gpurun_out/b384_z?.json gpurun_out/cfg4_z0*.json
DH_B200_LIB=$T/libdh_z0trace.so python profiles/scripts/unit_trace.py cfg5 384 > gpurun_out/trace_b384_bin.txt 2>&1
DH_G_NOBIN=1 DH_B200_LIB=$T/libdh_z0trace.so python profiles/scripts/unit_trace.py cfg5 384 > gpurun_out/trace_b384_nobin.txt 2>&1
DH_B200_LIB=$T/libdh_z0trace.so python profiles/scripts/unit_trace.py cfg3 > gpurun_out/trace_cfg3_bin.txt 2>&1
DH_G_NOBIN=1 DH_B200_LIB=$T/libdh_z0trace.so python profiles/scripts/unit_trace.py cfg3 > gpurun_out/trace_cfg3_nobin.txt 2>&1
cat gpurun_out/trace_b384_bin.txt gpurun_out/trace_b384_nobin.txt
