# round 2, call 18: host entry, small inputs: gather launch vs the renderer reading the pinned arrays in place
set -x
B="python bench.py --no-cpu-baseline --no-side-legs"
T=profiles/trace_build
for g in 1 2; do
  DH_HOST_GATHER=$g DH_B200_LIB=$T/libdh_cur.so $B --workload cfg5 --batch 384 --steps 200 --warmup 10 > gpurun_out/e2e_b384_gather$g.json 2>/dev/null
  DH_HOST_GATHER=$g DH_B200_LIB=$T/libdh_cur.so $B --workload cfg3 --steps 200 --warmup 10 > gpurun_out/e2e_cfg3_gather$g.json 2>/dev/null
  DH_HOST_GATHER=$g DH_B200_LIB=$T/libdh_cur.so $B --workload cfg1 --steps 200 --warmup 10 > gpurun_out/e2e_cfg1_gather$g.json 2>/dev/null
done
DH_HOST_GATHER=2 DH_B200_LIB=$T/libdh_cur.so python -m pytest tests/test_gpu_properties.py -m gpu -x -q -k "host" 2>&1 | tail -3
