# round 2, call 28: the lean layout (4 CTAs per SM, one staging buffer per warp) for the store-bound configs too, x unit size
set -x
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
T=profiles/trace_build
for lt in 256 0; do
  for ub in 196608 131072 98304; do
    DH_G_LEAN_T=$lt DH_G_UNIT_BYTES=$ub DH_B200_LIB=$T/libdh_cur.so $B --workload cfg3 --steps 300 --warmup 20 > gpurun_out/cfg3_lt${lt}_$ub.json 2>/dev/null
    DH_G_LEAN_T=$lt DH_G_UNIT_BYTES=$ub DH_B200_LIB=$T/libdh_cur.so $B --workload cfg5 --batch 384 --steps 150 --warmup 10 > gpurun_out/b384_lt${lt}_$ub.json 2>/dev/null
    DH_G_LEAN_T=$lt DH_G_UNIT_BYTES=$ub DH_B200_LIB=$T/libdh_cur.so $B --workload cfg5 --steps 40 --warmup 5 > gpurun_out/cfg5_lt${lt}_$ub.json 2>/dev/null
  done
done
DH_G_LEAN_T=0 DH_B200_LIB=$T/libdh_cur.so $B --workload cfg2 --steps 300 --warmup 20 > gpurun_out/cfg2_lt0.json 2>/dev/null
DH_G_LEAN_T=0 DH_B200_LIB=$T/libdh_cur.so $B --workload cfg1 --steps 300 --warmup 20 > gpurun_out/cfg1_lt0.json 2>/dev/null
python profiles/show_bench.py gpurun_out/cfg3_lt*.json gpurun_out/b384_lt*.json gpurun_out/cfg5_lt*.json gpurun_out/cfg[12]_lt0.json | awk '{print $1, $4, $5}'
