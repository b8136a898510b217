# round 2, call 5: cfg4 — row-fused sweep vs round-1 sweep (bench A/B), ncu full of both on the single-launch renderer;
# accumulate path: binned vs single-launch
set -x
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
T=profiles/trace_build
for v in swold swnew; do
  DH_G_NOBIN=1 DH_B200_LIB=$T/libdh_$v.so $B --workload cfg4 --steps 100 --warmup 10 > gpurun_out/cfg4_${v}_nobin.json 2>/dev/null
  DH_B200_LIB=$T/libdh_$v.so $B --workload cfg4 --steps 100 --warmup 10 > gpurun_out/cfg4_${v}_bin.json 2>/dev/null
  DH_G_NOBIN=1 DH_B200_LIB=$T/libdh_$v.so $B --workload cfg3 --steps 200 --warmup 20 > gpurun_out/cfg3_${v}_nobin.json 2>/dev/null
done
python profiles/show_bench.py gpurun_out/cfg4_sw*.json gpurun_out/cfg3_sw*.json
for v in swold swnew; do
  DH_G_NOBIN=1 DH_B200_LIB=$T/libdh_$v.so $B --workload cfg4 --steps 3 --warmup 3 > gpurun_out/plain_$v.log 2>&1 &&
  DH_G_NOBIN=1 DH_B200_LIB=$T/libdh_$v.so ncu --set full --clock-control none --import-source on -k regex:groups_kernel -s 3 -c 1 -o gpurun_out/cfg4_$v $B --workload cfg4 --steps 3 --warmup 3 > gpurun_out/ncu_$v.log 2>&1
done
DH_B200_LIB=$T/libdh_swnew.so python profiles/scripts/accum_bench.py > gpurun_out/accum_bin.txt 2>&1
DH_G_NOBIN=1 DH_B200_LIB=$T/libdh_swnew.so python profiles/scripts/accum_bench.py > gpurun_out/accum_nobin.txt 2>&1
cat gpurun_out/accum_bin.txt gpurun_out/accum_nobin.txt
