# round 2, call 3: where does the two-launch path lose time? per-kernel durations (ncu launch list) on the 8-GPU slice and cfg3
set -x
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
$B --workload cfg5 --batch 384 --steps 6 --warmup 3 > gpurun_out/plain_b384.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_b384_bin.csv $B --workload cfg5 --batch 384 --steps 6 --warmup 3 > gpurun_out/ncu1.log 2>&1
DH_G_NOBIN=1 $B --workload cfg5 --batch 384 --steps 6 --warmup 3 > gpurun_out/plain_b384_nobin.log 2>&1 &&
DH_G_NOBIN=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 30 --csv --log-file gpurun_out/launches_b384_nobin.csv $B --workload cfg5 --batch 384 --steps 6 --warmup 3 > gpurun_out/ncu2.log 2>&1
$B --workload cfg3 --steps 6 --warmup 3 > gpurun_out/plain_cfg3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_cfg3_bin.csv $B --workload cfg3 --steps 6 --warmup 3 > gpurun_out/ncu3.log 2>&1
# same binary, PDL off: does the dependent-launch chain itself cost?
for pdl in 0 1; do
  DH_G_PDL=$pdl $B --workload cfg5 --batch 384 --steps 100 --warmup 10 > gpurun_out/b384_bin_pdl$pdl.json 2>/dev/null
  DH_G_PDL=$pdl $B --workload cfg3 --steps 200 --warmup 20 > gpurun_out/cfg3_bin_pdl$pdl.json 2>/dev/null
done
python profiles/show_bench.py gpurun_out/*_bin_pdl*.json
grep -h "groups\|bin_kernel" gpurun_out/launches_b384_bin.csv | awk -F, '{print $5, $NF}' | tail -12
grep -h "groups" gpurun_out/launches_b384_nobin.csv | awk -F, '{print $5, $NF}' | tail -6
grep -h "groups\|bin_kernel" gpurun_out/launches_cfg3_bin.csv | awk -F, '{print $5, $NF}' | tail -12
