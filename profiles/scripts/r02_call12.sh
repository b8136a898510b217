# round 2, call 12: ncu --set full of the current groups_kernel: cfg3 fresh, 8-GPU slice of cfg5 fresh, cfg4, accumulate
set -x
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
$B --workload cfg3 --steps 4 --warmup 3 > gpurun_out/plain_cfg3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:groups_kernel -s 4 -c 1 -o gpurun_out/r02_groups_cfg3 $B --workload cfg3 --steps 4 --warmup 3 > gpurun_out/ncu_cfg3.log 2>&1
$B --workload cfg5 --batch 384 --steps 4 --warmup 3 > gpurun_out/plain_b384.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:groups_kernel -s 4 -c 1 -o gpurun_out/r02_groups_b384 $B --workload cfg5 --batch 384 --steps 4 --warmup 3 > gpurun_out/ncu_b384.log 2>&1
$B --workload cfg4 --steps 4 --warmup 3 > gpurun_out/plain_cfg4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:groups_kernel -s 4 -c 1 -o gpurun_out/r02_groups_cfg4 $B --workload cfg4 --steps 4 --warmup 3 > gpurun_out/ncu_cfg4.log 2>&1
python profiles/scripts/accum_one.py > gpurun_out/plain_accum.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:groups_kernel -s 3 -c 1 -o gpurun_out/r02_groups_accum_b1536 python profiles/scripts/accum_one.py > gpurun_out/ncu_accum.log 2>&1
ls -la gpurun_out/*.ncu-rep
