# session 12, call F: ncu --set full of the accumulate launch (half of cfg5) and of the bf16 launch (cfg5) on the current kernels
set -x
python profiles/scripts/accum_one.py > gpurun_out/plain_accum.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:groups_kernel -s 3 -c 1 -o gpurun_out/s12_accum_b1536 python profiles/scripts/accum_one.py > gpurun_out/ncu_accum.log 2>&1
python profiles/scripts/lowp_one.py > gpurun_out/plain_lowp.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:groups_kernel -s 3 -c 1 -o gpurun_out/s12_bf16_cfg5 python profiles/scripts/lowp_one.py > gpurun_out/ncu_lowp.log 2>&1
ls -la gpurun_out/*.ncu-rep
