# round 2, final evidence run on the product build: tests, smoke, bench (both arms), ncu launch list of the bench command,
# ncu --set full of the dominant kernel on cfg5 / cfg3 / cfg4 / accumulate, side benches of the widened rows
set -x
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python bench.py > gpurun_out/bench_cfg5_n1.json 2> gpurun_out/bench_cfg5_n1.err
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
$B --steps 3 --warmup 3 > gpurun_out/plain_cfg5.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_cfg5_n1.csv $B --steps 3 --warmup 3 > gpurun_out/ncu_l.log 2>&1
$B --steps 3 --warmup 3 > gpurun_out/plain_cfg5b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:groups_kernel -s 4 -c 1 -o gpurun_out/final_groups_cfg5 $B --steps 3 --warmup 3 > gpurun_out/ncu_cfg5.log 2>&1
$B --workload cfg3 --steps 4 --warmup 3 > gpurun_out/plain_cfg3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:groups_kernel -s 4 -c 1 -o gpurun_out/final_groups_cfg3 $B --workload cfg3 --steps 4 --warmup 3 > gpurun_out/ncu_cfg3.log 2>&1
$B --workload cfg4 --steps 4 --warmup 3 > gpurun_out/plain_cfg4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:groups_kernel -s 4 -c 1 -o gpurun_out/final_groups_cfg4 $B --workload cfg4 --steps 4 --warmup 3 > gpurun_out/ncu_cfg4.log 2>&1
python profiles/scripts/accum_one.py > gpurun_out/plain_accum.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:groups_kernel -s 3 -c 1 -o gpurun_out/final_groups_accum_b1536 python profiles/scripts/accum_one.py > gpurun_out/ncu_accum.log 2>&1
python profiles/scripts/lowp_bench.py > gpurun_out/lowp_planes.jsonl 2>&1
python profiles/scripts/accum_bench.py > gpurun_out/accum_final.txt 2>&1
python profiles/scripts/next_rows_bench.py > gpurun_out/next_rows.jsonl 2>&1
ls -la gpurun_out/*.ncu-rep
