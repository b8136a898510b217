set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1
for w in cfg1 cfg2 cfg3 cfg4; do python bench.py --workload $w --steps 200 --warmup 10 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; done
python bench.py > gpurun_out/bench_cfg5_n1.json 2> gpurun_out/bench_cfg5.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_cfg5_n1.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:groups_kernel -s 4 -c 1 -o gpurun_out/groups_cfg5_full -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
