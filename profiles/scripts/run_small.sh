python -m pytest tests/test_gpu_properties.py -m gpu -x -q -k "full_size" 2>&1 | tail -5
for w in cfg1 cfg2 cfg3 cfg4; do python bench.py --workload $w --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; tail -2 gpurun_out/bench_$w.err; done
