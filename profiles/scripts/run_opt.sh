python -m pytest tests/test_gpu_parity.py tests/test_lowp_planes.py tests/test_gpu_properties.py tests/test_gaussians_dali.py -m gpu -x -q 2>&1 | tail -5
python profiles/scripts/lowp_bench.py > gpurun_out/lowp_bench.jsonl 2> gpurun_out/lowp_bench.err; tail -3 gpurun_out/lowp_bench.err; cat gpurun_out/lowp_bench.jsonl
for w in cfg1 cfg2 cfg3 cfg4; do python bench.py --workload $w --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; tail -2 gpurun_out/bench_$w.err; python -c "
import json,sys; d=json.loads(open('gpurun_out/bench_$w.json').read().strip().splitlines()[-1]); print('$w', d['ms_per_step'], d.get('roofline',{}).get('cuda_graph_replay_ms_per_step'))"; done
