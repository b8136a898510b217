python -m pytest tests/test_lowp_planes.py -m gpu -x -q 2>&1 | tail -15
python profiles/scripts/lowp_bench.py > gpurun_out/lowp_bench.jsonl 2> gpurun_out/lowp_bench.err; tail -3 gpurun_out/lowp_bench.err; cat gpurun_out/lowp_bench.jsonl
