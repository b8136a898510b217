python -m pytest tests/test_focal_loss.py tests/test_new_rows_edges.py -m gpu -x -q 2>&1 | tail -5
DH_ROWS=N3 python profiles/scripts/next_rows_bench.py > gpurun_out/n3_bench.jsonl 2> gpurun_out/n3_bench.err || tail -5 gpurun_out/n3_bench.err
cat gpurun_out/n3_bench.jsonl
