# session 12, call R: scan bound follows the sample size for slot arrays wider than one block step — tests + padded-T probe + cfg4
python -m pytest tests/test_tiles_path.py tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_round2_surface.py tests/test_new_rows_edges.py -m gpu -q -x 2>&1 | tail -3 | tee gpurun_out/pytest_scan.log
python profiles/scripts/padded_t_probe.py 2>&1 | tee gpurun_out/padded_t_scan.txt
python bench.py --no-cpu-baseline --no-side-legs --no-e2e --workload cfg4 --steps 200 --warmup 10 > gpurun_out/plain_cfg4.json 2>/dev/null
