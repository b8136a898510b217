# session 12, call V: FRESH register path with one offset per chunk (digest index included) — tests + default bench (e2e) + b384 e2e
python -m pytest tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_lowp_planes.py tests/test_handout_order.py tests/test_two_render_paths.py -m gpu -q -x 2>&1 | tail -3 | tee gpurun_out/pytest_off.log
B="python bench.py --no-cpu-baseline --no-side-legs"
for rep in 1 2; do
  for wl in "--steps 50 --warmup 5" "--batch 384 --steps 100 --warmup 10" "--workload cfg3 --steps 400 --warmup 20"; do
    $B $wl 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$wl', round(d['ms_per_step']*1000,2), 'us  e2e', round(d['e2e']['ms_per_step']*1000,2), 'us')" | tee -a gpurun_out/off.txt
  done
done
python profiles/scripts/lowp_bench.py 2>&1 | head -3 | cut -c1-130 | tee -a gpurun_out/off.txt
