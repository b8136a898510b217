# session 12, call T: occupancy mask as two 32-bit shared atomics — tests + cfg3 / b384 / cfg5 / accumulate
python -m pytest tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_handout_order.py tests/test_tiles_path.py tests/test_lowp_planes.py -m gpu -q -x 2>&1 | tail -3 | tee gpurun_out/pytest_mask.log
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
for rep in 1 2; do
  for wl in "--workload cfg3 --steps 400 --warmup 20" "--workload cfg5 --batch 384 --steps 100 --warmup 10" "--workload cfg5 --steps 30 --warmup 5" "--workload cfg1 --steps 400 --warmup 20"; do
    $B $wl 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$wl'.split()[1], '$wl'.split()[2], round(d['ms_per_step']*1000,2), 'us')" | tee -a gpurun_out/mask32.txt
  done
done
python profiles/scripts/accum_bench.py 2>&1 | head -3 | tee -a gpurun_out/mask32.txt
