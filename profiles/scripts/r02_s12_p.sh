# session 12, call P: hand-out order built lazily (first wave outside the rearranged window) — tests + A/B on cfg3 / b384 / cfg5
python -m pytest tests/test_handout_order.py tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_lowp_planes.py tests/test_gaussians_dali.py -m gpu -q -x 2>&1 | tail -3 | tee gpurun_out/pytest_lazy.log
export DH_B200_LIB=profiles/trace_build/libdh_tune.so
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
for rep in 1 2 3; do for z in 0 1; do
  for wl in "--workload cfg3 --steps 400 --warmup 20" "--workload cfg5 --batch 384 --steps 100 --warmup 10" "--workload cfg5 --steps 30 --warmup 5"; do
    DH_G_PERM_LAZY=$z $B $wl 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('lazy $z', '$wl'.split()[1], '$wl'.split()[2], round(d['ms_per_step']*1000,2), 'us', d['digest'][2])" | tee -a gpurun_out/perm_lazy.txt
  done
done; done
