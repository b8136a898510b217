# session 12, call Y: first-wave mix of the hand-out order (one heavy unit per SM, lighter ones beside it) — A/B, same box
export DH_B200_LIB=profiles/trace_build/libdh_tune.so
python -m pytest tests/test_handout_order.py tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -2 | tee gpurun_out/pytest_mix.log
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
for rep in 1 2 3; do for z in 0 1; do
  for wl in "--workload cfg3 --steps 400 --warmup 20" "--workload cfg5 --batch 48 --steps 300 --warmup 10" "--workload cfg5 --batch 96 --steps 200 --warmup 10"; do
    DH_G_PERM_MIX=$z $B $wl 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('mix $z', '$wl'.split()[1], '$wl'.split()[2], '$wl'.split()[3], round(d['ms_per_step']*1000,2), 'us', d['digest'][2])" | tee -a gpurun_out/perm_mix.txt
  done
done; done
