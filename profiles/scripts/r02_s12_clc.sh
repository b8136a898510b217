# session 12: when a CTA claims its next unit through cluster launch control (DH_G_CLC_MODE), same box, interleaved
export DH_B200_LIB=profiles/trace_build/libdh_tune.so
python -m pytest tests/test_handout_order.py tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -1
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
for rep in 1 2; do for z in 1 2 3 5 7 11; do
  for wl in "--workload cfg3 --steps 400 --warmup 20" "--workload cfg5 --batch 384 --steps 100 --warmup 10" "--workload cfg5 --steps 30 --warmup 5"; do
    DH_G_CLC_MODE=$z $B $wl 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('clc_mode $z', '$wl'.split()[1], '$wl'.split()[2], round(d['ms_per_step']*1000,2), 'us', d['digest'][2])" | tee -a gpurun_out/clc_mode.txt
  done
done; done
