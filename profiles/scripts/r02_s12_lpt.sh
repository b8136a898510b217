# session 12: heaviest-first window (DH_G_LPT = 2 x waves) once more under late claiming
export DH_B200_LIB=profiles/trace_build/libdh_tune.so
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
for rep in 1 2; do for z in 10 6 14; do
  for wl in "--workload cfg5 --batch 192 --steps 100 --warmup 10" "--workload cfg5 --batch 384 --steps 100 --warmup 10" "--workload cfg5 --steps 30 --warmup 5"; do
    DH_G_LPT=$z $B $wl 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('lpt $z', '$wl'.split()[1], '$wl'.split()[2], '$wl'.split()[3], round(d['ms_per_step']*1000,2), 'us')" | tee -a gpurun_out/lpt_window.txt
  done
done; done
