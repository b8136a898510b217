# session 12, call E: hand-out order A/B (DH_G_PERM_ZIP: percent of the reordered window handed out heavy/light alternately)
export DH_B200_LIB=profiles/trace_build/libdh_tune.so
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
for rep in 1 2; do
for z in 0 20 40 60 100; do
  for wl in "--workload cfg3 --steps 400 --warmup 20" "--workload cfg5 --batch 384 --steps 100 --warmup 10"; do
    DH_G_PERM_ZIP=$z $B $wl 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('zip $z', '$wl'.split()[1], '$wl'.split()[2], round(d['ms_per_step']*1000,2), 'us', d['digest'][2])" | tee -a gpurun_out/perm_zip.txt
  done
done
done
