# session 12, call U: 32-bit against 64-bit unit-index divisions, same box, interleaved (measurement builds)
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
for rep in 1 2 3; do for v in div64 div32; do
  for wl in "--workload cfg3 --steps 400 --warmup 20" "--workload cfg5 --batch 384 --steps 100 --warmup 10" "--workload cfg5 --steps 30 --warmup 5"; do
    DH_B200_LIB=profiles/trace_build/libdh_$v.so $B $wl 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', '$wl'.split()[1], '$wl'.split()[2], round(d['ms_per_step']*1000,2), 'us')" | tee -a gpurun_out/div32.txt
  done
done; done
for v in div64 div32 div64 div32; do echo "== $v"; DH_B200_LIB=profiles/trace_build/libdh_$v.so python profiles/scripts/accum_bench.py 2>&1 | head -2; done | tee -a gpurun_out/div32.txt
