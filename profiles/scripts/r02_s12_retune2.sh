# session 12: tail tier under late claiming — size of its units (DIV), its length (WAVES_X2) and its share of the launch (CAP_DIV)
export DH_B200_LIB=profiles/trace_build/libdh_tune.so
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
run() {
  label=$1; shift
  for wl in "--workload cfg3 --steps 400 --warmup 20" "--workload cfg5 --batch 96 --steps 200 --warmup 10" "--workload cfg5 --batch 192 --steps 100 --warmup 10" "--workload cfg5 --batch 384 --steps 100 --warmup 10" "--workload cfg5 --steps 30 --warmup 5"; do
    env "$@" $B $wl 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$label', '$wl'.split()[1], '$wl'.split()[2], '$wl'.split()[3], round(d['ms_per_step']*1000,2), 'us')" | tee -a gpurun_out/retune2.txt
  done
}
for rep in 1 2; do
  run div2_cap2 DH_X=0
  run div4_cap2 DH_G_TAIL_DIV=4
  run div2_cap8 DH_G_TAIL_CAP_DIV=8
  run div4_cap8 DH_G_TAIL_DIV=4 DH_G_TAIL_CAP_DIV=8
  run div4_cap16 DH_G_TAIL_DIV=4 DH_G_TAIL_CAP_DIV=16
  run div4_w1 DH_G_TAIL_DIV=4 DH_G_TAIL_WAVES_X2=1
  run div1 DH_G_TAIL_DIV=1
done
