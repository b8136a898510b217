# session 12: with the next unit claimed late (truly dynamic hand-out), the knobs tuned under the old regime once more
export DH_B200_LIB=profiles/trace_build/libdh_tune.so
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
run() {  # label, env...
  label=$1; shift
  for wl in "--workload cfg3 --steps 400 --warmup 20" "--workload cfg5 --batch 384 --steps 100 --warmup 10" "--workload cfg5 --steps 30 --warmup 5"; do
    env "$@" $B $wl 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$label', '$wl'.split()[1], '$wl'.split()[2], round(d['ms_per_step']*1000,2), 'us')" | tee -a gpurun_out/retune.txt
  done
}
for rep in 1 2; do
  run base DH_X=0
  run early_claim DH_G_CLC_MODE=2
  run tail_div1 DH_G_TAIL_DIV=1
  run tail_div4 DH_G_TAIL_DIV=4
  run tail_waves4 DH_G_TAIL_WAVES_X2=4
  run lpt0 DH_G_LPT=0
  run lpt20 DH_G_LPT=20
  run unit128k DH_G_UNIT_BYTES=131072
  run unit96k DH_G_UNIT_BYTES=98304
done
