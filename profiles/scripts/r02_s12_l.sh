# session 12, call L: unit height of dense units on cfg4 (DH_G_HEAVY_PER_SLOT: units per resident CTA slot when T >= 256)
export DH_B200_LIB=profiles/trace_build/libdh_tune.so
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs --workload cfg4 --steps 200 --warmup 10"
for rep in 1 2; do for v in 8 4 3 2 1; do
  DH_G_HEAVY_PER_SLOT=$v $B 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('per_slot $v', d['ms_per_step'])" | tee -a gpurun_out/dense_unit_rows.txt
done; done
