# round 2, call 6: is it the write window? unit-size sweep for the single-launch renderer and for groups2 with / without
# a per-unit block barrier (8-GPU slice of cfg5, cfg3)
set -x
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
T=profiles/trace_build
for ub in 196608 131072 98304 65536 49152; do
  for v in g2sync g2free; do
    DH_G_UNIT_BYTES=$ub DH_B200_LIB=$T/libdh_$v.so $B --workload cfg5 --batch 384 --steps 100 --warmup 10 > gpurun_out/b384_${v}_$ub.json 2>/dev/null
    DH_G_UNIT_BYTES=$ub DH_B200_LIB=$T/libdh_$v.so $B --workload cfg3 --steps 200 --warmup 20 > gpurun_out/cfg3_${v}_$ub.json 2>/dev/null
  done
  DH_G_NOBIN=1 DH_G_UNIT_BYTES=$ub DH_B200_LIB=$T/libdh_g2free.so $B --workload cfg5 --batch 384 --steps 100 --warmup 10 > gpurun_out/b384_legacy_$ub.json 2>/dev/null
  DH_G_NOBIN=1 DH_G_UNIT_BYTES=$ub DH_B200_LIB=$T/libdh_g2free.so $B --workload cfg3 --steps 200 --warmup 20 > gpurun_out/cfg3_legacy_$ub.json 2>/dev/null
done
for f in gpurun_out/b384_*_*.json gpurun_out/cfg3_*_*.json; do python - "$f" <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print(f"{sys.argv[1].split('/')[-1]:34s} {d['ms_per_step']*1e3:8.1f} us  {d['roofline']['achieved']:7.0f} GB/s")
PY
done
