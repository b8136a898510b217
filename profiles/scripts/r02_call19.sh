# round 2, call 19: why did e2e (host entry, 4 pipelined chunks) not follow the kernel? PDL / heaviest-first on the chunk launches
set -x
B="python bench.py --no-cpu-baseline --no-side-legs"
T=profiles/trace_build
for v in "DH_G_PDL=1 DH_G_LPT=10" "DH_G_PDL=0 DH_G_LPT=10" "DH_G_PDL=1 DH_G_LPT=0" "DH_G_PDL=0 DH_G_LPT=0"; do
  n=$(echo $v | tr ' =' '__')
  env $v DH_B200_LIB=$T/libdh_cur.so $B --workload cfg5 --steps 40 --warmup 5 > gpurun_out/e2e_cfg5_$n.json 2>/dev/null
done
for f in gpurun_out/e2e_cfg5_DH*.json; do python - $f <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print(sys.argv[1].split('/')[-1], 'dev', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4))
PY
done
