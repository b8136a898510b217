# round 2, call 17: host entry (e2e): input-chunk growth factor and pipelining threshold
set -x
B="python bench.py --no-cpu-baseline --no-side-legs"
T=profiles/trace_build
for g in 4 8 16 32; do
  DH_HOST_GROWTH=$g DH_B200_LIB=$T/libdh_lean.so $B --workload cfg5 --steps 40 --warmup 5 > gpurun_out/e2e_cfg5_g$g.json 2>/dev/null
done
for kib in 1536 512 256; do
  DH_HOST_PIPELINE_KIB=$kib DH_B200_LIB=$T/libdh_lean.so $B --workload cfg5 --batch 384 --steps 200 --warmup 10 > gpurun_out/e2e_b384_k$kib.json 2>/dev/null
  DH_HOST_PIPELINE_KIB=$kib DH_HOST_GROWTH=8 DH_B200_LIB=$T/libdh_lean.so $B --workload cfg5 --batch 384 --steps 200 --warmup 10 > gpurun_out/e2e_b384_k${kib}_g8.json 2>/dev/null
done
DH_HOST_GATHER=0 DH_B200_LIB=$T/libdh_lean.so $B --workload cfg5 --batch 384 --steps 200 --warmup 10 > gpurun_out/e2e_b384_nogather.json 2>/dev/null
python profiles/show_bench.py gpurun_out/e2e_*.json | awk '{print $1, "dev", $4, "e2e", $13}'
