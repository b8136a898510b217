# round 2, call 11: heaviest-first window, wider values; side check on cfg2 / cfg4 / accumulate
set -x
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
for lpt in 0 8 10 12 16 24; do
  DH_G_LPT=$lpt $B --workload cfg3 --steps 300 --warmup 20 > gpurun_out/cfg3_lpt$lpt.json 2>/dev/null
  DH_G_LPT=$lpt $B --workload cfg5 --batch 384 --steps 150 --warmup 10 > gpurun_out/b384_lpt$lpt.json 2>/dev/null
  DH_G_LPT=$lpt $B --workload cfg5 --steps 40 --warmup 5 > gpurun_out/cfg5_lpt$lpt.json 2>/dev/null
done
python profiles/show_bench.py gpurun_out/cfg3_lpt*.json gpurun_out/b384_lpt*.json gpurun_out/cfg5_lpt*.json | awk '{print $1, $4, $5}'
for lpt in 0 10; do DH_G_LPT=$lpt python profiles/scripts/accum_bench.py 2>&1 | head -3; done
