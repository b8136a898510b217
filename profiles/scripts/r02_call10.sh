# round 2, call 10: heaviest-first hand-out restricted to the last k/2 waves
set -x
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
for lpt in 0 2 3 4 6 10; do
  DH_G_LPT=$lpt $B --workload cfg3 --steps 200 --warmup 20 > gpurun_out/cfg3_lpt$lpt.json 2>/dev/null
  DH_G_LPT=$lpt $B --workload cfg5 --batch 384 --steps 100 --warmup 10 > gpurun_out/b384_lpt$lpt.json 2>/dev/null
  DH_G_LPT=$lpt $B --workload cfg5 --steps 30 --warmup 5 > gpurun_out/cfg5_lpt$lpt.json 2>/dev/null
done
python profiles/show_bench.py gpurun_out/cfg3_lpt*.json gpurun_out/b384_lpt*.json gpurun_out/cfg5_lpt*.json | awk '{print $1, $4, $5}'
