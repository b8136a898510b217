# round 2, call 9: heaviest-first hand-out (in-kernel counting sort of the samples by object count) on/off
set -x
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs"
python -m pytest tests/test_gpu_parity.py tests/test_two_render_paths.py tests/test_guard_bands.py tests/test_lowp_planes.py tests/test_gpu_properties.py tests/test_gaussians_dali.py -m gpu -x -q 2>&1 | tail -3
for lpt in 0 1; do
  for cfg in cfg2 cfg3 cfg4; do
    DH_G_LPT=$lpt $B --workload $cfg --steps 200 --warmup 20 > gpurun_out/${cfg}_lpt$lpt.json 2>/dev/null
  done
  DH_G_LPT=$lpt $B --workload cfg5 --batch 384 --steps 100 --warmup 10 > gpurun_out/b384_lpt$lpt.json 2>/dev/null
  DH_G_LPT=$lpt $B --workload cfg5 --batch 768 --steps 100 --warmup 10 > gpurun_out/b768_lpt$lpt.json 2>/dev/null
  DH_G_LPT=$lpt $B --workload cfg5 --steps 30 --warmup 5 > gpurun_out/cfg5_lpt$lpt.json 2>/dev/null
done
python profiles/show_bench.py gpurun_out/*_lpt0.json gpurun_out/*_lpt1.json
