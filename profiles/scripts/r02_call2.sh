# round 2, call 2: K2 pre-pass + warp-autonomous renderer (groups2) vs the single-launch renderer, all configs
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for nobin in 1 0; do
  for cfg in cfg1 cfg2 cfg3 cfg4; do
    DH_G_NOBIN=$nobin python bench.py --workload $cfg --steps 200 --warmup 20 --no-cpu-baseline --no-e2e > gpurun_out/${cfg}_nobin$nobin.json 2> gpurun_out/${cfg}_nobin$nobin.err
  done
  DH_G_NOBIN=$nobin python bench.py --workload cfg5 --batch 384 --steps 100 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/b384_nobin$nobin.json 2> gpurun_out/b384_nobin$nobin.err
  DH_G_NOBIN=$nobin python bench.py --workload cfg5 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/cfg5_nobin$nobin.json 2> gpurun_out/cfg5_nobin$nobin.err
done
python profiles/show_bench.py gpurun_out/*_nobin*.json
