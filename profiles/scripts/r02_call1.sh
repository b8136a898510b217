# round 2, call 1: new parity tests + PDL / lazy-prologue A/B on cfg3, the 8-GPU slice and cfg5
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
for pdl in 0 1; do
  DH_G_PDL=$pdl python bench.py --workload cfg3 --steps 200 --warmup 20 --no-cpu-baseline --no-e2e > gpurun_out/cfg3_pdl$pdl.json 2> gpurun_out/cfg3_pdl$pdl.err
  DH_G_PDL=$pdl python bench.py --workload cfg5 --batch 384 --steps 100 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/b384_pdl$pdl.json 2> gpurun_out/b384_pdl$pdl.err
  DH_G_PDL=$pdl python bench.py --workload cfg5 --steps 30 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/cfg5_pdl$pdl.json 2> gpurun_out/cfg5_pdl$pdl.err
done
python profiles/show_bench.py gpurun_out/*_pdl*.json
