set -x
python -m pytest tests/test_gpu_properties.py -m gpu -x -q -k "host" > gpurun_out/pytest_host.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_host.log
python bench.py --no-cpu-baseline > gpurun_out/e2e_cfg5.json 2> gpurun_out/e2e_cfg5.err
python bench.py --no-cpu-baseline --batch 384 --steps 300 > gpurun_out/e2e_cfg5_b384.json 2>> gpurun_out/e2e_cfg5.err
for ub in 32768 49152 65536 98304 131072; do
  for w in cfg2 cfg3; do
    DH_G_UNIT_BYTES=$ub python bench.py --workload $w --steps 300 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/ub_${w}_$ub.json 2>> gpurun_out/ub.err
  done
done
