# session 12, call I: per-unit dense renderer inside groups_kernel<..., DENSE>: tests, padded-T probe, cfg4, threshold A/B
set -x
python -m pytest tests/test_tiles_path.py tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_round2_surface.py -m gpu -q -x 2>&1 | tail -4 | tee gpurun_out/pytest_dense.log
python profiles/scripts/padded_t_probe.py 2>&1 | tee gpurun_out/padded_t_dense.txt
for m in 16 32 48 96 200; do
  echo "== DH_G_DENSE_MIN=$m" | tee -a gpurun_out/padded_t_dense.txt
  DH_B200_LIB=profiles/trace_build/libdh_tune.so DH_G_DENSE_MIN=$m python profiles/scripts/padded_t_probe.py 2>&1 | grep -E "cfg4|\"T\": 256" | tee -a gpurun_out/padded_t_dense.txt
done
python bench.py --no-cpu-baseline --no-e2e --no-side-legs --workload cfg4 --steps 200 --warmup 10 > gpurun_out/plain_cfg4.json 2>/dev/null
