# session 12, call G: ACCUMULATE band size A/B (groups per band striped over the warps)
python -m pytest tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_gpu_python_api.py tests/test_guard_bands.py tests/test_handout_order.py -m gpu -q -x 2>&1 | tail -2
for v in product ab1 ab3 ab4 product ab1; do
  if [ $v = product ]; then unset DH_B200_LIB; else export DH_B200_LIB=profiles/trace_build/libdh_$v.so; fi
  echo "== $v" | tee -a gpurun_out/accum_band.txt
  python profiles/scripts/accum_bench.py 2>&1 | head -3 | tee -a gpurun_out/accum_band.txt
done
