# round 2, call 22: accumulate: request the existing pixel before the arithmetic
set -x
T=profiles/trace_build
for v in accbase accload; do
  DH_B200_LIB=$T/libdh_$v.so python -m pytest tests/test_gpu_parity.py tests/test_gpu_properties.py -m gpu -x -q 2>&1 | tail -1
  DH_B200_LIB=$T/libdh_$v.so python profiles/scripts/accum_bench.py > gpurun_out/accum_$v.txt 2>&1
  cat gpurun_out/accum_$v.txt
done
