# session 12, call C: tile geometry A/B on cfg4 (rows per strip x resident CTAs per SM), measurement builds
B="python bench.py --no-cpu-baseline --no-e2e --no-side-legs --workload cfg4 --steps 200 --warmup 10"
for v in r8m4 r8m3 r12m3 r16m3 r16m4 r8m4 r16m3; do
  DH_B200_LIB=profiles/trace_build/libdh_$v.so $B 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', d['ms_per_step'], d['digest'])" | tee -a gpurun_out/tiles_geometry.txt
done
DH_B200_LIB=profiles/trace_build/libdh_r16m3.so python -m pytest tests/test_tiles_path.py -m gpu -q 2>&1 | tail -3 | tee -a gpurun_out/tiles_geometry.txt
