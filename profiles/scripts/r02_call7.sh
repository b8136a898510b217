# round 2, call 7: full GPU test-suite on the current defaults + the full bench line (all legs) + smoke
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
( time python bench.py > gpurun_out/bench_cfg5_n1.json 2> gpurun_out/bench_cfg5_n1.err ) 2>&1 | tail -3
tail -3 gpurun_out/bench_cfg5_n1.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_cfg5_n1.json') if l.startswith('{')][-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'frac', d['roofline']['frac'], 'e2e', d['e2e']['ms_per_step'], 'launches', d['gpu_launches'])
for k in ('configs','accumulate','reference_gpu'):
    print(k, json.dumps(d.get(k))[:1500])
PY
