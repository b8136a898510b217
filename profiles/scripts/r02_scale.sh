# round 2: 1 / 2 / 4 / 8 GPU sweep of the default bench (cfg5 sharded by samples), one 8-GPU box
set -x
python bench.py --no-cpu-baseline --no-side-legs > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2950$n bench.py --gpus $n --steps 100 --warmup 5 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
done
for n in 1 2 4 8; do python - gpurun_out/scale_n$n.json <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print(d['n_gpus'], 'ms', round(d['ms_per_step'],4), 'planes/s', round(d['value']/1e6,2), 'M  e2e ms', round(d['e2e']['ms_per_step'],4), 'e2e planes/s', round(d['e2e']['value']/1e6,2), 'M digest', d['digest'], 'clk', d['clocks']['sm_mhz'])
PY
done
