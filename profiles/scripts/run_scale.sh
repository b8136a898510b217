set -x
for n in 1 2 4 8; do
  if [ $n = 1 ]; then python bench.py > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2950$n bench.py --gpus $n --steps 100 --warmup 5 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err; fi
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --impl reference --steps 3 --warmup 1 > gpurun_out/scale_ref_n8.json 2> gpurun_out/scale_ref_n8.err
