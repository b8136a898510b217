# session 12: 2-GPU sanity run of both bench arms on the final kernels (torchrun, NCCL), as the driver launches them
set -x
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 100 --warmup 5 > gpurun_out/scale2_n2.json 2> gpurun_out/scale2_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/scale2_ref_n2.json 2> gpurun_out/scale2_ref_n2.err
python -m pytest tests -m gpu -q -k "two_device or 2gpu or two_gpu or multi" 2>&1 | tail -3
