set -x
nvidia-smi -L | wc -l
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --steps 1 --warmup 1 > gpurun_out/bench_8gpu.json 2> gpurun_out/bench_8gpu.err
echo "rc=$?"
