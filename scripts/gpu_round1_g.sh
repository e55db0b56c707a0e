set -x
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_dist_nccl.py -m gpu -x -q > gpurun_out/pytest_nccl.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_nccl.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 1 --warmup 1 --prompts 2048 --batch-prompts 2048 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err
python bench.py --gpus 1 --steps 1 --warmup 1 --prompts 2048 --batch-prompts 2048 --no-cpu-baseline > gpurun_out/bench_1gpu_2048.json 2> gpurun_out/bench_1gpu_2048.err
