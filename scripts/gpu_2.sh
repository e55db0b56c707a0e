# 2 GPUs: NCCL parity test, the new sweep test, the default bench (weak scaling) and the vocabulary-sharded projection bench
set -x
nvidia-smi -L | wc -l
timeout 600 python -m pytest tests/test_gpu_dist_nccl.py tests/test_gpu_sweeps.py -m gpu -x -q > gpurun_out/pytest_2gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_2gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 2 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 2 --steps 1 --warmup 1 --no-cpu-baseline --workload projection-vocab-shard --prompts 1024 > gpurun_out/bench_proj2.json 2> gpurun_out/bench_proj2.err
timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --workload projection-vocab-shard --prompts 1024 > gpurun_out/bench_proj1.json 2> gpurun_out/bench_proj1.err
