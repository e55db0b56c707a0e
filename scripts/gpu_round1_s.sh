set -x
timeout 600 python -m pytest tests/test_gpu_dist_nccl.py -m gpu -x -q > gpurun_out/pytest_nccl.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_nccl.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 1 --warmup 1 --prompts 1024 --batch-prompts 1024 --workload projection-vocab-shard > gpurun_out/bench_proj2.json 2> gpurun_out/bench_proj2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 1 --warmup 1 --prompts 2048 --batch-prompts 2048 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > gpurun_out/bench_ref2.json 2> gpurun_out/bench_ref2.err
