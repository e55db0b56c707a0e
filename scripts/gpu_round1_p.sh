set -x
for a in 0.5 0.7 0.85; do
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --prompts 2048 --batch-prompts 2048 --ll-alpha $a > gpurun_out/bench_alpha$a.json 2> gpurun_out/bench_alpha$a.err
done
