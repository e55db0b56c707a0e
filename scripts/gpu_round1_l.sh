set -x
for ls in 0 1 2 3; do
TB_LOCKSTEP=$ls timeout 600 python scripts/lens_sweep.py --rows 75776 --seconds 2.5 --chunks 4,8 --modes 4,2 --no-cublas > gpurun_out/lens_sweep_ls$ls.jsonl 2> gpurun_out/lens_sweep_ls$ls.err
done
