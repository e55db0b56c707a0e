# same SM count (120), different operand sharing: 2x1 clusters limited to 30 vs 15 clusters of 2x2
TB_MAX_CLUSTERS=30 python scripts/lens_sweep.py --rows 75776 --seconds 3 --modes 2 --chunks 2,4 --no-cublas > gpurun_out/lens_sweep_g.jsonl 2> gpurun_out/lens_sweep_g.err
python scripts/lens_sweep.py --rows 75776 --seconds 3 --modes 4 --chunks 2,4 --no-cublas >> gpurun_out/lens_sweep_g.jsonl 2>> gpurun_out/lens_sweep_g.err
TB_MAX_CLUSTERS=60 python scripts/lens_sweep.py --rows 75776 --seconds 3 --modes 1 --chunks 2,4 --no-cublas >> gpurun_out/lens_sweep_g.jsonl 2>> gpurun_out/lens_sweep_g.err
