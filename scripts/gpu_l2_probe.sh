set -x
python scripts/lens_l2_probe.py > gpurun_out/l2probe_plain.log 2>&1 &&
ncu --metrics dram__bytes_read.sum,lts__t_sector_hit_rate.pct,gpu__time_duration.sum,l1tex__m_xbar2l1tex_read_bytes.sum,sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.per_cycle_elapsed,gpc__cycles_elapsed.avg.per_second,lts__t_bytes.sum --clock-control none --kernel-name-base demangled -k regex:LensEpi --csv --log-file gpurun_out/l2probe.csv python scripts/lens_l2_probe.py > gpurun_out/l2probe_ncu.log 2>&1
