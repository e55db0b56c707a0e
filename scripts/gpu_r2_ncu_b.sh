#!/bin/bash
# ncu --set full of the lens kernel with the logits store and of the tiled column-sum kernel (second invocation of each)
mkdir -p gpurun_out
python scripts/dense_ll_ncu.py > gpurun_out/plain_dense.log 2>&1 && cat gpurun_out/plain_dense.log &&
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"LensDenseEpi|colsum_tiled" -s 2 -c 2 -o gpurun_out/r02z_dense_ll python scripts/dense_ll_ncu.py > gpurun_out/ncu_dense.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_dense.log; ls -la gpurun_out | grep r02z_dense
