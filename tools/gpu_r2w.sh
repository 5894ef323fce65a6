#!/bin/bash
# round 2u: ncu --set full of the default K1 forward and backward kernels (source-level instruction mix)
mkdir -p gpurun_out
timeout 600 ncu --set full --import-source on --clock-control none -k regex:"k_unproject_fwd|k_unproject_bwd" -c 2 -o gpurun_out/r2w_k1 python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > gpurun_out/r2w_ncu.log 2>&1
ls -la gpurun_out/r2w_k1.ncu-rep
