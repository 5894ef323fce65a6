#!/bin/bash
# round 2j: ncu --set full of the TMA forward kernel
mkdir -p gpurun_out
export LT_B200_K1FWD=tma
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_unproject_fwd_tma -c 1 -o gpurun_out/r2j_k1fwd_tma python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/r2j_ncu.log 2>&1
ls -la gpurun_out/*.ncu-rep
