#!/bin/bash
# round 2ad: ncu --set full + source of the tile backward (tile shape 1: 8x4x4, 2 CTAs/SM)
mkdir -p gpurun_out
export LT_B200_K1BWD=tile LT_B200_K1BWD_TILE=1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_unproject_bwd_tile -c 1 -o gpurun_out/r2ad_tile python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > gpurun_out/r2ad_ncu.log 2>&1
ls -la gpurun_out/r2ad_tile.ncu-rep
