#!/bin/bash
mkdir -p gpurun_out
bash tools/gpu_variants.sh 2>&1 | tee gpurun_out/r2z_variants.log
