#!/bin/bash
# round 2x: K1 backward with the tap loads of 2 / 4 views in flight before the first blend, 2 or 3 CTAs per SM (variants)
mkdir -p gpurun_out
bash tools/gpu_variants.sh 2>&1 | tee gpurun_out/r2x_variants.log
for v in b4c2 b2c2; do
LT_B200_LIB=$PWD/learnable-triangulation-pytorch_b200/variants/$v.so LT_B200_TORCH_EXT=0 timeout 600 python -m pytest tests/test_gpu_unproject.py tests/test_gpu_fuzz_shapes.py -x -q -m gpu 2>&1 | tail -1
done
