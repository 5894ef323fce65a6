#!/bin/bash
mkdir -p gpurun_out
export LT_B200_K2FWD=cluster
for lib in "" $PWD/learnable-triangulation-pytorch_b200/variants/k2c_nocoord.so; do
LT_B200_LIB=$lib timeout 400 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/ab.json 2> gpurun_out/ab.err
python - "$lib" <<'PY'
import json, sys
d=json.loads(open("gpurun_out/ab.json").read().strip().splitlines()[-1])
k=d["roofline"]["kernels"]
print((sys.argv[1].split('/')[-1] or 'cluster'), "K2 fwd %.4f ms" % k["k2_softargmax_fwd"]["ms"])
PY
done
