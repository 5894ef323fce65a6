#!/bin/bash
mkdir -p gpurun_out
export LT_B200_K2FWD=fused
python -m pytest tests/test_gpu_softargmax.py tests/test_gpu_coordspec.py -m gpu -q -x 2>&1 | tail -1
for lag in 1 2 3 5 15; do
  export LT_B200_K2LAG=$lag
  python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2g_bench_$lag.json 2> gpurun_out/r2g_bench_$lag.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2g_bench_$lag.json"))
    k=d["roofline"]["kernels"]
    print("lag $lag", round(d["value"]), k["k2_softargmax_fwd"]["ms"], k["k2_softargmax_bwd"]["ms"])
except Exception as e: print("$lag failed", e)
PY
done
export LT_B200_K2LAG=2
ncu --set full --import-source on --clock-control none -k regex:k_sam_fwd_fused -c 1 -o gpurun_out/r2g_k2fused python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/r2g_ncu.log 2>&1
