#!/bin/bash
mkdir -p gpurun_out
for v in fused split; do
  export LT_B200_K2FWD=$v
  python -m pytest tests/test_gpu_softargmax.py tests/test_gpu_coordspec.py tests/test_gpu_guard_bands.py tests/test_gpu_fuzz_shapes.py tests/test_gpu_unproject_tile.py -m gpu -q -x 2>&1 | tail -2
  python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2f_bench_$v.json 2> gpurun_out/r2f_bench_$v.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2f_bench_$v.json"))
    k=d["roofline"]["kernels"]
    print("$v", round(d["value"]), {n:k[n]["ms"] for n in k})
except Exception as e: print("$v failed", e)
PY
done
