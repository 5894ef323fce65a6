#!/bin/bash
# round 2p (8 consumer warps x 4 ch, all fixes): per-role wait accounting of the TMA forward kernel (library built with -DLT_TMA_PROF)
mkdir -p gpurun_out
export LT_B200_K1FWD=tma
timeout 600 python -m pytest tests/test_gpu_unproject.py tests/test_gpu_coordspec.py tests/test_gpu_fuzz_shapes.py -x -q -m gpu 2>&1 | tail -2
for d in 0 1 2; do
export LT_B200_TMA_DBG=$d
timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2q_bench_$d.json 2> gpurun_out/r2q_bench_$d.err
grep "tma prof" gpurun_out/r2q_bench_$d.err | head -2
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2q_bench_$d.json").read().strip().splitlines()[-1])
    k=d["roofline"]["kernels"]
    print("dbg $d", round(d["value"]), k["k1_unproject_fwd"]["ms"])
except Exception as e: print("$d failed", e)
PY
done
