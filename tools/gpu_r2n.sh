#!/bin/bash
# round 2n: TMA forward v4 (16 consumer warps, 2 channels per lane): parity, timing (+ dbg 1/2), ncu
mkdir -p gpurun_out
export LT_B200_K1FWD=tma
timeout 600 python -m pytest tests/test_gpu_unproject.py tests/test_gpu_coordspec.py tests/test_gpu_fuzz_shapes.py tests/test_gpu_guard_bands.py tests/test_gpu_pipeline.py -x -q -m gpu 2>&1 | tail -5 | tee gpurun_out/r2n_pytest.log
for d in 0 1 2; do
export LT_B200_TMA_DBG=$d
timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2n_bench_$d.json 2> gpurun_out/r2n_bench_$d.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2n_bench_$d.json").read().strip().splitlines()[-1])
    k=d["roofline"]["kernels"]
    print("dbg $d", round(d["value"]), k["k1_unproject_fwd"]["ms"])
except Exception as e: print("$d failed", e)
PY
done
export LT_B200_TMA_DBG=0
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_unproject_fwd_tma -c 1 -o gpurun_out/r2n_k1fwd_tma python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/r2n_ncu.log 2>&1
