#!/bin/bash
# round 2ac: tile backward with the address-arithmetic fix: parity + timing per tile shape
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_unproject_tile.py -x -q -m gpu 2>&1 | tail -2
for v in 0 1 2 3; do
export LT_B200_K1BWD=tile LT_B200_K1BWD_TILE=$v
python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2ac_bench_$v.json 2> gpurun_out/r2ac_bench_$v.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2ac_bench_$v.json").read().strip().splitlines()[-1])
print("tile $v", round(d["value"]), d["roofline"]["kernels"]["k1_unproject_bwd"]["ms"])
PY
done
