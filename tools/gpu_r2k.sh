#!/bin/bash
# round 2k: where does the TMA forward kernel wait? dbg 1 = no copies, dbg 2 = no gather (timing only)
mkdir -p gpurun_out
export LT_B200_K1FWD=tma
for d in 0 1 2; do
export LT_B200_TMA_DBG=$d
timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2k_bench_$d.json 2> gpurun_out/r2k_bench_$d.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2k_bench_$d.json").read().strip().splitlines()[-1])
    k=d["roofline"]["kernels"]
    print("dbg $d", round(d["value"]), k["k1_unproject_fwd"]["ms"])
except Exception as e: print("$d failed", e)
PY
done
