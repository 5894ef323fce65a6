#!/bin/bash
# round 2ae: K0 with pointer walking instead of re-derived 64-bit addresses: parity (everything goes through K0) + timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_unproject.py tests/test_gpu_guard_bands.py tests/test_gpu_fuzz_shapes.py tests/test_gpu_process_features.py tests/test_gpu_pipeline.py -x -q -m gpu 2>&1 | tail -2
python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/r2ae_bench.json 2> gpurun_out/r2ae_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2ae_bench.json").read().strip().splitlines()[-1])
k=d["roofline"]["kernels"]
print(round(d["value"]), d["ms_per_step"], {x:(k[x]["ms"], k[x]["frac"]) for x in k if x.startswith("k0")})
PY
