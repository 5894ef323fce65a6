#!/bin/bash
# round 2v: descriptors carry the view offset (32-bit, relative to the item) and the texel pitch is an immediate:
# parity of everything that shares build_descs / gather4 / scatter4, then the bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_unproject.py tests/test_gpu_unproject_tile.py tests/test_gpu_unproject_tma.py tests/test_gpu_coordspec.py tests/test_gpu_fuzz_shapes.py tests/test_gpu_guard_bands.py tests/test_gpu_pipeline.py tests/test_gpu_torch_ops.py -x -q -m gpu 2>&1 | tail -3
timeout 300 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/r2v_bench.json 2> gpurun_out/r2v_bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2v_bench.json").read().strip().splitlines()[-1])
k=d["roofline"]["kernels"]
print(round(d["value"]), d["ms_per_step"], {x:k[x]["ms"] for x in k})
PY
