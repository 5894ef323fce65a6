#!/bin/bash
# round 2i: K1 forward, persistent TMA-staged kernel (LT_B200_K1FWD=tma): parity on every K1 test, then timing
mkdir -p gpurun_out
export LT_B200_K1FWD=tma
timeout 600 python -m pytest tests/test_gpu_unproject.py tests/test_gpu_coordspec.py tests/test_gpu_fuzz_shapes.py tests/test_gpu_guard_bands.py tests/test_gpu_unproject_tile.py tests/test_gpu_pipeline.py -x -q -m gpu 2>&1 | tail -15 | tee gpurun_out/r2i_pytest.log
timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2i_bench_tma.json 2> gpurun_out/r2i_bench_tma.err
unset LT_B200_K1FWD
timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2i_bench_base.json 2> gpurun_out/r2i_bench_base.err
python - <<'PY'
import json
for n in ("tma","base"):
    try:
        d=json.loads(open("gpurun_out/r2i_bench_%s.json"%n).read().strip().splitlines()[-1])
        k=d["roofline"]["kernels"]
        print(n, round(d["value"]), {x:k[x]["ms"] for x in k})
    except Exception as e: print(n,"failed",e)
PY
tail -3 gpurun_out/r2i_bench_tma.err
