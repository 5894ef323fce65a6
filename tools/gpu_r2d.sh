#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2d_pytest.log
tail -12 gpurun_out/r2d_pytest.log
grep -n "parameter gradients\|algebraic net" gpurun_out/r2d_pytest.log | head
for v in fused split; do
  export LT_B200_K2FWD=$v
  python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2d_bench_$v.json 2> gpurun_out/r2d_bench_$v.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2d_bench_$v.json"))
    k=d["roofline"]["kernels"]
    print("$v", round(d["value"]), {n:k[n]["ms"] for n in k})
except Exception as e: print("$v failed", e)
PY
done
unset LT_B200_K2FWD
python bench.py --steps 100 --warmup 5 > gpurun_out/r2d_bench_full.json 2> gpurun_out/r2d_bench_full.err; python -c "
import json; d=json.load(open('gpurun_out/r2d_bench_full.json')); print('full', round(d['value']), 'e2e', round(d['e2e']['value']), d['e2e']['cuda_graph'], 'cpu', d['cpu_baseline'])"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2d_bench_ref.json 2> gpurun_out/r2d_bench_ref.err; cat gpurun_out/r2d_bench_ref.json | cut -c1-400
