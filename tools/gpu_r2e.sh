#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2e_pytest.log
tail -12 gpurun_out/r2e_pytest.log
grep -n "parameter gradients\|algebraic net\|public call" gpurun_out/r2e_pytest.log | head
for v in fused split; do
  export LT_B200_K2FWD=$v
  python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2e_bench_$v.json 2> gpurun_out/r2e_bench_$v.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2e_bench_$v.json"))
    k=d["roofline"]["kernels"]
    print("$v", round(d["value"]), {n:k[n]["ms"] for n in k})
except Exception as e: print("$v failed", e)
PY
done
unset LT_B200_K2FWD
for w in cfg3 cfg1; do
python bench.py --steps 100 --warmup 5 --workload $w --no-cpu-baseline > gpurun_out/r2e_bench_$w.json 2> gpurun_out/r2e_bench_$w.err; python -c "
import json; d=json.load(open('gpurun_out/r2e_bench_$w.json')); print('$w', round(d['value']), 'e2e graph', round(d['e2e']['value']), d['e2e']['cuda_graph'])"
python bench.py --steps 100 --warmup 5 --workload $w --no-cpu-baseline --no-graph > gpurun_out/r2e_bench_${w}_eager.json 2> gpurun_out/r2e_bench_${w}_eager.err; python -c "
import json; d=json.load(open('gpurun_out/r2e_bench_${w}_eager.json')); print('$w', round(d['value']), 'e2e eager', round(d['e2e']['value']), d['e2e']['cuda_graph'])"
done
ncu --set full --import-source on --clock-control none -k regex:k_unproject_fwd -c 1 -o gpurun_out/r2e_k1fwd python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/r2e_ncu.log 2>&1
