#!/bin/bash
mkdir -p gpurun_out
python tools/diag/nets_diag.py > gpurun_out/r2b_nets_diag.log 2>&1; tail -6 gpurun_out/r2b_nets_diag.log
python -m pytest tests -m gpu -q --deselect tests/test_gpu_reference_nets.py > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
tail -15 gpurun_out/r2b_pytest.log
for v in scatter 0 1 2 3; do
  if [ "$v" = scatter ]; then export LT_B200_K1BWD=scatter; else export LT_B200_K1BWD=tile LT_B200_K1BWD_TILE=$v; fi
  python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2b_bench_$v.json 2> gpurun_out/r2b_bench_$v.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2b_bench_$v.json"))
    k=d["roofline"]["kernels"]
    print("$v", round(d["value"]), {n:k[n]["ms"] for n in k})
except Exception as e: print("$v failed", e)
PY
done
for v in 0 1; do
export LT_B200_K1BWD=tile LT_B200_K1BWD_TILE=$v
ncu --set full --import-source on --clock-control none -k regex:k_unproject_bwd_tile -c 1 -o gpurun_out/r2b_tile$v python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/r2b_ncu$v.log 2>&1
done
ls -la gpurun_out/*.ncu-rep
