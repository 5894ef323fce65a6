#!/bin/bash
# round 2, first GPU pass: parity of everything (incl. the tile backward + new cfg4/cfg5/reference-net tests), then A/B
# of the K1 backward formulations at cfg3, then an ncu launch list.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
for v in scatter 0 1 2 3; do
  if [ "$v" = scatter ]; then export LT_B200_K1BWD=scatter; else export LT_B200_K1BWD=tile LT_B200_K1BWD_TILE=$v; fi
  python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2a_bench_$v.json 2> gpurun_out/r2a_bench_$v.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2a_bench_$v.json"))
    k=d["roofline"]["kernels"]
    print("$v", round(d["value"]), {n:k[n]["ms"] for n in k})
except Exception as e: print("$v failed", e)
PY
done
unset LT_B200_K1BWD LT_B200_K1BWD_TILE
ncu --metrics gpu__time_duration.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2a_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/r2a_ncu.log 2>&1
echo done
