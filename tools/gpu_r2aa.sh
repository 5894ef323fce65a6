#!/bin/bash
# round 2aa: K2 forward, one read per volume with thread-block clusters (LT_B200_K2FWD=cluster): timing incl. spec path
mkdir -p gpurun_out
for m in cluster split; do
if [ $m = split ]; then unset LT_B200_K2FWD; else export LT_B200_K2FWD=cluster; fi
timeout 400 python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2aa_bench_$m.json 2> gpurun_out/r2aa_bench_$m.err
python - $m <<'PY'
import json, sys
m=sys.argv[1]
d=json.loads(open("gpurun_out/r2aa_bench_%s.json"%m).read().strip().splitlines()[-1])
k=d["roofline"]["kernels"]; s=d["roofline"]["next_rows"]["in_kernel_coordinates"]["kernels"]
print(m, round(d["value"]), "e2e", round(d["e2e"]["value"]), "K2 fwd tensor %.4f ms" % k["k2_softargmax_fwd"]["ms"], "spec %.4f ms" % s["k2_softargmax_fwd"]["ms"])
PY
done
export LT_B200_K2FWD=cluster
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_sam_fwd_cluster -c 2 -o gpurun_out/r2aa_k2cluster python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/r2aa_ncu.log 2>&1
