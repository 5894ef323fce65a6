#!/bin/bash
# round 2 multi-GPU evidence: NCCL slab parity on real ranks, the bench contract at N ranks (cfg3 weak with the sharded
# cfg5 side record, cfg5 strong), the H2D-only ceiling of the e2e arm
N=${1:-2}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_slab.py -m gpu -q > gpurun_out/r2m_slab_x$N.log 2>&1; echo "slab pytest rc=$?" | tee -a gpurun_out/r2m_slab_x$N.log; tail -3 gpurun_out/r2m_slab_x$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29511 tools/microbench/h2d_ranks.py > gpurun_out/r2m_h2d_x$N.json 2> gpurun_out/r2m_h2d_x$N.err; cat gpurun_out/r2m_h2d_x$N.json
$TR --master-port 29512 bench.py --gpus $N --steps 100 --warmup 5 > gpurun_out/r2m_bench_cfg3_x$N.json 2> gpurun_out/r2m_bench_cfg3_x$N.err
$TR --master-port 29513 bench.py --gpus $N --steps 30 --warmup 3 --workload cfg5 > gpurun_out/r2m_bench_cfg5_x$N.json 2> gpurun_out/r2m_bench_cfg5_x$N.err
python - <<PY
import json
for w in ("cfg3","cfg5"):
    try:
        d=json.loads(open("gpurun_out/r2m_bench_%s_x$N.json"%w).read().strip().splitlines()[-1])
        print(w, "value", round(d["value"]), "e2e", round(d["e2e"]["value"]) if d.get("e2e") else None, d["config"]["parallelism"][:60])
        print("   sharded:", json.dumps(d["roofline"].get("next_rows",{}).get("cfg5_sharded_softargmax"))[:600])
    except Exception as e: print(w, "failed", e)
PY
