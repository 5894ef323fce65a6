# round-2 evidence pass (one B200): parity suite, smoke, default bench (both arms), the other BASELINE configs, DLT,
# ncu launch list + one full-set capture of every kernel of the step (each ncu run only after the same command exited 0)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
python -m pytest tests -m gpu -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2_pytest_gpu.log; tail -3 gpurun_out/r2_pytest_gpu.log
python -m pytest tests/test_gpu_reference_nets.py tests/test_gpu_torch_ops.py -m gpu -q -s 2>&1 | grep -E "parameter gradients|algebraic net|public call|stock net|passed|failed" > gpurun_out/r2_nets_and_binding.log; cat gpurun_out/r2_nets_and_binding.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2_smoke.log
python bench.py > gpurun_out/r2_bench_cfg3.json 2> gpurun_out/r2_bench_cfg3.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err
python bench.py --no-graph --no-cpu-baseline --steps 100 > gpurun_out/r2_bench_cfg3_eager.json 2> /dev/null
for w in cfg1 cfg4 cfg5; do
  python bench.py --workload $w --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench_$w.json 2> gpurun_out/r2_bench_$w.err
done
python bench.py --workload cfg1 --steps 100 --warmup 5 --no-cpu-baseline --no-graph > gpurun_out/r2_bench_cfg1_eager.json 2> /dev/null
python tools/bench_dlt.py > gpurun_out/r2_bench_dlt.json 2> gpurun_out/r2_bench_dlt.err
python - <<PY
import json
for n in ("cfg3","cfg3_eager","cfg1","cfg1_eager","cfg4","cfg5"):
    try:
        d=json.loads(open("gpurun_out/r2_bench_%s.json"%n).read().strip().splitlines()[-1])
        k=d["roofline"]["kernels"]
        print(n, round(d["value"],1), "e2e", round(d["e2e"]["value"],1), d["e2e"]["cuda_graph"], {a:k[a]["ms"] for a in k}, "step frac", d["roofline"]["whole_step"]["frac"])
    except Exception as e: print(n, "failed", e)
PY
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'lt::|k_' -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu1.log 2>&1
echo "ncu1 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_unproject|k_sam_stats|k_sam_final|k_sam_bwd|k_feat' -s 7 -c 7 -o gpurun_out/r2_prof_round python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu2.log 2>&1
echo "ncu2 rc=$?"
