# final evidence of the round: full GPU suite, smoke, default bench (both arms), launch list + full-set ncu of the step,
# K7 bench + ncu
set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_cfg3.json 2> gpurun_out/bench_cfg3.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>&1
bash tools/gpu_k7.sh > gpurun_out/k7_final.log 2>&1; grep "_us" gpurun_out/k7_final.log | head -12
bash tools/gpu_k7_ncu.sh > gpurun_out/k7_ncu.log 2>&1; tail -2 gpurun_out/k7_ncu.log
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu1.log 2>&1
echo "ncu1 rc=$?"
