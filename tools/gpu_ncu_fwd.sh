python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_unproject_fwd' -s 1 -c 1 -o gpurun_out/prof_fwd python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_fwd.log 2>&1
echo "ncu rc=$?"
