# round-1d evidence for K7 on top of tools/gpu_round1.sh: full GPU suite, K7 bench (tensor-core and FFMA paths), K7 ncu,
# forward load-policy variants, one bench.py line with the K7 side measurement
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
bash tools/gpu_k7.sh > gpurun_out/k7_final.log 2>&1; grep "_us" gpurun_out/k7_final.log | head -12
bash tools/gpu_k7_ncu.sh > gpurun_out/k7_ncu.log 2>&1; tail -3 gpurun_out/k7_ncu.log
if ls learnable-triangulation-pytorch_b200/variants/*.so > /dev/null 2>&1; then bash tools/gpu_k7_variants.sh > gpurun_out/variants.log 2>&1; grep fwd_us gpurun_out/variants.log; fi
python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_quick.json'))
print('value',round(d['value'],1),'e2e',d['e2e'] and round(d['e2e']['value'],1)); print(d['roofline'].get('next_rows'))
PY
