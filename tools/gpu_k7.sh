# K7 (process_features) parity + measurement
python -m pytest tests/test_gpu_process_features.py -x -q > gpurun_out/pytest_k7.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_k7.log
timeout 300 python tools/bench_process_features.py > gpurun_out/bench_k7.json 2> gpurun_out/bench_k7.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_k7.json'))
for k,v in d.items(): print(k, round(v,3) if isinstance(v,float) else v)
PY
tail -3 gpurun_out/bench_k7.err
