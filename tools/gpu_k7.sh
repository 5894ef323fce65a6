# K7 (process_features) parity + measurement, tensor-core forward (default) and FFMA forward (LT_B200_PWCONV=ffma)
timeout 300 python -m pytest tests/test_gpu_process_features.py -x -q > gpurun_out/pytest_k7.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_k7.log
LT_B200_PWCONV=ffma timeout 300 python -m pytest tests/test_gpu_process_features.py -x -q > gpurun_out/pytest_k7_ffma.log 2>&1; echo "pytest(ffma) rc=$?"; tail -3 gpurun_out/pytest_k7_ffma.log
timeout 300 python tools/bench_process_features.py > gpurun_out/bench_k7.json 2> gpurun_out/bench_k7.err; echo "bench rc=$?"
LT_B200_PWCONV=ffma timeout 300 python tools/bench_process_features.py > gpurun_out/bench_k7_ffma.json 2> gpurun_out/bench_k7_ffma.err; echo "bench(ffma) rc=$?"
python - <<'PY'
import json
for f in ('gpurun_out/bench_k7.json','gpurun_out/bench_k7_ffma.json'):
    try:
        d=json.load(open(f))
    except Exception as e:
        print(f, 'unreadable', e); continue
    print(f)
    for k,v in d.items():
        if k.startswith('k7') or k.startswith('stock'): print('  ',k, round(v,3) if isinstance(v,float) else v)
PY
tail -3 gpurun_out/bench_k7.err
