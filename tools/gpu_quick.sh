# quick GPU check: parity suite + one bench line (no CPU baseline)
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 20 --warmup 3 --no-cpu-baseline ${BENCH_ARGS:-} > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_quick.json'))
print('value',round(d['value'],1),'e2e',d['e2e'] and round(d['e2e']['value'],1),'ms/step',round(d['ms_per_step'],4))
for k,v in d['roofline']['kernels'].items(): print(k,v['ms'],v['frac'])
PY
tail -3 gpurun_out/bench_quick.err
