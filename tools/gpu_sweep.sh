run() {
  echo "== $*"
  env "$@" python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/sweep.json 2> gpurun_out/sweep.err || tail -3 gpurun_out/sweep.err
  python - <<'PY'
import json
d=json.load(open('gpurun_out/sweep.json'))
k=d['roofline']['kernels']
print('value',round(d['value'],1),' '.join('%s=%.4f'%(n.split('_',1)[1],v['ms']) for n,v in k.items()))
PY
}
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run LT_DUMMY=1
python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/bench_quick.json 2>gpurun_out/bench_quick.err; python -c "
import json; d=json.load(open('gpurun_out/bench_quick.json')); print('value',d['value'],'e2e',d['e2e'])"
