# one line per BASELINE config + the DLT line; each stored under gpurun_out/
for wl in cfg1 cfg4 cfg5; do
  python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err || tail -3 gpurun_out/bench_$wl.err
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_$wl.json'))
print('$wl value',round(d['value'],2),'e2e',d['e2e'] and round(d['e2e']['value'],2),'ms/step',round(d['ms_per_step'],4), {k:v['ms'] for k,v in d['roofline']['kernels'].items()})
PY
done
python tools/bench_dlt.py > gpurun_out/bench_dlt.json 2> gpurun_out/bench_dlt.err; cat gpurun_out/bench_dlt.json; tail -2 gpurun_out/bench_dlt.err
