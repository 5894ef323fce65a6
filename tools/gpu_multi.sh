# N-GPU smoke of the bench contract: cfg3 (batch-sharded, weak) and cfg5 (z-slabs, one all_gather)
N=${1:-2}
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for wl in cfg3 cfg5; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload $wl --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${wl}_x$N.json 2> gpurun_out/bench_${wl}_x$N.err || tail -5 gpurun_out/bench_${wl}_x$N.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_${wl}_x$N.json').read().strip().splitlines()[-1])
print('$wl x$N value',round(d['value'],2),'e2e',d['e2e'] and round(d['e2e']['value'],2),'ms/step',round(d['ms_per_step'],4), 'n_gpus', d['n_gpus'])
PY
done
python bench.py --workload cfg5 --steps 20 --warmup 3 --no-cpu-baseline --no-e2e | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('cfg5 x1', round(d['value'],2), {k:v['ms'] for k,v in d['roofline']['kernels'].items()})"
