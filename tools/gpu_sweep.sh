# env sweeps of the K1 tile kernels: prints per-kernel ms for each setting
run() {
  echo "== $*"
  env "$@" python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/sweep.json 2> gpurun_out/sweep.err || tail -3 gpurun_out/sweep.err
  python - <<'PY'
import json
d=json.load(open('gpurun_out/sweep.json'))
k=d['roofline']['kernels']
print('value',round(d['value'],1),' fwd',k['k1_unproject_fwd']['ms'],' bwd',k['k1_unproject_bwd']['ms'])
PY
}
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run LT_K1_TILE_ZT=4 LT_K1_TILE_OCC=2
run LT_K1_TILE_ZT=4 LT_K1_TILE_OCC=3
run LT_K1_TILE_ZT=8 LT_K1_TILE_OCC=2
run LT_K1_TILE_ZT=8 LT_K1_TILE_OCC=1
