# A/B differently built libraries (learnable-triangulation-pytorch_b200/variants/*.so, see csrc/Makefile EXTRA=) on cfg3
run() {
  LT_B200_LIB=$1 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/variant.json 2> gpurun_out/variant.err || tail -3 gpurun_out/variant.err
  python - "$1" <<'PY'
import json, sys
d = json.load(open('gpurun_out/variant.json'))
k = d['roofline']['kernels']
print(sys.argv[1].split('/')[-1] or 'default', 'value', round(d['value'], 1), ' '.join('%s=%.4f' % (n.split('_', 1)[1], v['ms']) for n, v in k.items()))
PY
}
run ""
for f in learnable-triangulation-pytorch_b200/variants/*.so; do run $PWD/$f; done
