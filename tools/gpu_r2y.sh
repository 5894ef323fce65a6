#!/bin/bash
# round 2y: A/B per aggregation at cfg3's shape: library variants (opaque row pointers per aggregation)
mkdir -p gpurun_out
for agg in sum conf max softmax; do
for lib in "" $PWD/learnable-triangulation-pytorch_b200/variants/*.so; do
LT_B200_LIB=$lib python bench.py --agg $agg --steps 30 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/ab.json 2> gpurun_out/ab.err || tail -3 gpurun_out/ab.err
python - "$agg" "$lib" <<'PY'
import json, sys
d = json.loads(open('gpurun_out/ab.json').read().strip().splitlines()[-1])
k = d['roofline']['kernels']
print(sys.argv[1].ljust(8), (sys.argv[2].split('/')[-1] or 'new').ljust(18), 'value', round(d['value'], 1), ' '.join('%s=%.4f' % (n.split('_', 1)[1], v['ms']) for n, v in k.items()))
PY
done; done
