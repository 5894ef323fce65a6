# timing-only A/B of K7 forward variants (variants/*.so built with EXTRA=-D...; results of experiment builds may be wrong)
for f in "" learnable-triangulation-pytorch_b200/variants/*.so; do
  LT_B200_LIB=${f:+$PWD/$f} timeout 200 python tools/bench_process_features.py > gpurun_out/v.json 2> gpurun_out/v.err || tail -3 gpurun_out/v.err
  python - "$f" <<'PY'
import json, sys
d = json.load(open('gpurun_out/v.json'))
print(sys.argv[1].split('/')[-1] or 'default', 'fwd_us', round(d['k7_fwd_us'], 1), 'bwd_data', round(d['k7_bwd_data_us'], 1), 'bwd_weight', round(d['k7_bwd_weight_us'], 1))
PY
done
