# Model-zoo networks through bench.py on one GPU: native kernels vs the same engine on PyTorch ops (both with the
# whole step in a CUDA graph).  Not a BASELINE.json config — breadth numbers for the zoo kernels.
#   bash scripts/gpu_sweep_zoo.sh        -> gpurun_out/zoo_<net>_<kernels>.json + one summary line each
mkdir -p gpurun_out
for cfg in "mobilenet_v1 224 64" "mobilenet_v2 224 64" "resnet_v1_50 224 32" "inception_v1 224 32" "nasnet_mobile 224 32" "inception_v4 299 16"; do
  set -- $cfg
  for k in native torch; do
    timeout 600 python bench.py --model $1 --image-size $2 --batch $3 --steps 10 --warmup 3 --kernels $k --no-e2e \
      > gpurun_out/zoo_$1_$k.json 2> gpurun_out/zoo_$1_$k.err
    echo "$1 b$3 $k rc=$? $(python -c "
import json
try:
    d=json.load(open('gpurun_out/zoo_$1_$k.json')); print(round(d['value']), 'img/s', round(d['ms_per_step'],2), 'ms loss', d['final_loss'])
except Exception as e: print('FAIL', open('gpurun_out/zoo_$1_$k.err').read()[-400:])
")"
  done
done
