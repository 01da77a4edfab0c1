mkdir -p gpurun_out
for cfg in "inception bsp" "inception ssp" "vgg asp" "vgg bsp" "alexnet bsp" "cifarnet bsp" "cnn bsp" "resnet ssp"; do
  set -- $cfg
  for k in native torch; do
    timeout 600 python bench.py --model $1 --mode $2 --steps 10 --warmup 3 --kernels $k --no-e2e > gpurun_out/sweep_$1_$2_$k.json 2> gpurun_out/sweep_$1_$2_$k.err
    echo "$1 $2 $k rc=$? $(python -c "
import json,sys
try:
    d=json.load(open('gpurun_out/sweep_$1_$2_$k.json')); print(round(d['value']), 'img/s', round(d['ms_per_step'],2), 'ms loss', d['final_loss'])
except Exception as e: print('FAIL', open('gpurun_out/sweep_$1_$2_$k.err').read()[-600:])
")"
  done
done
