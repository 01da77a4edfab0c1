mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_models.py -q -m gpu --timeout 600 -k "inception or branch" 2>&1 | tail -4
for m in "inception bsp" "inception ssp"; do set -- $m
timeout 600 python bench.py --model $1 --mode $2 --steps 10 --warmup 3 --no-e2e > gpurun_out/inc_v12_$2.json 2> gpurun_out/inc_v12_$2.err; python -c "
import json; d=json.load(open('gpurun_out/inc_v12_$2.json')); print('$1 $2', round(d['value'],1), round(d['ms_per_step'],3), d['final_loss'])" || tail -5 gpurun_out/inc_v12_$2.err
done
timeout 600 python bench.py --model inception --steps 10 --warmup 3 > gpurun_out/inc_v12_e2e.json 2> gpurun_out/inc_v12_e2e.err; python -c "
import json; d=json.load(open('gpurun_out/inc_v12_e2e.json')); print('e2e', d['e2e'])"
