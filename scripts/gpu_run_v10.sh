mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_models.py -q -m gpu --timeout 600 -x -k "inception" 2>&1 | tail -12
for flag in "" "--no-branch-streams"; do
  timeout 600 python bench.py --model inception --steps 10 --warmup 3 --no-e2e $flag > gpurun_out/inc_v10$flag.json 2> gpurun_out/inc_v10$flag.err; python -c "
import json; d=json.load(open('gpurun_out/inc_v10$flag.json')); print('inception $flag', round(d['value'],1), round(d['ms_per_step'],3), d['final_loss'])" || tail -5 gpurun_out/inc_v10$flag.err
done
timeout 600 python bench.py --model inception --steps 10 --warmup 3 --no-e2e --kernels torch > gpurun_out/inc_v10_torch.json 2> gpurun_out/inc_v10_torch.err; python -c "
import json; d=json.load(open('gpurun_out/inc_v10_torch.json')); print('inception torch', round(d['value'],1), round(d['ms_per_step'],3), d['final_loss'])" || tail -5 gpurun_out/inc_v10_torch.err
