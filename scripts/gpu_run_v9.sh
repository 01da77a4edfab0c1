mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_zoo_ops.py tests/test_gpu_ops.py -q -m gpu --timeout 300 -x 2>&1 | tail -8
timeout 600 python scripts/gpu_check_zoo_ops.py > gpurun_out/zoo_ops_perf2.jsonl 2> gpurun_out/zoo_ops_perf2.err; tail -2 gpurun_out/zoo_ops_perf2.err
timeout 600 python bench.py --model inception --steps 10 --warmup 3 --no-e2e > gpurun_out/inc_v9.json 2> gpurun_out/inc_v9.err; python -c "
import json; d=json.load(open('gpurun_out/inc_v9.json')); print('inception', d['value'], d['ms_per_step'])"
timeout 600 python bench.py --steps 50 --warmup 10 > gpurun_out/bench_v9.json 2> gpurun_out/bench_v9.err; python -c "
import json; d=json.load(open('gpurun_out/bench_v9.json')); print('resnet', d['value'], d['ms_per_step'], d['e2e'])"
