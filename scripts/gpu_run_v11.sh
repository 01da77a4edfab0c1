mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu --timeout 600 2>&1 | tail -6 > gpurun_out/pytest_gpu_v11.log; tail -3 gpurun_out/pytest_gpu_v11.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --steps 50 --warmup 10 > gpurun_out/bench_v11.json 2> gpurun_out/bench_v11.err; python -c "
import json; d=json.load(open('gpurun_out/bench_v11.json')); print('resnet', d['value'], d['ms_per_step'], d['e2e'])"
bash scripts/gpu_sweep1.sh 2>&1 | tail -20
