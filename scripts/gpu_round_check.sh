# One-call GPU check used at the end of a work session: full GPU test-suite, smoke, headline bench, zoo-op timings,
# and a per-kernel launch list of one Inception-v3 step (ncu, duration metric only).  Everything lands in gpurun_out/.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu --timeout 600 -x 2>&1 | tail -15 > gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py --steps 50 --warmup 10 > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err; tail -c 900 gpurun_out/bench_native.json
timeout 600 python scripts/gpu_check_zoo_ops.py > gpurun_out/zoo_ops_perf.jsonl 2> gpurun_out/zoo_ops_perf.err; tail -3 gpurun_out/zoo_ops_perf.jsonl; tail -3 gpurun_out/zoo_ops_perf.err
timeout 600 python bench.py --model inception --steps 3 --warmup 3 --no-graph --no-e2e > gpurun_out/inc_nograph.json 2> gpurun_out/inc_nograph.err
timeout 800 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_inception.csv python bench.py --model inception --steps 1 --warmup 3 --no-graph --no-e2e > gpurun_out/ncu_inc.log 2>&1
tail -2 gpurun_out/ncu_inc.log; wc -l gpurun_out/launches_inception.csv
