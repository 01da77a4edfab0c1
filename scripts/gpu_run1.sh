set -x
python -m pytest tests/test_gpu_ops.py -x -q -m gpu --timeout 300 2>&1 | tail -30
python -m pytest tests/test_gpu_models.py -q -m gpu --timeout 600 2>&1 | tail -40
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
timeout 300 python bench.py --steps 20 --warmup 5 2>&1 | tail -5
