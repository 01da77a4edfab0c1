# Convenience targets; everything here is a one-liner you can also type yourself (README.md "Use").
PY ?= python

.PHONY: build test gpu-test bench bench8 bench-input sass sass-check clean

build:            ## nvcc -gencode arch=compute_100a,code=sm_100a for every csrc/*.cu, in-tree .so
	$(PY) -c "import __graft_entry__ as g; g.build()"

test: build       ## CPU suite: protocols (gloo / thread cluster), models, slim, data, checkpoints, CLI
	$(PY) -m pytest tests -q -m "not gpu"

gpu-test: build   ## B200: every kernel and model against plain PyTorch fp32
	$(PY) -m pytest tests -q -m gpu

bench: build      ## headline: ResNet-32 b128 BSP on one GPU, one JSON line
	$(PY) bench.py --gpus 1 --steps 50 --warmup 10

bench8: build     ## weak scaling on 8 GPUs
	$(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29500 \
	    bench.py --gpus 8 --steps 50 --warmup 10

bench-input: build  ## host input pipelines (no GPU): CIFAR-10 and ImageNet, native passes vs torch / NumPy
	$(PY) scripts/bench_input_pipeline.py

sass: build       ## regenerate profiles/sass/*.sass + MNEMONICS.txt from the current sources
	python scripts/regen_sass.py

sass-check:       ## fail when a committed SASS listing is older than its source
	python scripts/regen_sass.py --check

clean:
	rm -rf distributed_tensorflow_models_b200/_native/build distributed_tensorflow_models_b200/_native/libdtfb200.so
