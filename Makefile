# Convenience targets; everything here is a one-liner you can also type yourself (README.md "Use").
PY ?= python

.PHONY: build test gpu-test bench bench8 bench8-all multi-check litmus convergence trace bench-input sass sass-check clean
NGPU ?= 2

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

bench8-all: build ## the round's 8-GPU table: headline + secondary, VGG-16 BSP / ASP, AlexNet, and the NCCL arms
	bash scripts/final_n8.sh

multi-check: build  ## NGPU GPUs: bucketed push + gated pull vs the NCCL arm, ASP, SSP bound (one JSON line, "ok")
	$(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node $(NGPU) --master-addr 127.0.0.1 --master-port 29511 \
	    scripts/multi_gpu_check.py

litmus: build     ## NGPU GPUs: no reader that saw an `updated` flag reads a value of the previous epoch
	$(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node $(NGPU) --master-addr 127.0.0.1 --master-port 29533 \
	    scripts/litmus_push_order.py 300 alexnet

convergence: build  ## one GPU: loss curves native TF32 / PyTorch fp32 / native 3xTF32 (gpurun_out/convergence_*.json)
	$(PY) scripts/convergence.py resnet 2000
	$(PY) scripts/convergence.py inception 300

trace: build      ## kernel timeline of one step (diagnosis only): make trace MODEL=vgg MODE=bsp
	$(PY) scripts/trace_step.py $(or $(MODEL),resnet) $(or $(MODE),bsp)

bench-input: build  ## host input pipelines (no GPU): CIFAR-10 and ImageNet, native passes vs torch / NumPy
	$(PY) scripts/bench_input_pipeline.py

sass: build       ## regenerate profiles/sass/*.sass + MNEMONICS.txt from the current sources
	python scripts/regen_sass.py

sass-check:       ## fail when a committed SASS listing is older than its source
	python scripts/regen_sass.py --check

clean:
	rm -rf distributed_tensorflow_models_b200/_native/build distributed_tensorflow_models_b200/_native/libdtfb200.so
