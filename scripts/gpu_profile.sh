set -x
mkdir -p gpurun_out
# baselines (PyTorch ops: cuDNN/cuBLAS, TF32 allowed) on the same engine, eager and graph
timeout 300 python bench.py --steps 30 --warmup 5 --kernels torch --no-graph > gpurun_out/bench_torch_eager.json 2> gpurun_out/bench_torch_eager.err
timeout 300 python bench.py --steps 30 --warmup 5 --kernels torch > gpurun_out/bench_torch_graph.json 2> gpurun_out/bench_torch_graph.err
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err
timeout 300 python bench.py --steps 30 --warmup 5 --no-fuse-bn > gpurun_out/bench_native_nofuse.json 2> gpurun_out/bench_native_nofuse.err
tail -n 2 gpurun_out/bench_*.json gpurun_out/bench_*.err | cut -c1-400
# launch list + top-kernel profile (plain run first, same command line)
python bench.py --steps 2 --warmup 3 --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 600 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-e2e > gpurun_out/ncu1.log 2>&1
python bench.py --steps 2 --warmup 3 --no-e2e > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv_gemm_kernel -s 100 -c 3 -o gpurun_out/prof_conv_gemm \
    python bench.py --steps 2 --warmup 3 --no-e2e > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out | tail -12
