# First GPU call of the next session: the tests of the kernels that were written without a GPU, then their timings
# and one ncu capture each (duration + memory throughput).  Everything lands in gpurun_out/.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_zy_gpu_new_kernels.py -q -m gpu --timeout 600 2>&1 | tail -25 > gpurun_out/pytest_new_kernels.log; tail -5 gpurun_out/pytest_new_kernels.log
timeout 900 python scripts/gpu_check_new_kernels.py > gpurun_out/new_kernels_perf.jsonl 2> gpurun_out/new_kernels_perf.err; cat gpurun_out/new_kernels_perf.jsonl; tail -3 gpurun_out/new_kernels_perf.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"cifar_distort|fake_quant|minmax" -c 12 \
  -o gpurun_out/new_kernels python -m pytest tests/test_zy_gpu_new_kernels.py -q -m gpu -k "fake_quant or device_kernel or range_update" > gpurun_out/ncu_new_kernels.log 2>&1
tail -2 gpurun_out/ncu_new_kernels.log; ls -la gpurun_out/new_kernels.ncu-rep
