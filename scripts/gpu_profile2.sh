set -x
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 700 -c 500 --csv --log-file gpurun_out/launches2.csv \
    python bench.py --steps 2 --warmup 3 --no-e2e > gpurun_out/ncu1.log 2>&1
python bench.py --steps 2 --warmup 3 --no-e2e > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv_wgrad_kernel -s 60 -c 2 -o gpurun_out/prof_wgrad \
    python bench.py --steps 2 --warmup 3 --no-e2e > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out | tail -5
