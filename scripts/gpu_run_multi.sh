set -x
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/multi_gpu_check.py 2>&1 | tail -25
for be in symm nccl_rs_ag; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 30 --warmup 5 --backend $be > gpurun_out/bench_n${N}_${be}.json 2> gpurun_out/bench_n${N}_${be}.err
  tail -c 1500 gpurun_out/bench_n${N}_${be}.json; tail -5 gpurun_out/bench_n${N}_${be}.err
done
