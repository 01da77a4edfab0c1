N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -q -m gpu --timeout 900 2>&1 | tail -3
run() { # name, extra args
  name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N "$@" > gpurun_out/bench_n${N}_${name}.json 2> gpurun_out/bench_n${N}_${name}.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_n${N}_${name}.json')); print('$name', round(d['value'],1), round(d['ms_per_step'],3), 'e2e', d['e2e'] and round(d['e2e']['value'],1), 'loss', d['final_loss'])" || tail -5 gpurun_out/bench_n${N}_${name}.err
}
run resnet_bsp --steps 30 --warmup 5
run inception_ssp --model inception --mode ssp --staleness 3 --steps 10 --warmup 3
run vgg_asp --model vgg --mode asp --steps 10 --warmup 3
