mkdir -p gpurun_out
timeout 600 python bench.py --steps 50 --warmup 10 > gpurun_out/bench_v13.json 2> gpurun_out/bench_v13.err; python -c "
import json; d=json.load(open('gpurun_out/bench_v13.json')); print('resnet', round(d['value']), d['ms_per_step'], d['e2e'])" || tail -5 gpurun_out/bench_v13.err
timeout 600 python bench.py --model inception --steps 10 --warmup 3 > gpurun_out/inc_v13.json 2> gpurun_out/inc_v13.err; python -c "
import json; d=json.load(open('gpurun_out/inc_v13.json')); print('inception', round(d['value']), d['ms_per_step'], d['e2e'])" || tail -5 gpurun_out/inc_v13.err
rm -rf /tmp/t13; timeout 300 python -m distributed_tensorflow_models_b200.trainers.cifar10_resnet_bsp --synthetic --max_steps 300 --log_every 100 --train_dir /tmp/t13 2>&1 | tail -4
