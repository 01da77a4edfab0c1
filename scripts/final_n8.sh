export DTF_PS_TIMEOUT_S=60
R() { name=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 20 --warmup 5 "$@" > gpurun_out/final_n8_$name.json 2> gpurun_out/final_n8_$name.err; echo "$name rc=$? $(python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/final_n8_$name.json").readline())
    print(round(d["value"],1),"img/s",round(d["ms_per_step"],4),"ms e2e",round(d["e2e"]["value"],1) if d.get("e2e") else None,[(s.get("metric","")[12:45],round(s.get("value",0),1),round(s.get("ms_per_step",0),3)) for s in d.get("secondary",[])])
except Exception as e: print("ERR",e)
PY
)"; }
R headline
R vgg_bsp --model vgg --mode bsp --no-secondary
R vgg_asp --model vgg --mode asp --no-secondary
R alexnet_bsp --model alexnet --mode bsp --no-secondary
R nccl_vgg_bsp --model vgg --mode bsp --no-secondary --no-e2e --backend nccl_rs_ag
R nccl_vgg_asp --model vgg --mode asp --no-secondary --no-e2e --backend nccl_rs_ag
R nccl_inception_bsp --model inception --mode bsp --no-secondary --no-e2e --backend nccl_rs_ag
R nccl_resnet_bsp --model resnet --mode bsp --no-secondary --no-e2e --backend nccl_rs_ag
