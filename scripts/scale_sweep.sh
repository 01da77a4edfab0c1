#!/bin/bash
# Scaling sweep on N GPUs of one box: headline + secondary configs, the comm-bound models, the push alone.
#   bash scripts/scale_sweep.sh N [tag]      -> gpurun_out/scale_<tag>_n<N>_*.json
N=${1:-2}; TAG=${2:-r2}; OUT=gpurun_out; mkdir -p $OUT
export DTF_PS_TIMEOUT_S=${DTF_PS_TIMEOUT_S:-60}
run() {  # name, extra bench args...
  local name=$1; shift
  if [ "$N" = "1" ]; then
    timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 "$@" > $OUT/scale_${TAG}_n${N}_${name}.json 2> $OUT/scale_${TAG}_n${N}_${name}.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 \
      bench.py --gpus $N --steps 20 --warmup 5 "$@" > $OUT/scale_${TAG}_n${N}_${name}.json 2> $OUT/scale_${TAG}_n${N}_${name}.err
  fi
  echo "$name rc=$? $(python - <<PY
import json
try:
    d=json.loads(open("$OUT/scale_${TAG}_n${N}_${name}.json").readline())
    print(round(d["value"],1),"img/s",round(d["ms_per_step"],3),"ms", "e2e",round(d["e2e"]["value"],1) if d.get("e2e") else None, [ (s.get("metric","")[12:40], round(s.get("value",0),1), round(s.get("ms_per_step",0),3)) for s in d.get("secondary",[])])
except Exception as e: print("ERR",e)
PY
)"
}
run headline
run vgg_bsp --model vgg --mode bsp --no-secondary
run vgg_asp --model vgg --mode asp --no-secondary
run alexnet_bsp --model alexnet --mode bsp --no-secondary
run cifarnet_bsp --model cifarnet --mode bsp --no-secondary
if [ "$N" != "1" ]; then
  DTF_PUSH_OVERLAP=0 run vgg_bsp_nooverlap --model vgg --mode bsp --no-secondary
  DTF_PUSH_OVERLAP=0 run inception_bsp_nooverlap --model inception --mode bsp --no-secondary
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 \
    scripts/time_push.py vgg alexnet inception resnet > $OUT/scale_${TAG}_n${N}_time_push.json 2> $OUT/scale_${TAG}_n${N}_time_push.err
  grep '^{' $OUT/scale_${TAG}_n${N}_time_push.json
fi
