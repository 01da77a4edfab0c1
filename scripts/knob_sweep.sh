#!/bin/bash
# A/B runs of the push / gate knobs on N GPUs:  bash scripts/knob_sweep.sh N model mode "ENV=.. ENV=.." ...
N=$1; MODEL=$2; MODE=$3; shift 3
export DTF_PS_TIMEOUT_S=${DTF_PS_TIMEOUT_S:-60}
mkdir -p gpurun_out
for knobs in "$@"; do
  out=$(env $knobs timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 \
    bench.py --gpus $N --steps 20 --warmup 5 --model $MODEL --mode $MODE --no-secondary --no-e2e 2> gpurun_out/knob.err | tail -1)
  echo "$MODEL $MODE N=$N [$knobs] $(echo "$out" | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.readline()); print(round(d['value'],1),'img/s',round(d['ms_per_step'],4),'ms',d['config'].get('push'))
except Exception as e: print('ERR',e)
")" | tee -a gpurun_out/knob_sweep.log
done
