#!/bin/bash
# Usage (same contract as the reference's train.sh):
#   export ps='host1 host2' workers='host3 host4 host5'
#   bash train.sh <model> <mode> [extra trainer flags]      model: cnn cifarnet alexnet vgg resnet inception
#                                                            mode : bsp asp ssp
# Logs: ps_<i>.log, worker_<j>.log, model_eval.log in $DTF_LOG_DIR (default: current directory).
exp_home="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
cd "$exp_home" || exit 1
if [ -z "$1" ] || [ -z "$2" ]; then
  echo "please specify <model> <mode>, e.g.: bash train.sh resnet bsp"
  exit 1
fi
exec python -m distributed_tensorflow_models_b200.launcher "$@"
