#!/bin/bash
# Stop everything train.sh started (recorded PIDs / process groups only), on worker AND ps hosts.
exp_home="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
cd "$exp_home" || exit 1
exec python -m distributed_tensorflow_models_b200.launcher --clear "$@"
