#!/bin/bash
# Run an app config on the GPUs of this machine, one process per GPU (server shard + worker):
#   script/gpu.sh <#gpus> -app_file example/linear/rcv1/batch_l1lr.conf [more flags]
if [ $# -lt 2 ]; then
  echo "usage: $0 num_gpus [flags of parameter_server_b200.apps.ps]"
  exit 1
fi
n=$1; shift
if [ "$n" -le 1 ]; then exec python -m parameter_server_b200.apps.ps "$@"; fi
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" \
  --master-addr 127.0.0.1 --master-port ${PS_BASE_PORT:-29400} \
  -m parameter_server_b200.apps.ps "$@"
