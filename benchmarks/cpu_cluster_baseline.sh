#!/bin/bash
# CPU arm of BASELINE.md B1 / B3: this repository's own host runtime (scheduler + 2 servers +
# 2 workers over TCP on localhost) on synthetic rcv1-shaped text data.  Not the reference binary:
# that cannot be built in this image.
#   benchmarks/cpu_cluster_baseline.sh [rows=80000] [workdir=/tmp/psb]
set -e
rows=${1:-80000}
work=${2:-/tmp/psb}
root=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p "${work}" && cd "${work}"
[ -d data/rcv1/train ] || python "${root}/example/linear/gen_data.py" rcv1 --rows "${rows}" --parts 8 --out data
mkdir -p model data/cache
for conf in batch_l1lr online_l1lr; do
  echo "== ${conf}"
  /usr/bin/time -f "wall %e s" env PS_BASE_PORT=$((23000 + RANDOM % 2000)) "${root}/script/local.sh" \
    "${root}/build/ps" 2 2 -app_file "${root}/example/linear/rcv1/${conf}.conf" -num_threads 2 2>&1 | grep -v '^\['
done
