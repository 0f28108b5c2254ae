#!/bin/bash
# One machine, one container per node (same calling convention as the reference's docker/local.sh):
#   docker/local.sh num_servers num_workers app_conf_file data_dir model_dir [args...]
# Containers share a user-defined bridge network; remove them with docker/rm_local.sh.
if [ $# -lt 5 ]; then
  echo "usage: $0 num_servers num_workers app_conf_file data_dir model_dir [args...]"
  exit 1
fi
ns=$1; nw=$2; app=$(realpath "$3"); data=$(realpath "$4"); model=$(realpath "$5"); shift 5
image=${PS_IMAGE:-psb200}
net=psb200-net
docker network inspect ${net} >/dev/null 2>&1 || docker network create ${net} >/dev/null
mounts="-v ${app}:/app.conf:ro -v ${data}:/data:ro -v ${model}:/model"
common="-e PS_NUM_SERVERS=${ns} -e PS_NUM_WORKERS=${nw} --network ${net} ${mounts}"
docker run -d --name ps-H ${common} -e PS_ROLE=scheduler -e PS_ID=H -e PS_PORT=8000 ${image} -app_file /app.conf "$@"
sched_ip=$(docker inspect -f '{{range .NetworkSettings.Networks}}{{.IPAddress}}{{end}}' ps-H)
sched="role:SCHEDULER,hostname:'${sched_ip}',port:8000,id:'H'"
for ((i = 0; i < ns; ++i)); do
  docker run -d --name ps-S${i} ${common} -e PS_ROLE=server -e PS_ID=S${i} -e PS_PORT=$((8100 + i)) \
    -e PS_SCHEDULER="${sched}" ${image} -app_file /app.conf "$@"
done
for ((i = 0; i < nw; ++i)); do
  docker run -d --name ps-W${i} ${common} -e PS_ROLE=worker -e PS_ID=W${i} -e PS_PORT=$((8200 + i)) \
    -e PS_SCHEDULER="${sched}" ${image} -app_file /app.conf "$@"
done
docker logs -f ps-H
