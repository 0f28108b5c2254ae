#!/bin/bash
# stop and remove the containers started by docker/local.sh
ids=$(docker ps -aq --filter "name=ps-")
[ -n "${ids}" ] && docker rm -f ${ids}
docker network rm psb200-net >/dev/null 2>&1 || true
