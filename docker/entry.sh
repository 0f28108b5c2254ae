#!/bin/bash
# Container entry point.  PS_ROLE selects what this container is:
#   scheduler | server | worker   one CPU node of build/ps (PS_ID, PS_PORT, PS_SCHEDULER, PS_NUM_SERVERS,
#                                 PS_NUM_WORKERS and the app flags in "$@")
#   gpu                           all GPUs of this container: script/gpu.sh $PS_NUM_GPUS "$@"
set -e
cd /opt/psb200
case "${PS_ROLE:-gpu}" in
  gpu)
    exec script/gpu.sh "${PS_NUM_GPUS:-1}" "$@" ;;
  scheduler|server|worker)
    role=$(echo "${PS_ROLE}" | tr a-z A-Z)
    ip=$(hostname -i | awk '{print $1}')
    me="role:${role},hostname:'${ip}',port:${PS_PORT:-8000},id:'${PS_ID:-H}'"
    if [ "${PS_ROLE}" = scheduler ]; then sched=${me}; else sched=${PS_SCHEDULER}; fi
    exec build/ps -my_node "${me}" -scheduler "${sched}" -num_servers "${PS_NUM_SERVERS}" \
      -num_workers "${PS_NUM_WORKERS}" "$@" ;;
  *)
    echo "unknown PS_ROLE=${PS_ROLE}"; exit 1 ;;
esac
