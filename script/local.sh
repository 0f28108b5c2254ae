#!/bin/bash
# Launch one scheduler, S servers and W workers of a node binary on this machine.
#   script/local.sh <binary> <#servers> <#workers> [flags forwarded to every node]
# Same calling convention as the reference's script/local.sh; PS_BASE_PORT moves the ports.
if [ $# -lt 3 ]; then
  echo "usage: $0 bin num_servers num_workers [args..]"
  exit 1
fi
bin=$1; shift
num_servers=$1; shift
num_workers=$1; shift
arg="-num_servers ${num_servers} -num_workers ${num_workers} $@"
base=${PS_BASE_PORT:-8000}
Sch="role:SCHEDULER,hostname:'127.0.0.1',port:${base},id:'H'"
pids=()
${bin} -my_node ${Sch} -scheduler ${Sch} ${arg} &
sched_pid=$!
for ((i = 0; i < num_servers; ++i)); do
  port=$((base + 1600 + i))
  N="role:SERVER,hostname:'127.0.0.1',port:${port},id:'S${i}'"
  ${bin} -my_node ${N} -scheduler ${Sch} ${arg} &
  pids+=($!)
done
for ((i = 0; i < num_workers; ++i)); do
  port=$((base + 1500 + i))
  N="role:WORKER,hostname:'127.0.0.1',port:${port},id:'W${i}'"
  ${bin} -my_node ${N} -scheduler ${Sch} ${arg} &
  pids+=($!)
done
# Wait for every node.  A node that dies (a port already taken, a failed CHECK) would leave the
# scheduler waiting for its registration forever: on the first non-zero exit stop the others.
all=(${sched_pid} "${pids[@]}")
rc=0
left=${#all[@]}
while [ ${left} -gt 0 ]; do
  wait -n
  st=$?
  left=$((left - 1))
  if [ ${st} -ne 0 ]; then
    rc=${st}
    kill "${all[@]}" 2>/dev/null
    sleep 1
    kill -9 "${all[@]}" 2>/dev/null   # whoever ignored the first signal
    wait
    break
  fi
done
exit $rc
