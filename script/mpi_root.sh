#!/bin/bash
# Start a CPU cluster through MPI:  script/mpi_root.sh <launch.conf>
# launch.conf is a shell fragment (see script/mpi.conf.example) naming the binary, the number
# of servers and workers, the interface / port of the scheduler, an optional hostfile and the
# arguments forwarded to every node.  Run it on the first machine of the hostfile: that is
# where rank 0, the scheduler, lands.  Same role as the reference's script/mpi_root.sh.
if [ $# -ne 1 ]; then
  echo "usage: $0 launch.conf"
  exit 1
fi
conf=$1
source "${conf}"
dir=$(dirname "$0")
root=$("${dir}/get_root_node.sh" "${network_interface:--}" "${network_port:-8000}")
np=$((num_workers + num_servers + 1 + ${num_unused:-0}))
launcher=${launcher:-mpirun}
if ! command -v "${launcher}" >/dev/null 2>&1; then
  echo "$0: ${launcher} not found; use script/local.sh on one machine"
  exit 1
fi
hosts=()
if [ -n "${hostfile}" ]; then hosts=(-hostfile "${hostfile}"); fi
exec "${launcher}" "${hosts[@]}" -np ${np} "${dir}/mpi_node.sh" "${root}" "${conf}"
