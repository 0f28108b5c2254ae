#!/bin/bash
# One rank of an mpirun / srun / torchrun launched CPU cluster:
#   script/mpi_node.sh "<scheduler node text>" <launch.conf>
# The rank comes from the launcher's environment (MPICH PMI_RANK, Open MPI
# OMPI_COMM_WORLD_RANK, Slurm SLURM_PROCID, torchrun RANK).  Rank 0 becomes the scheduler,
# ranks 1..num_workers the workers, the next num_servers the servers, the rest hot spares; every
# rank finds its own address and port (Postoffice::start with -my_rank, runtime.cc).
# Same role as the reference's script/mpi_node.sh.
if [ $# -ne 2 ]; then
  echo "usage: $0 scheduler_node launch.conf"
  exit 1
fi
scheduler=$1
source "$2"
for v in PMI_RANK OMPI_COMM_WORLD_RANK SLURM_PROCID RANK; do
  if [ -n "${!v}" ]; then my_rank=${!v}; break; fi
done
for v in PMI_SIZE OMPI_COMM_WORLD_SIZE SLURM_NTASKS WORLD_SIZE; do
  if [ -n "${!v}" ]; then size=${!v}; break; fi
done
if [ -z "${my_rank}" ] || [ -z "${size}" ]; then
  echo "$0: cannot find my rank / the number of ranks in the environment"
  exit 1
fi
need=$((num_workers + num_servers + 1))
if [ "${size}" -lt "${need}" ]; then
  echo "$0: ${size} ranks, but 1 scheduler + ${num_servers} servers + ${num_workers} workers need ${need}"
  exit 1
fi
iface=()
if [ -n "${network_interface}" ]; then iface=(-interface "${network_interface}"); fi
exec ${bin} -num_servers ${num_servers} -num_workers ${num_workers} \
  -num_unused $((size - need)) -scheduler "${scheduler}" -my_rank ${my_rank} "${iface[@]}" ${arg}
