# Reproduces profiles/README.md section 0.5 on a box with 2 GPUs (results under gpurun_out/).
run() { tag=$1; shift; env "$@" timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 benchmarks/peer_table_probe.py > gpurun_out/probe_$tag.log 2>&1; echo "== $tag rc=$?"; grep -E '"peer, (one|all)' gpurun_out/probe_$tag.log; }
run granule_4k PSB_ARENA_GRANULE=4096
run granule_2m PSB_ARENA_GRANULE=2097152
echo "== standalone, 8 GB + 4 KB allocation"
PEER_ODD=1 timeout 60 build/micro/peer_random > gpurun_out/peer_random_odd.txt 2>&1; echo rc=$?; grep -E "^peer +round|^peer +2048 MB +200000 +1 +8" gpurun_out/peer_random_odd.txt
