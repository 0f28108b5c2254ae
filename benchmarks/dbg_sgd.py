import sys, torch, faulthandler
sys.path.insert(0, "/root/repo")
faulthandler.dump_traceback_later(40, exit=True)
from parameter_server_b200.models.async_sgd import AsyncSGDConfig, AsyncSGDEngine
from parameter_server_b200.parallel.comm import run_threads
from parameter_server_b200.utils.synthetic import make_rcv1_like
dev = torch.device("cuda:0")
world = int(sys.argv[1]) if len(sys.argv) > 1 else 1
cfg = AsyncSGDConfig(algo="FTRL", minibatch=256, alpha=0.5, beta=1.0, lambda_l1=0.02, lambda_l2=0.01, table_capacity=1 << 14)
shards = [make_rcv1_like(12 * 256, 3000, 24.0, seed=4, shard=r, device=dev) for r in range(world)]
for s in shards:
    s.keys = (s.keys * ((1 << 63) // 3000 * 2 // 1)).contiguous()
def build(comm):
    e = AsyncSGDEngine(cfg, comm, dev); e.load(shards[comm.rank]); return e
engines = run_threads(world, build)
print("built", flush=True)
for step in range(12):
    for e in engines:
        e.worker_step(step)
    if len(sys.argv) > 2: torch.cuda.synchronize(); print("step", step, "workers done", [e.b_lflag.tensor.tolist() for e in engines], flush=True)
    for e in engines:
        e.server_step()
    torch.cuda.synchronize(); print("step", step, "servers done", [e.b_ack.tensor.tolist() for e in engines], flush=True)
