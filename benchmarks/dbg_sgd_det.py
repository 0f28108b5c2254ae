"""Is the lock-step emulated async-SGD run reproducible run to run? (float atomics in the gradient)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from parameter_server_b200.models.async_sgd import AsyncSGDConfig, AsyncSGDEngine
from parameter_server_b200.parallel.comm import run_threads
from parameter_server_b200.utils.synthetic import make_rcv1_like
dev = torch.device("cuda:0")
steps, world = 10, 2
cfg = AsyncSGDConfig(algo="FTRL", minibatch=256, alpha=0.5, beta=1.0, lambda_l1=0.02, lambda_l2=0.01,
                     max_delay=0, table_capacity=1 << 14)
def shard(r):
    s = make_rcv1_like(steps * 256, 3000, 24.0, seed=4, shard=r, device=dev)
    s.keys = (s.keys * ((1 << 63) // 3000 * 2)).contiguous()
    return s
def run():
    shards = [shard(r) for r in range(world)]
    def build(comm):
        e = AsyncSGDEngine(cfg, comm, dev, out=open(os.devnull, "w")); e.load(shards[comm.rank]); return e
    engines = run_threads(world, build)
    for i in range(steps):
        for e in engines: e.worker_step(i)
        for e in engines: e.server_step()
    torch.cuda.synchronize()
    m = {}
    for e in engines:
        k, w = e.shard_model(); m.update(dict(zip(k.tolist(), w.tolist()))); e.close()
    return m
a, b = run(), run()
keys = set(a) | set(b)
d = sorted(((abs(a.get(k, 0) - b.get(k, 0)), a.get(k, 0), b.get(k, 0)) for k in keys), reverse=True)
print("nnz", len(a), len(b), "same keys", set(a) == set(b), "top diffs", d[:5])
