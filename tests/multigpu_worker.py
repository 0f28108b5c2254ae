"""torchrun worker: train a small Darlin problem on WORLD_SIZE GPUs with the fused kernel and
compare the objective trajectory with the emulated lock-step run done by rank 0 alone."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--passes", type=int, default=4)
    args = ap.parse_args()
    from parameter_server_b200.models.darlin import DarlinConfig, DarlinEngine, run_pass_lockstep
    from parameter_server_b200.parallel.comm import TorchComm, run_threads
    from parameter_server_b200.utils.synthetic import make_rcv1_like
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    dist.init_process_group("nccl", device_id=dev)
    cfg = DarlinConfig(lambda_l1=0.5, feature_block_ratio=0.5, seg_len=64, max_block_delay=1, seed=3)
    shard = make_rcv1_like(4000, 900, 24.0, seed=3, shard=rank, device=dev)
    eng = DarlinEngine(cfg, TorchComm(), dev)
    eng.load(shard)
    rng = np.random.RandomState(1)
    thr, reset, objs = 1e20, False, []
    n_ex = eng.num_ex_global
    orders = []
    for it in range(args.passes):
        order = list(range(len(eng.blocks)))
        rng.shuffle(order)
        orders.append(order)
        p = eng.run_pass(order, thr, reset)
        objs.append((p.objective, p.nnz_w, p.nnz_active_set, p.violation))
        thr = p.violation / n_ex * 0.5
    ok = True
    detail = {}
    if rank == 0 and args.check:
        shards = [make_rcv1_like(4000, 900, 24.0, seed=3, shard=r, device=dev) for r in range(world)]
        cfg2 = DarlinConfig(lambda_l1=0.5, feature_block_ratio=0.5, seg_len=64, max_block_delay=1,
                            seed=3, use_graph=False, fused=False)

        def build(comm):
            e = DarlinEngine(cfg2, comm, dev)
            e.load(shards[comm.rank])
            return e
        engines = run_threads(world, build)
        thr, ref = 1e20, []
        for it in range(args.passes):
            p = run_pass_lockstep(engines, orders[it], thr, False)
            ref.append((p.objective, p.nnz_w, p.nnz_active_set, p.violation))
            thr = p.violation / n_ex * 0.5
        for a, b in zip(objs, ref):
            # the persistent kernel evaluates exp / reciprocal with its own <=1.5 ulp routines, the
            # split kernels with the library's: trajectories agree to rounding noise, not bitwise
            ok &= abs(a[0] - b[0]) <= 1e-7 * abs(b[0]) and a[1] == b[1] and a[2] == b[2]
        detail = {"multi": objs, "emulated": ref}
    dist.barrier()
    if rank == 0:
        print(json.dumps({"ok": bool(ok), **detail}))
    eng.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
