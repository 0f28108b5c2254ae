"""NCCL-collective variant of the Darlin step: the comparison baseline BASELINE.json asks for.

Same split kernels, same arithmetic, same scheduler; only the transport differs.  Where the
peer-memory engine stores gradients straight into the owners' landing slots and reads weights
back from peer memory, this variant keeps every kernel on local buffers and moves the data with
NCCL between the launches:

    grad kernel -> local staging -> all_to_all (block range of every owner)   = push  (C1)
    server_update kernel (fixed worker order reduce, as always)
    all_to_all of the owners' updated (w, dw) ranges -> local replicas        = pull  (C2)
    pull kernel on the replicas

ref: the reference sends the same slices through ZeroMQ (sliceKeyOrderedMsg, message.h:120-159;
KVBufferedVector::setValue, kv_buffered_vector.h:41-79; KVVector::getValue, kv_vector.h:78-93).
Host driven (one collective per direction and step, no graph): it measures what a step costs when
the collective library sits between the kernels.
"""
from __future__ import annotations

import torch

from ..ops import darlin as K


def _setup(e) -> None:
    W, cap, dev = e.world, e.land_cap, e.device
    f64 = dict(dtype=torch.float64, device=dev)
    e._n_stage_g = [torch.zeros(cap, **f64) for _ in range(W)]
    e._n_stage_u = [torch.zeros(cap, **f64) for _ in range(W)]
    e._n_rep_w = [torch.zeros(cap, **f64) for _ in range(W)]
    e._n_rep_dw = [torch.zeros(cap, **f64) for _ in range(W)]
    e._n_flags = torch.zeros(4 * W, dtype=torch.int64, device=dev)  # push | ver | sink | sink
    wk = type(e.wk).from_buffer_copy(bytes(e.wk))
    sv = type(e.sv).from_buffer_copy(bytes(e.sv))
    sink = e._n_flags.data_ptr() + 2 * W * 8
    for s in range(W):
        wk.land_g[s], wk.land_u[s] = e._n_stage_g[s].data_ptr(), e._n_stage_u[s].data_ptr()
        wk.land_cap[s] = 0                      # slot index = column offset in the owner's shard
        wk.push_flag[s] = sink                  # the kernel's release flags go nowhere
        wk.srv_w[s], wk.srv_dw[s] = e._n_rep_w[s].data_ptr(), e._n_rep_dw[s].data_ptr()
        sv.ver_flag[s] = sink + 8 * W
    wk.ver_flag = e._n_flags.data_ptr() + W * 8
    sv.push_flag = e._n_flags.data_ptr()
    e._n_wk, e._n_sv = wk, sv
    e.b_land_g.tensor.zero_(); e.b_land_u.tensor.zero_()


def _exchange(e, send: list, recv: list) -> None:
    if e.world == 1:
        recv[0].copy_(send[0])
        return
    import torch.distributed as dist
    dist.all_to_all(recv, send)


def enqueue_pass_nccl(e, n_steps: int) -> None:
    if not hasattr(e, "_n_wk"):
        _setup(e)
    W, cap, r = e.world, e.land_cap, e.rank
    wk, sv = e._n_wk, e._n_sv
    cp, op = e.ctrl_dev.data_ptr(), e.order_dev.data_ptr()
    order = e.order_host[:n_steps * e.bps].view(n_steps, e.bps).tolist()
    land_g = e.b_land_g.tensor.view(W, cap)
    land_u = e.b_land_u.tensor.view(W, cap)
    w_loc, dw_loc = e.b_w.tensor, e.b_dw.tensor
    K.pass_begin(wk, sv, cp)
    seq = e._seq
    launches = 1
    for i in range(n_steps):
        seq += 1
        K.grad_push(wk, cp, op, i, e.bps, e.grid_worker)
        for b in order[i]:
            if b < 0:
                continue
            rng = e.blk_srv_rng[b]
            lo_me, hi_me = int(rng[r, 0]), int(rng[r, 1])
            for stage, land in ((e._n_stage_g, land_g), (e._n_stage_u, land_u)):
                send = [stage[s][int(rng[s, 0]):int(rng[s, 1])] for s in range(W)]
                recv = [land[q, lo_me:hi_me] for q in range(W)]
                _exchange(e, send, recv)
                launches += 1
        e._n_flags[:W].fill_(seq)
        K.server_update(sv, cp, op, i, e.bps, e.grid_server)
        for b in order[i]:
            if b < 0:
                continue
            rng = e.blk_srv_rng[b]
            lo_me, hi_me = int(rng[r, 0]), int(rng[r, 1])
            for src, rep in ((w_loc, e._n_rep_w), (dw_loc, e._n_rep_dw)):
                send = [src[lo_me:hi_me] for _ in range(W)]
                recv = [rep[s][int(rng[s, 0]):int(rng[s, 1])] for s in range(W)]
                _exchange(e, send, recv)
                launches += 1
        e._n_flags[W:2 * W].fill_(seq)
        K.pull_update(wk, cp, op, i, e.bps, e.grid_worker)
        launches += 5
    K.evaluate(wk, sv, cp, e.arena.peer_ptr(r, e.b_progress, 6 * r), e.prog_peer.data_ptr(), W, 0,
               e.eval_scratch.data_ptr(), e.grid_eval)
    e._nccl_launches = launches + 1
    if W > 1:
        import torch.distributed as dist
        dist.barrier()
