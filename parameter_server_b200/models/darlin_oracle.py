"""Plain PyTorch (fp64, CPU or GPU) restatement of the reference's Darlin loops.

Used as the numerical oracle of the CUDA kernels and as the documented semantics of the
algorithm.  It follows the reference literally — it keeps dual_i = exp(y_i <x_i, w>) and
multiplies it, sends nothing clever — so agreement with the device path (which keeps the log
of dual and pre-multiplied values) is a real check.

ref: src/app/linear_method/darlin.h:206-246 (updateWeight), :381-427 (computeGradient),
     :429-502 (updateDual), :248-265 / :504-511 (evaluate).
"""
from __future__ import annotations

import math

import torch


class DarlinOracle:
    def __init__(self, col_ptr, row_idx, val, y, n_rows, lambda_l1=1.0, eta=1.0,
                 delta_init=1.0, delta_max=5.0):
        """CSC of the (plain, not label-folded) matrix of ONE worker; several oracles can
        share a server through `OracleServer`."""
        self.col_ptr = col_ptr.long()
        self.row = row_idx.long()
        self.val = None if val is None else val.double()
        self.y = y.double()
        n_cols = col_ptr.numel() - 1
        dev = col_ptr.device
        self.col_of = torch.repeat_interleave(torch.arange(n_cols, device=dev),
                                              self.col_ptr[1:] - self.col_ptr[:-1])
        self.dual = torch.ones(n_rows, dtype=torch.float64, device=dev)
        self.w = torch.zeros(n_cols, dtype=torch.float64, device=dev)
        self.delta = torch.full((n_cols,), delta_init, dtype=torch.float64, device=dev)
        self.active = torch.ones(n_cols, dtype=torch.bool, device=dev)
        self.delta_max = delta_max

    def gradient(self, lo: int, hi: int):
        """G, U of columns [lo,hi); NaN for inactive columns (darlin.h:400-406)."""
        a, b = int(self.col_ptr[lo]), int(self.col_ptr[hi])
        row, col = self.row[a:b], self.col_of[a:b] - lo
        tau = 1.0 / (1.0 + self.dual[row])
        y = self.y[row]
        n = hi - lo
        G = torch.zeros(n, dtype=torch.float64, device=row.device)
        U = torch.zeros(n, dtype=torch.float64, device=row.device)
        d = self.delta[lo:hi][col]
        if self.val is None:
            G.index_add_(0, col, -y * tau)
            U.index_add_(0, col, torch.minimum(tau * (1 - tau) * torch.exp(d), torch.tensor(.25, dtype=torch.float64, device=row.device)))
        else:
            v = self.val[a:b]
            G.index_add_(0, col, -y * tau * v)
            U.index_add_(0, col, torch.minimum(tau * (1 - tau) * torch.exp(v.abs() * d),
                                               torch.tensor(.25, dtype=torch.float64, device=row.device)) * v * v)
        inactive = ~self.active[lo:hi]
        G[inactive] = float("nan")
        U[inactive] = float("nan")
        return G, U

    def update_dual(self, lo: int, hi: int, new_w: torch.Tensor) -> None:
        cur = self.w[lo:hi]
        marked = torch.isnan(new_w)
        dw = torch.where(marked, torch.zeros_like(new_w), new_w - cur)
        self.active[lo:hi] &= ~marked
        self.delta[lo:hi] = torch.where(marked, self.delta[lo:hi],
                                        torch.clamp(2 * dw.abs() + .1, max=self.delta_max))
        self.w[lo:hi] = torch.where(marked, torch.zeros_like(cur), new_w)
        a, b = int(self.col_ptr[lo]), int(self.col_ptr[hi])
        row, col = self.row[a:b], self.col_of[a:b] - lo
        wd = torch.where(self.active[lo:hi], dw, torch.zeros_like(dw))[col]
        x = 1.0 if self.val is None else self.val[a:b]
        # dual_i *= exp(y_i wd x_ij), several columns may hit one row
        expo = torch.zeros_like(self.dual).index_add_(0, row, self.y[row] * wd * x)
        self.dual *= torch.exp(expo)

    def loss(self) -> float:
        return float(torch.log(1 + 1 / self.dual).sum())


class OracleServer:
    def __init__(self, n_keys, device, lambda_l1=1.0, eta=1.0, delta_init=1.0, delta_max=5.0):
        f = dict(dtype=torch.float64, device=device)
        self.w = torch.zeros(n_keys, **f)
        self.delta = torch.full((n_keys,), delta_init, **f)
        self.active = torch.ones(n_keys, dtype=torch.bool, device=device)
        self.lambda_l1, self.eta, self.delta_max = lambda_l1, eta, delta_max
        self.violation = 0.0
        self.kkt_threshold = 1e20

    def begin_pass(self, kkt_threshold: float, reset: bool) -> None:
        self.kkt_threshold = kkt_threshold
        self.violation = 0.0
        if reset:
            self.active[:] = True

    def update_weight(self, lo: int, hi: int, G: torch.Tensor, U: torch.Tensor) -> None:
        """Literal per-coordinate loop semantics, vectorised (darlin.h:206-246)."""
        lam, thr = self.lambda_l1, self.kkt_threshold
        w = self.w[lo:hi].clone()
        act = self.active[lo:hi].clone()
        delta = self.delta[lo:hi].clone()
        g = G
        u = U / self.eta + 1e-10
        gp, gn = g + lam, g - lam
        is0 = w == 0
        vio = torch.zeros_like(w)
        vio = torch.where(is0 & (gp < 0), -gp, vio)
        vio = torch.where(is0 & ~(gp < 0) & (gn > 0), gn, vio)
        filt = act & is0 & ~(gp < 0) & ~(gn > 0) & (gp > thr) & (gn < -thr)
        upd = act & ~filt
        if bool(upd.any()):
            self.violation = max(self.violation, float(vio[upd].max()))
        d = -w
        d = torch.where(gp <= u * w, -gp / u, d)
        d = torch.where(~(gp <= u * w) & (gn >= u * w), -gn / u, d)
        # NaN weights revived by a reset: std::max/min on NaN collapse to d = -delta
        nanw = torch.isnan(w)
        d = torch.where(nanw, -delta, d)
        d = torch.minimum(delta, torch.maximum(-delta, d))
        new_delta = torch.clamp(2 * d.abs() + .1, max=self.delta_max)
        self.delta[lo:hi] = torch.where(upd, new_delta, delta)
        w_new = torch.where(upd, w + d, w)
        w_new = torch.where(filt, torch.full_like(w, float("nan")), w_new)
        self.w[lo:hi] = w_new
        self.active[lo:hi] = act & ~filt

    def evaluate(self):
        ok = ~torch.isnan(self.w) & (self.w != 0)
        return dict(penalty=float(self.w[ok].abs().sum()) * self.lambda_l1,
                    nnz_w=int(ok.sum()), nnz_active=int(self.active.sum()),
                    violation=self.violation)


def run_oracle_pass(workers: list, server: OracleServer, blocks: list, order: list,
                    kkt_threshold: float, reset: bool, blocks_per_step: int = 1):
    """One pass with the reference's semantics; `blocks` = [(lo,hi)] in global columns (all
    workers share the global column space).  Returns merged progress dict."""
    server.begin_pass(kkt_threshold, reset)
    if reset:
        for wk in workers:
            wk.active[:] = True
    for s in range(0, len(order), blocks_per_step):
        step = order[s:s + blocks_per_step]
        grads = []
        for b in step:
            lo, hi = blocks[b]
            G = U = None
            for wk in workers:
                g, u = wk.gradient(lo, hi)
                g = torch.nan_to_num(g, nan=0.0)  # server skips inactive coords anyway
                u = torch.nan_to_num(u, nan=0.0)
                G = g if G is None else G + g
                U = u if U is None else U + u
            grads.append((G, U))
        for b, (G, U) in zip(step, grads):
            lo, hi = blocks[b]
            server.update_weight(lo, hi, G, U)
        for b in step:
            lo, hi = blocks[b]
            for wk in workers:
                wk.update_dual(lo, hi, server.w[lo:hi].clone())
    ev = server.evaluate()
    loss = sum(wk.loss() for wk in workers)
    ev["objective"] = loss + ev["penalty"]
    ev["loss"] = loss
    return ev
