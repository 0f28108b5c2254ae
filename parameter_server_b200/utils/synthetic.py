"""Seeded synthetic sparse datasets of the shapes named in BASELINE.json.

There is no network, hence no rcv1 / CTR download; the generators reproduce the *shape*:

rcv1-shaped  47 236 features, 677 399 examples (the reference trains on the big split,
             example/linear/rcv1/download.sh:26-36), ~73 non-zeros per row, skewed feature
             popularity with permuted ids, positive tf-idf-like values with unit row norm,
             labels from a planted sparse model plus noise.
ctr-shaped   hashed 64-bit keys in several feature groups, binary features.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

RCV1_FEATURES = 47236
RCV1_EXAMPLES = 677399
RCV1_NNZ_PER_ROW = 73.2


@dataclass
class SparseCSR:
    """Row-major sparse examples; keys are uint64 bit patterns stored in int64."""
    row_ptr: torch.Tensor          # int64 [n_rows+1]
    keys: torch.Tensor             # int64 [nnz]
    vals: torch.Tensor | None      # float32 [nnz]; None => binary
    y: torch.Tensor                # float32 [n_rows] in {-1,+1}
    n_features: int                # key space upper bound (max_key+1)

    @property
    def n_rows(self) -> int:
        return self.row_ptr.numel() - 1

    @property
    def nnz(self) -> int:
        return self.keys.numel()

    def to(self, device) -> "SparseCSR":
        return SparseCSR(self.row_ptr.to(device), self.keys.to(device),
                         None if self.vals is None else self.vals.to(device),
                         self.y.to(device), self.n_features)


def _planted_model(n_features: int, seed: int, density: float = 0.02, device="cpu"):
    g = torch.Generator(device="cpu").manual_seed(seed * 7919 + 13)
    w = torch.randn(n_features, generator=g, dtype=torch.float64)
    mask = torch.rand(n_features, generator=g) < density
    return (w * mask).to(device)


def make_rcv1_like(n_rows: int = RCV1_EXAMPLES, n_features: int = RCV1_FEATURES,
                   nnz_per_row: float = RCV1_NNZ_PER_ROW, seed: int = 0, shard: int = 0,
                   device="cpu", binary: bool = False, noise: float = 0.5,
                   model_density: float | None = None) -> SparseCSR:
    """One worker's shard of an rcv1-shaped dataset.

    `seed` fixes the feature popularity permutation and the planted model (shared by all
    shards); `shard` varies the examples.
    """
    dev = torch.device(device)
    g = torch.Generator(device=dev).manual_seed(seed * 1000003 + shard * 101 + 1)
    gperm = torch.Generator(device="cpu").manual_seed(seed * 31 + 5)
    perm = torch.randperm(n_features, generator=gperm).to(dev)
    # row lengths: log-normal around the target mean, at least 1
    sigma = 0.6
    mu = math.log(nnz_per_row * 1.12) - sigma * sigma / 2  # 1.12 ~ compensates de-duplication
    ln = torch.exp(mu + sigma * torch.randn(n_rows, generator=g, device=dev))
    ln = ln.clamp_(1, max(2.0, min(n_features / 2, 20 * nnz_per_row))).round_().to(torch.int64)
    row = torch.repeat_interleave(torch.arange(n_rows, device=dev), ln)
    total = row.numel()
    # skewed popularity: rank = floor(P * u^3)  (density ~ rank^(-2/3)), then permute ids
    u = torch.rand(total, generator=g, device=dev, dtype=torch.float64)
    rank = (u * u * u * n_features).to(torch.int64).clamp_(max=n_features - 1)
    fid = perm[rank]
    comb = torch.unique(row * n_features + fid)  # sorted, duplicates inside a row removed
    row = torch.div(comb, n_features, rounding_mode="floor")
    fid = comb - row * n_features
    counts = torch.bincount(row, minlength=n_rows)
    row_ptr = torch.zeros(n_rows + 1, dtype=torch.int64, device=dev)
    row_ptr[1:] = torch.cumsum(counts, 0)
    nnz = fid.numel()
    if binary:
        vals = None
        xv = torch.ones(nnz, dtype=torch.float64, device=dev)
    else:
        v = torch.rand(nnz, generator=g, device=dev, dtype=torch.float32) + 0.05
        sq = torch.zeros(n_rows, dtype=torch.float32, device=dev).index_add_(0, row, v * v)
        vals = v / sq.sqrt().clamp_min(1e-12)[row]
        xv = vals.double()
    if model_density is None:  # enough planted features that nearly every row carries signal
        model_density = min(0.5, max(0.02, 8.0 / max(nnz_per_row, 1.0) ))
    w_true = _planted_model(n_features, seed, density=model_density, device=dev)
    margin = torch.zeros(n_rows, dtype=torch.float64, device=dev).index_add_(0, row, xv * w_true[fid])
    margin = margin / margin.std().clamp_min(1e-12) * 2.0
    margin = margin + noise * torch.randn(n_rows, generator=g, device=dev, dtype=torch.float64)
    y = torch.where(margin >= 0, 1.0, -1.0).float()
    return SparseCSR(row_ptr, fid, vals, y, n_features)


def make_ctr_like(n_rows: int, n_groups: int = 8, keys_per_group: int = 1 << 20, seed: int = 0,
                  shard: int = 0, device="cpu", nnz_per_group: float = 3.0):
    """CTR-shaped binary data: per feature group, a few hashed 64-bit keys per example.

    Returns {group_id: SparseCSR} sharing `y`; keys carry the group id in the top 10 bits the
    way the reference's TERAFEA parser lays them out (src/data/text_parser.cc:129-181).
    """
    dev = torch.device(device)
    g = torch.Generator(device=dev).manual_seed(seed * 1000003 + shard * 101 + 7)
    margin = torch.zeros(n_rows, dtype=torch.float64, device=dev)
    out = {}
    for grp in range(1, n_groups + 1):
        ln = torch.poisson(torch.full((n_rows,), nnz_per_group, device=dev), generator=g).clamp_(min=1).long()
        row = torch.repeat_interleave(torch.arange(n_rows, device=dev), ln)
        u = torch.rand(row.numel(), generator=g, device=dev, dtype=torch.float64)
        rank = (u ** 4 * keys_per_group).to(torch.int64).clamp_(max=keys_per_group - 1)
        comb = torch.unique(row * keys_per_group + rank)
        row = torch.div(comb, keys_per_group, rounding_mode="floor")
        rank = comb - row * keys_per_group
        # multiplicative hash into 54 bits, group id in the top 10 bits
        h = (rank * (0x9E3779B97F4A7C15 - (1 << 64)) + ((grp * 0x632BE59BD9B4E019) & ((1 << 54) - 1))) & ((1 << 54) - 1)
        keys = h | (grp << 54)
        wg = _planted_model(keys_per_group, seed + grp, density=0.01, device=dev)
        margin.index_add_(0, row, wg[rank])
        counts = torch.bincount(row, minlength=n_rows)
        row_ptr = torch.zeros(n_rows + 1, dtype=torch.int64, device=dev)
        row_ptr[1:] = torch.cumsum(counts, 0)
        out[grp] = SparseCSR(row_ptr, keys, None, None, 1 << 63)
    margin = margin - margin.median() + 1.5  # CTR-like class imbalance
    noise = torch.randn(n_rows, generator=g, device=dev, dtype=torch.float64)
    y = torch.where(margin / margin.std().clamp_min(1e-12) + 0.5 * noise < 0, 1.0, -1.0).float()
    for grp in out:
        out[grp].y = y
    return out


def write_libsvm(data: SparseCSR, path: str) -> None:
    """Dump as LIBSVM text (label idx:val ...), ascending feature ids per line."""
    rp = data.row_ptr.cpu().tolist()
    keys = data.keys.cpu().tolist()
    vals = data.vals.cpu().tolist() if data.vals is not None else None
    y = data.y.cpu().tolist()
    with open(path, "w") as f:
        for r in range(data.n_rows):
            items = sorted((keys[e], 1.0 if vals is None else vals[e]) for e in range(rp[r], rp[r + 1]))
            f.write(("+1" if y[r] > 0 else "-1") + "".join(f" {k}:{v:.6g}" for k, v in items) + "\n")


def write_ps_sparse_binary(groups: dict, path: str) -> None:
    """Dump {group: SparseCSR} as PS SPARSE_BINARY text: `label; grp key key ...; grp ...;`."""
    gids = sorted(groups)
    y = groups[gids[0]].y.cpu().tolist()
    rps = {g: groups[g].row_ptr.cpu().tolist() for g in gids}
    keys = {g: [k & 0xFFFFFFFFFFFFFFFF for k in groups[g].keys.cpu().tolist()] for g in gids}
    with open(path, "w") as f:
        for r in range(len(y)):
            parts = ["1" if y[r] > 0 else "-1"]
            for g in gids:
                ks = keys[g][rps[g][r]:rps[g][r + 1]]
                if ks:
                    parts.append(f" {g} " + " ".join(str(k) for k in sorted(ks)))
            f.write(";".join(parts) + ";\n")
