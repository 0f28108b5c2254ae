"""GPU front end of the linear-method apps: the device counterpart of `build/ps`.

    python -m parameter_server_b200.apps.ps --app_file example/linear/rcv1/batch_l1lr.conf
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        -m parameter_server_b200.apps.ps --app_file example/linear/ctr/online_l1lr.conf

One process per GPU; every rank is a worker and a server shard (rank 0 also plays the scheduler
and prints the progress tables).  The same protobuf text configs as the reference's
`script/local.sh build/ps ... -app_file` are accepted: `darlin {}` -> models/darlin.py,
`async_sgd {}` -> models/async_sgd.py, `validation_data + model_input` -> model evaluation.
ref: src/app/main/main.cc:15-39, src/app/linear_method/linear.cc:8-34
"""
from __future__ import annotations

import argparse
import os
import re
import sys

import numpy as np
import torch

from ..runtime import host
from ..utils.data_io import load_slot_data
from ..utils.synthetic import SparseCSR


def search_files(patterns: list) -> list:
    """Expand `dir/regex` entries (ref src/data/common.cc:100-150)."""
    out = []
    for p in patterns:
        d, base = os.path.split(p)
        d = d or "."
        if not os.path.isdir(d):
            continue
        rx = re.compile(base)
        out += [os.path.join(d, f) for f in sorted(os.listdir(d)) if rx.fullmatch(f) or f == base]
    return sorted(set(out))


def data_conf_for(conf: str, section: str, files: list) -> str:
    g = lambda k, d="": host.textproto_get(conf, f"linear_method.{section}.{k}") or d  # noqa: E731
    body = f'format: {g("format", "TEXT")} text: {g("text", "LIBSVM")}'
    if g("ignore_feature_group") == "true":
        body += " ignore_feature_group: true"
    return body + "".join(f' file: "{f}"' for f in files)


def conf_files(conf: str, section: str) -> list:
    out, i = [], 0
    while True:
        f = host.textproto_get(conf, f"linear_method.{section}.file[{i}]")
        if not f:
            return out
        out.append(f)
        i += 1


def merge_groups(groups: dict) -> SparseCSR:
    """All feature groups as one key space (async SGD treats the model as one KV store)."""
    gids = sorted(groups)
    if len(gids) == 1:
        return groups[gids[0]]
    n = groups[gids[0]].n_rows
    rows = [torch.repeat_interleave(torch.arange(n), groups[g].row_ptr[1:] - groups[g].row_ptr[:-1]) for g in gids]
    keys = torch.cat([groups[g].keys for g in gids])
    binary = all(groups[g].vals is None for g in gids)
    vals = None if binary else torch.cat([groups[g].vals if groups[g].vals is not None
                                          else torch.ones(groups[g].nnz) for g in gids])
    row = torch.cat(rows)
    order = torch.argsort(row, stable=True)
    rp = torch.zeros(n + 1, dtype=torch.int64)
    rp[1:] = torch.cumsum(torch.bincount(row, minlength=n), 0)
    return SparseCSR(rp, keys[order], None if vals is None else vals[order], groups[gids[0]].y, 1 << 63)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--app_file", "-app_file", required=True)
    ap.add_argument("--cache", default="", help="prefix of the columnar slot cache (default: from the config)")
    a = ap.parse_args(argv)
    conf = open(a.app_file).read()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    from ..parallel.comm import SingleComm, TorchComm
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        comm = TorchComm()
    else:
        comm = SingleComm()
    has = lambda k: bool(host.textproto_get(conf, "linear_method." + k))  # noqa: E731
    model_out = host.textproto_get(conf, "linear_method.model_output.file[0]")
    cache = a.cache or host.textproto_get(conf, "linear_method.darlin.local_cache.file[0]") or "/tmp/psb200_"
    cache = f"{cache}_r{rank}_"

    if has("darlin") or has("async_sgd"):
        files = search_files(conf_files(conf, "training_data"))
        mine = [f for i, f in enumerate(files) if i % world == rank]  # Postmaster::partitionData
        if rank == 0:
            print(f"Found {len(files)} files, {len(mine)} for every worker", file=sys.stderr)
        groups = load_slot_data(data_conf_for(conf, "training_data", mine), cache)
        if has("darlin"):
            from ..models.darlin import DarlinConfig, DarlinEngine
            cfg = DarlinConfig.from_proto_text(conf)
            eng = DarlinEngine(cfg, comm, dev)
            eng.load(groups)
            eng.train()
        else:
            from ..models.async_sgd import AsyncSGDConfig, AsyncSGDEngine
            cfg = AsyncSGDConfig.from_proto_text(conf)
            eng = AsyncSGDEngine(cfg, comm, dev)
            eng.load(merge_groups(groups))
            eng.train()
        if model_out:
            path = eng.save_model(model_out)
            print(f"S{rank} written the model to {path}", file=sys.stderr)
        eng.close()
    elif has("validation_data"):
        # ModelEvaluation (ref model_evaluation.h:19-74) with the K17 lookup SpMV + K14 AUC kernels
        from ..ops import evaluation as E
        from ..utils.model_io import read_model_text
        if rank == 0:
            mfiles = search_files(conf_files(conf, "model_input"))
            model = read_model_text(mfiles)
            keys = np.array(sorted(model), dtype=np.uint64)
            w = torch.tensor([model[int(k)] for k in keys], dtype=torch.float64, device=dev)
            dk = torch.from_numpy(keys.view(np.int64)).to(dev)
            files = search_files(conf_files(conf, "validation_data"))
            groups = load_slot_data(data_conf_for(conf, "validation_data", files) + " ignore_feature_group: true",
                                    cache + "val_")
            d = merge_groups(groups).to(dev)
            xw = E.spmv_lookup(d.row_ptr, d.keys, d.vals, dk, w)
            acc, ll = E.accuracy_logloss(d.y, xw)
            print(f"load {len(model)} model entries\nauc: {E.auc(d.y, xw):.6f}\naccuracy: {acc:.6f}\nlogloss: {ll:.6f}",
                  file=sys.stderr)
    else:
        print("nothing to do: the config has neither darlin, async_sgd nor validation_data", file=sys.stderr)
        return 1
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
