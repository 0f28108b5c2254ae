"""Load training data through the native SlotReader (text parsing + columnar cache in C++) into
the SparseCSR groups the device engines consume."""
from __future__ import annotations

import numpy as np
import torch

from ..runtime import host
from .synthetic import SparseCSR


def parse_example_info(info_text: str) -> dict:
    """ExampleInfo text -> {slot_id: dict(format,min_key,max_key,nnz_ele,nnz_ex)}, num_ex."""
    slots, i = {}, 0
    while True:
        s = host.textproto_get(info_text, f"slot[{i}]")
        if not s:
            break
        get = lambda k, d="0": host.textproto_get(info_text, f"slot[{i}].{k}") or d  # noqa: E731
        slots[int(get("id"))] = dict(format=get("format", "SPARSE"), min_key=int(get("min_key")),
                                     max_key=int(get("max_key")), nnz_ele=int(get("nnz_ele")),
                                     nnz_ex=int(get("nnz_ex")))
        i += 1
    num_ex = int(host.textproto_get(info_text, "num_ex") or 0)
    return slots, num_ex


def load_slot_data(data_conf: str, cache_prefix: str, device="cpu") -> dict:
    """{group: SparseCSR} for every feature slot of the files named by the DataConfig text."""
    info_text = host.slot_reader_info(data_conf, cache_prefix)
    slots, num_ex = parse_example_info(info_text)
    _, _, yv = host.slot_reader_arrays(data_conf, cache_prefix, 0, cap=max(num_ex, 1) + 16)
    y = torch.from_numpy(np.where(yv > 0, 1.0, -1.0).astype(np.float32))
    out = {}
    for sid, s in sorted(slots.items()):
        if sid == 0:
            continue
        cap = s["nnz_ele"] + num_ex + 16
        idx, off, val = host.slot_reader_arrays(data_conf, cache_prefix, sid, cap=cap)
        if off.size == 0:
            off = np.zeros(num_ex + 1, dtype=np.int64)
        keys = torch.from_numpy(idx.view(np.int64).copy())
        vals = torch.from_numpy(val.copy()) if val.size == idx.size and val.size else None
        out[sid] = SparseCSR(torch.from_numpy(off.copy()).to(device), keys.to(device),
                             None if vals is None else vals.to(device), y.to(device), s["max_key"])
    return out
