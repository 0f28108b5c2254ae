#!/usr/bin/env python
"""Write seeded synthetic datasets in the layout the example configs expect (no network here:
this replaces the reference's download.sh).

    python example/linear/gen_data.py rcv1 [--rows 20000] [--parts 8]
    python example/linear/gen_data.py ctr  [--rows 20000] [--parts 8]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from parameter_server_b200.utils.synthetic import (make_ctr_like, make_rcv1_like, write_libsvm,
                                                   write_ps_sparse_binary)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("shape", choices=["rcv1", "ctr"])
    ap.add_argument("--rows", type=int, default=20000)
    ap.add_argument("--parts", type=int, default=8)
    ap.add_argument("--out", default="data")
    a = ap.parse_args()
    for split, seed_shift in (("train", 0), ("test", 1000)):
        d = os.path.join(a.out, a.shape, split)
        os.makedirs(d, exist_ok=True)
        for p in range(a.parts if split == "train" else 1):
            path = os.path.join(d, f"part-{p:03d}")
            if a.shape == "rcv1":
                write_libsvm(make_rcv1_like(a.rows // a.parts, seed=0, shard=p + seed_shift), path)
            else:
                write_ps_sparse_binary(make_ctr_like(a.rows // a.parts, seed=0, shard=p + seed_shift), path)
            print("wrote", path)


if __name__ == "__main__":
    main()
