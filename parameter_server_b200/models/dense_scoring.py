"""Score several linear models on a DENSE feature group in one pass over the examples.

The reference's ModelEvaluation (src/app/linear_method/model_evaluation.h:19-74) streams the
validation data and evaluates ONE model with a hash-lookup SpMV; for PS-format DENSE slots
(src/data/text_parser.cc:200-250) the matrix is a plain [examples x features] array, and scoring K
candidate models - a regularisation path, the rounds of a hyper-parameter search - is a
[n x d] x [d x K] product.  That is the one GEMM-shaped operation of this framework and it runs on
the tcgen05 tensor cores (ops/dense.py, csrc/cuda/dense_tile.cu), the logistic loss of every model
reduced inside the kernel's epilogue.
"""
from __future__ import annotations

import torch

from ..ops import evaluation as E
from ..ops.dense import dense_margins, pad_features


def score_dense_models(x: torch.Tensor, label: torch.Tensor, weights: torch.Tensor, with_auc: bool = True) -> list:
    """x: [n, d] dense features (any float dtype, on the GPU), label: [n] in {-1, +1},
    weights: [K, d].  Returns one dict per model: logloss, accuracy and (optionally) exact AUC."""
    n, d = x.shape
    xb = x if (x.dtype == torch.bfloat16 and x.stride(0) % 8 == 0 and x.stride(1) == 1) else pad_features(x)
    wb = pad_features(weights)
    y, loss = dense_margins(xb, wb, label, d=d)
    out = []
    lab = label.float().contiguous()
    pos = (lab > 0)[:, None]
    acc = ((y > 0) == pos).float().mean(0).tolist()
    for k in range(weights.shape[0]):
        rec = {"logloss": float(loss[k]) / n, "accuracy": acc[k]}
        if with_auc:
            rec["auc"] = E.auc(lab, y[:, k].contiguous())
        out.append(rec)
    return out
