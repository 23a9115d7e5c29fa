"""Validation metrics on the GPU behind the reference's `get_metrics` surface.

Reference: `explainn/architectures/__init__.py:458-461` returns
`dict(aucROC=roc_auc_score, aucPR=average_precision_score)` for binary data and
`dict(Pearson=pearsonr, Spearman=spearmanr)` otherwise; `selene/__init__.py:359-366` calls every entry as
`metric(y_true.flatten(), y_score.flatten())` and takes `score[0]` when the result is not a float;
`pwm-scoring.py:115-120` calls them once per PWM column (750 sklearn calls over the same `y_true`).

Here the arrays stay on the device: one segmented radix sort per column batch (`xnn_sort_columns`, csrc/sort.cu) and the kernels of
`csrc/metrics.cu`.  `get_metrics()` keeps the reference's names and call signature (numpy or torch inputs, float /
`(r, pvalue)` results); `column_metrics` is the batched form for the per-PWM loop.  No CPU fallback: inputs are moved
to the CUDA device and the kernels raise without one.
"""
import numpy as np
import torch

from . import ops


def _dev(a, dtype, device="cuda"):
    t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
    return t.to(device=device, dtype=dtype).contiguous()


def _binary(y_true):
    t = y_true if isinstance(y_true, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(y_true))
    vals = torch.unique(t)
    if vals.numel() > 2 or not bool(torch.all((vals == 0) | (vals == 1))):
        # sklearn accepts any two labels and multiclass targets; the reference only ever passes 0/1 (train.py targets)
        raise ValueError("explainn_b200.metrics: y_true must hold 0/1 labels")
    return t


def roc_auc_score(y_true, y_score):
    """`sklearn.metrics.roc_auc_score(y_true, y_score)` for 0/1 labels (one flattened column).  One class only raises
    ValueError as the reference's pinned scikit-learn 0.23.1 does (`conda/create-conda-env.sh:20`)."""
    y = _binary(y_true)
    auroc, _ = ops.metric_auc(_dev(y_score, torch.float32).reshape(1, -1), _dev(y, torch.uint8).reshape(-1))
    v = float(auroc[0])
    if v != v:
        raise ValueError("Only one class present in y_true. ROC AUC score is not defined in that case.")
    return v


def average_precision_score(y_true, y_score):
    """`sklearn.metrics.average_precision_score(y_true, y_score)` for 0/1 labels."""
    y = _binary(y_true)
    _, aupr = ops.metric_auc(_dev(y_score, torch.float32).reshape(1, -1), _dev(y, torch.uint8).reshape(-1))
    v = float(aupr[0])
    return 0.0 if v != v else v     # sklearn: no positive class -> recall set to one, AP 0.0 (with a warning)


def pearsonr(x, y):
    """`scipy.stats.pearsonr(x, y)`: returns `(r, None)` -- the reference reads `score[0]` only."""
    xt = _dev(x, torch.float64).reshape(1, -1)
    r = ops.metric_pearson(xt, _dev(y, torch.float64).reshape(-1))
    return float(r[0]), None


def spearmanr(x, y):
    """`scipy.stats.spearmanr(x, y)`: Pearson r of the average ranks; returns `(rho, None)`."""
    rx = ops.metric_ranks(_dev(x, torch.float32).reshape(1, -1))
    ry = ops.metric_ranks(_dev(y, torch.float32).reshape(1, -1))
    r = ops.metric_pearson(rx, ry)
    return float(r[0]), None


def get_metrics(input_data="binary"):
    """Same keys and call signature as `explainn/architectures/__init__.py:458-461`."""
    if input_data == "binary":
        return dict(aucROC=roc_auc_score, aucPR=average_precision_score)
    return dict(Pearson=pearsonr, Spearman=spearmanr)


def column_metrics(y_true, scores, input_data="binary"):
    """All columns of `scores` (N,G) against one `y_true` (N,) or per-column `y_true` (N,G) in one pass -- the
    per-PWM loop of `pwm-scoring.py:115-120`.  Returns {name: float64 numpy array (G,)} with `get_metrics`' keys."""
    s = scores if isinstance(scores, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(scores))
    if s.dim() != 2:
        raise ValueError("scores: expected (N,G)")
    cols = s.to(device="cuda", dtype=torch.float32).t().contiguous()          # (G,N): columns contiguous
    yt = y_true if isinstance(y_true, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(y_true))
    per_col = yt.dim() == 2
    if input_data == "binary":
        y = _dev(_binary(yt), torch.uint8)
        y = y.t().contiguous() if per_col else y.reshape(-1)
        auroc, aupr = ops.metric_auc(cols, y)
        return dict(aucROC=auroc.cpu().numpy(), aucPR=aupr.cpu().numpy())
    y32 = _dev(yt, torch.float32)
    y32 = y32.t().contiguous() if per_col else y32.reshape(1, -1)
    pear = ops.metric_pearson(cols.double(), (y32.double() if per_col else y32.double().reshape(-1)))
    ry = ops.metric_ranks(y32)
    spear = ops.metric_pearson(ops.metric_ranks(cols), ry if per_col else ry.reshape(-1))
    return dict(Pearson=pear.cpu().numpy(), Spearman=spear.cpu().numpy())
