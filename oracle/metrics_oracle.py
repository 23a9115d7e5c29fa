"""TEST INFRASTRUCTURE ONLY (imported by tests/ alone): the reference's metric arithmetic.

`explainn/architectures/__init__.py:458-461` returns the third-party functions themselves
(`sklearn.metrics.roc_auc_score`, `average_precision_score`, `scipy.stats.pearsonr`, `spearmanr`; the reference
pins scikit-learn=0.23.1 in `conda/create-conda-env.sh:20`, this image has scikit-learn 1.9 / scipy 1.18: the same
arithmetic; the one behavioural difference -- one class only in `y_true` raises ValueError in 0.23.1 and returns NaN
with a warning in 1.9 -- is tested against the pinned behaviour), and `selene/__init__.py:359-366` / `pwm-scoring.py:115-120` call them on flattened arrays / per column.
The oracle is therefore those same functions; parity is pinned by calling them, not by a restatement.
"""
from scipy.stats import pearsonr, spearmanr
from sklearn.metrics import average_precision_score, roc_auc_score


def get_metrics(input_data="binary"):
    if input_data == "binary":
        return dict(aucROC=roc_auc_score, aucPR=average_precision_score)
    return dict(Pearson=pearsonr, Spearman=spearmanr)


def column_metrics(y_true, scores, input_data="binary"):
    """The per-column loop of pwm-scoring.py:115-120."""
    out = {}
    for name, fn in get_metrics(input_data).items():
        vals = []
        for g in range(scores.shape[1]):
            yt = y_true[:, g] if y_true.ndim == 2 else y_true
            v = fn(yt, scores[:, g])
            vals.append(v if isinstance(v, float) else v[0])
        out[name] = vals
    return out
