"""GPU: device metrics (csrc/metrics.cu through the C ABI) against the functions the reference calls --
sklearn.metrics.roc_auc_score / average_precision_score, scipy.stats.pearsonr / spearmanr
(explainn/architectures/__init__.py:458-461; call sites selene/__init__.py:359-366, pwm-scoring.py:115-120).

Tolerance: 1e-12 absolute on AUROC (a ratio of exactly representable integers here and a trapezoid sum in sklearn),
1e-12 on AUPR and on Pearson / Spearman of float64 inputs (different fp64 summation order only).
"""
import numpy as np
import pytest
import torch

from oracle import metrics_oracle as MO

pytestmark = pytest.mark.gpu


def _cases():
    rng = np.random.default_rng(7)
    out = []
    for N, ties in [(1, None), (2, None), (37, None), (1000, None), (1000, 8), (4097, 3), (200003, None), (200003, 50)]:
        y = (rng.random(N) < 0.3).astype(np.int64)
        if N >= 2:
            y[0], y[1] = 0, 1                      # both classes present
        s = rng.standard_normal(N).astype(np.float32) + 0.7 * y.astype(np.float32)
        if ties:
            s = np.round(s * ties) / ties           # many exact ties, runs that mix both classes
        out.append((y, s.astype(np.float32)))
    return out


def test_auc_single_column_vs_sklearn():
    from explainn_b200 import metrics as M
    ref = MO.get_metrics("binary")
    mine = M.get_metrics("binary")
    assert list(mine) == list(ref) == ["aucROC", "aucPR"]
    for y, s in _cases():
        if len(y) < 2:
            continue
        for name in ref:
            a, b = mine[name](y, s), ref[name](y, s)
            assert isinstance(a, float)
            assert abs(a - b) <= 1e-12, (name, len(y), a, b)
    # torch inputs on the device, 2-D arrays flattened by the caller as in selene/__init__.py:360-361
    y, s = _cases()[3]
    a = mine["aucROC"](torch.tensor(y).cuda(), torch.tensor(s).cuda())
    assert abs(a - ref["aucROC"](y, s)) <= 1e-12


def test_auc_edge_cases_follow_sklearn():
    from explainn_b200 import metrics as M
    s = np.array([0.1, 0.4, 0.35, 0.8], dtype=np.float32)
    with pytest.raises(ValueError):
        M.roc_auc_score(np.zeros(4, dtype=np.int64), s)          # sklearn: only one class present
    with pytest.raises(ValueError):
        M.roc_auc_score(np.ones(4, dtype=np.int64), s)
    with pytest.raises(ValueError):
        M.roc_auc_score(np.array([0, 1, 2, 1]), s)               # not 0/1 labels
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        assert M.average_precision_score(np.zeros(4, dtype=np.int64), s) == MO.get_metrics()["aucPR"](np.zeros(4, dtype=np.int64), s)
    assert M.average_precision_score(np.ones(4, dtype=np.int64), s) == 1.0
    # all scores equal: one threshold
    y = np.array([0, 1, 1, 0, 1])
    c = np.full(5, 0.25, dtype=np.float32)
    assert abs(M.roc_auc_score(y, c) - 0.5) <= 1e-15
    assert abs(M.average_precision_score(y, c) - MO.get_metrics()["aucPR"](y, c)) <= 1e-15


def test_column_metrics_vs_per_column_loop():
    """pwm-scoring.py:115-120: one y_true, G score columns."""
    from explainn_b200 import metrics as M
    rng = np.random.default_rng(11)
    N, G = 5000, 23
    y = (rng.random(N) < 0.4).astype(np.int64)
    scores = (rng.standard_normal((N, G)) + 0.3 * y[:, None] * rng.random(G)[None, :]).astype(np.float32)
    scores[:, 5] = np.round(scores[:, 5] * 4) / 4            # a column full of ties
    scores[:, 6] = 1.0                                       # a constant column
    got = M.column_metrics(y, scores, "binary")
    want = MO.column_metrics(y, scores, "binary")
    for name in want:
        assert np.abs(got[name] - np.asarray(want[name])).max() <= 1e-12, name
    # per-column targets (the flattened (N,F) case split by feature)
    y2 = (rng.random((N, G)) < 0.5).astype(np.int64)
    got = M.column_metrics(y2, scores, "binary")
    want = MO.column_metrics(y2, scores, "binary")
    for name in want:
        assert np.abs(got[name] - np.asarray(want[name])).max() <= 1e-12, name


def test_pearson_spearman_vs_scipy():
    from explainn_b200 import metrics as M
    rng = np.random.default_rng(3)
    ref = MO.get_metrics("other")
    mine = M.get_metrics("other")
    assert list(mine) == list(ref) == ["Pearson", "Spearman"]
    for N, ties in [(2, None), (3, None), (1000, None), (1000, 5), (100003, None), (100003, 20)]:
        x = rng.standard_normal(N).astype(np.float32)
        y = (0.6 * x + rng.standard_normal(N)).astype(np.float32)
        if ties:
            x = (np.round(x * ties) / ties).astype(np.float32)
            y = (np.round(y * ties) / ties).astype(np.float32)
        for name in ref:
            a = mine[name](y, x)
            b = ref[name](y.astype(np.float64), x.astype(np.float64))
            assert abs(a[0] - float(b[0])) <= 1e-12, (name, N, a[0], float(b[0]))
    # batched regression columns
    N, G = 3000, 9
    yv = rng.standard_normal(N).astype(np.float32)
    sc = (rng.standard_normal((N, G)) + yv[:, None] * np.linspace(-1, 1, G)[None, :]).astype(np.float32)
    got = M.column_metrics(yv, sc, "other")
    want = MO.column_metrics(yv.astype(np.float64), sc.astype(np.float64), "other")
    for name in want:
        assert np.abs(got[name] - np.asarray(want[name], dtype=np.float64)).max() <= 1e-12, name
    # a constant input has no correlation: scipy returns nan
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        assert np.isnan(mine["Pearson"](np.ones(10, dtype=np.float32), np.arange(10, dtype=np.float32))[0])


def test_ranks_vs_scipy_rankdata():
    from scipy.stats import rankdata
    from explainn_b200 import ops
    rng = np.random.default_rng(5)
    v = np.round(rng.standard_normal((4, 20001)) * 3).astype(np.float32) / 3
    v[1] = 2.0
    v[2] = rng.standard_normal(20001).astype(np.float32)
    r = ops.metric_ranks(torch.tensor(v).cuda()).cpu().numpy()
    for g in range(4):
        assert np.array_equal(r[g], rankdata(v[g], method="average"))


@pytest.mark.parametrize("shape", [(1, 1), (3, 1023), (5, 1024), (7, 1025), (37, 10007), (750, 4000)])
def test_sort_columns_matches_torch_sort(shape):
    """xnn_sort_columns (one block per column, stable 8-bit LSD radix sort) against torch.sort: values identical,
    the permutation is a permutation that reproduces them, equal keys keep their original order (stability)"""
    from explainn_b200 import ops
    G, N = shape
    g = torch.Generator().manual_seed(G * 1000 + N)
    v = torch.randn(G, N, generator=g)
    v[:, ::7] = torch.round(v[:, ::7] * 2) / 2            # many ties, including +-0
    v[0, :min(N, 5)] = torch.tensor([0.0, -0.0, 3e38, -3e38, 1e-42])[:min(N, 5)]
    vd = v.cuda()
    for desc in (False, True):
        sv, perm = ops.sort_columns(vd, descending=desc)
        ref = torch.sort(v, dim=1, descending=desc, stable=True)
        assert torch.equal(sv.cpu(), ref.values)
        p = perm.cpu()
        assert torch.equal(torch.gather(v, 1, p), ref.values)
        assert torch.equal(torch.sort(p, dim=1).values, torch.arange(N).expand(G, N))
        # stability: inside a run of equal values the original indices ascend (-0.0 and +0.0 are distinct keys here:
        # compare with the bit pattern torch's stable sort would group only by value, so test value-equal AND sign-equal)
        same = (sv.cpu()[:, 1:] == sv.cpu()[:, :-1]) & (torch.signbit(sv.cpu()[:, 1:]) == torch.signbit(sv.cpu()[:, :-1]))
        assert bool(((p[:, 1:] > p[:, :-1]) | ~same).all())
