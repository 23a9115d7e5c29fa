"""CPU: host-side logic and the C-ABI surface (no compute calls -- there is no GPU here)."""
import ctypes
import io
import os

import numpy as np
import pytest
import torch

import _util as U
from explainn_b200 import _lib


def test_library_exports_every_declared_symbol():
    protos = _lib.parse_header()
    assert len(protos) >= 30
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in protos:
        assert hasattr(lib, name), f"{name} declared in include/explainn_b200.h but not exported"
    assert _lib.load().xnn_version() >= 100


def test_work_size_queries_need_no_gpu():
    lib = _lib.load()
    assert lib.xnn_bn_work_bytes(4096, 30000) > 0
    assert lib.xnn_fc1_bwd_work_bytes(4096, 300, 83) > 0
    # slices x units x (4k+2) doubles; the slice count follows the SM count (148 when no device is visible)
    assert lib.xnn_frontend_bwd_work_bytes(4096, 300, 19) % (300 * 78 * 8) == 0
    assert lib.xnn_frontend_bwd_work_bytes(4096, 300, 19) >= 64 * 300 * 78 * 8
    assert lib.xnn_mlp_bn2_work_bytes(4096, 300) >= 2 * 2 * 30000 * 8
    assert lib.xnn_mlp_supported(83) == 1 and lib.xnn_mlp_supported(97) == 0


def test_module_surface_matches_reference_state_dict():
    from explainn_b200.architectures import ExplaiNN
    d = U.load_npz("eval_small.npz")
    Uu, k, L, F, _ = [int(v) for v in d["opts"]]
    m = ExplaiNN(Uu, k, L, F)
    sd = m.state_dict()
    ref = U.golden_sd(d)
    assert list(sd.keys()) == list(ref.keys())
    for key in sd:
        assert tuple(sd[key].shape) == tuple(ref[key].shape), key
    m.load_state_dict(ref)
    assert m._options == {"cnn_units": Uu, "kernel_size": k, "sequence_length": L, "n_features": F,
                          "weights_file": None}
    assert m.linears[0].weight.shape == (Uu, 4, k) and m.final.weight.shape == (F, Uu)
    assert m.linears[:3]._kind() == "acts" and m.linears[:4]._kind() == "pool"
    # same construction order as the reference -> identical default initialisation under the same seed
    torch.manual_seed(0)
    m2 = ExplaiNN(100, 19, 200, 1)
    c = U.load_npz("eval_cfg1.npz")
    assert np.array_equal(m2.linears[0].weight.detach().numpy(), c["sd.linears.0.weight"])
    assert np.array_equal(m2.final.weight.detach().numpy(), c["sd.final.weight"])


def test_kernel_size_limits_agree_with_the_library():
    from explainn_b200 import ops
    from explainn_b200.architectures import ExplaiNN
    assert ops.kernel_size_limit(0) == ops.KERNEL_SIZE_MAX and ops.kernel_size_limit(1) == ops.KERNEL_SIZE_MAX_TRAIN
    ExplaiNN(3, ops.KERNEL_SIZE_MAX, 120, 1)                        # constructs (eval / interpret paths)
    with pytest.raises(NotImplementedError, match="kernel_size"):
        ExplaiNN(3, ops.KERNEL_SIZE_MAX + 1, 120, 1)


def test_reference_call_sites_on_the_module_object(tmp_path):
    """what the reference's trainer does to the model object besides calling it.
    copy.deepcopy (selene/__init__.py:375), the checkpoint dict (:376-384) and its reload (:208-221),
    `_Model.load_weights` (architectures/__init__.py:27-39), weight re-assignment (train.py:316-324)."""
    import copy
    from explainn_b200.architectures import ExplaiNN, get_optimizer
    torch.manual_seed(0)
    m = ExplaiNN(6, 19, 67, 2)
    m2 = copy.deepcopy(m)
    assert type(m2) is ExplaiNN and m2._options == m._options and m2._options is not m._options
    for (k1, v1), (k2, v2) in zip(m.state_dict().items(), m2.state_dict().items()):
        assert k1 == k2 and torch.equal(v1, v2) and v1.data_ptr() != v2.data_ptr()
    # checkpoint dict round trip through torch.save / torch.load, optimizer state included
    opt = get_optimizer(m.parameters(), 0.003)
    ckpt = {"step": 7, "arch": m.__class__.__name__, "options": m._options, "state_dict": m.state_dict(),
            "min_loss": 0.5, "optimizer": copy.deepcopy(opt.state_dict())}
    path = tmp_path / "best_model.pth.tar"
    torch.save(ckpt, path)
    back = torch.load(path, weights_only=False)
    assert back["arch"] == "ExplaiNN" and back["options"] == m._options
    m3 = ExplaiNN(**{k: v for k, v in back["options"].items() if k != "weights_file"})
    m3.load_state_dict(back["state_dict"])
    get_optimizer(m3.parameters(), 0.003).load_state_dict(back["optimizer"])
    # the same state dict loads into a stock torch.optim.Adam and vice versa (FusedAdam keeps Adam's layout)
    ref_opt = torch.optim.Adam(m3.parameters(), lr=0.003)
    ref_opt.load_state_dict(back["optimizer"])
    assert ref_opt.state_dict()["param_groups"][0]["lr"] == 0.003
    get_optimizer(m3.parameters(), 0.003).load_state_dict(ref_opt.state_dict())
    # legacy positional loader: values are matched to keys by POSITION and a trailing size-1 dimension of every
    # >= 2-d value is squeezed -- so a legacy file (one extra trailing dimension) loads under any key names ...
    legacy = {f"w{i}": (v.unsqueeze(-1) if v.dim() >= 1 else v) for i, v in enumerate(m.state_dict().values())}
    wpath = tmp_path / "legacy.pth"
    torch.save(legacy, wpath)
    m4 = ExplaiNN(6, 19, 67, 2, weights_file=str(wpath))
    assert m4._options["weights_file"] == str(wpath)
    for v1, v2 in zip(m.state_dict().values(), m4.state_dict().values()):
        assert torch.equal(v1, v2)
    # ... and, exactly like the reference's loader, the module's OWN state dict does not (its (100U,n,1) weights
    # lose their last dimension): same exception type as the reference raises
    torch.save(m.state_dict(), wpath)
    with pytest.raises(RuntimeError, match="size mismatch"):
        ExplaiNN(6, 19, 67, 2, weights_file=str(wpath))
    # transfer learning: `linears[0].weight = nn.Parameter(...)` replaces the registered parameter
    w = m.linears[0].weight.data
    w[0] = 1.0
    w.requires_grad = True
    m.linears[0].weight = torch.nn.Parameter(w)
    assert "linears.0.weight" in dict(m.named_parameters())
    assert float(m.state_dict()["linears.0.weight"][0].min()) == 1.0


def test_no_cpu_fallback():
    from explainn_b200.architectures import ExplaiNN, PWM
    m = ExplaiNN(4, 5, 30, 1).eval()
    x = torch.zeros(2, 4, 30)
    with pytest.raises(RuntimeError):
        m(x)
    with pytest.raises(RuntimeError):
        m.linears[0](x)           # layers are parameter holders, not a torch-library path
    with pytest.raises(RuntimeError):
        PWM(np.zeros((2, 4, 5)), 30)(x)


def test_get_pwms_matches_reference():
    from explainn_b200 import pwm_scoring
    d = U.load_npz("pwm_small.npz")
    pw, names = pwm_scoring.get_pwms(os.path.join(U.GOLDEN, "small.meme"), resize_pwms=True, return_log=True)
    assert list(names) == list(d["names"])
    assert np.array_equal(np.asarray(pw), d["pwms_log"])   # bit-exact: same float ops on the host
    raw, _ = pwm_scoring.get_pwms(os.path.join(U.GOLDEN, "small.meme"))
    assert [len(p[0]) for p in raw] == list(d["raw_widths"])


def test_resize_tie_rule():
    from explainn_b200.pwm_scoring import resize_pwm
    cols = [[.7, .1, .1, .1], [.25, .25, .25, .25]]
    out = resize_pwm(cols, 4)
    assert len(out) == 4 and out.count([.25, .25, .25, .25]) == 3


def test_sequence_helpers():
    from explainn_b200 import sequence as S
    from oracle import explainn_oracle as O
    s = "ACGTNacgtnRYA"
    assert np.array_equal(S.one_hot_encode(s), O.one_hot_encode(s))
    assert S.one_hot_decode(S.one_hot_encode(s)) == "ACGTNACGTNNNA"
    assert S.rc("ACGTNNa") == "tNNACGT"
    e = S.one_hot_encode("AACGT")
    assert np.array_equal(S.rc_one_hot_encoding(e), S.one_hot_encode("ACGTT"))


def test_fasta_and_matrix_readers():
    from explainn_b200 import pwmscan
    g = U.pwmscan_golden()
    headers, codes, offsets = pwmscan.read_fasta(io.StringIO(g["fasta"]))
    assert "withN" in headers and "multi" in headers and len(headers) == len(offsets) - 1
    i = headers.index("multi")
    assert offsets[i + 1] - offsets[i] == 45          # two sequence lines joined
    i = headers.index("lower")
    assert set(codes[offsets[i]:offsets[i + 1]].tolist()) <= {0, 1, 2, 3}
    assert (codes[offsets[headers.index("withN")]:offsets[headers.index("withN") + 1]] == 4).sum() == 5
    tmp = os.path.join(U.ROOT, "tests", "golden", "_tmp.lpm")
    try:
        open(tmp, "w").write(g["lpm"])
        m = pwmscan.read_matrix(tmp)
        assert m.shape == (4, 19) and np.allclose(m, U.parse_matrix_text(g["lpm"]))
    finally:
        os.remove(tmp)


def test_jaspar_writer_matches_fixture():
    import json
    from explainn_b200.interpret import pfm_to_jaspar
    from oracle import explainn_oracle as O
    fx = json.load(open(os.path.join(U.GOLDEN, "sites_jaspar_fixtures.json")))
    for name, pair in fx.items():
        pfm = O.parse_jaspar(pair["jaspar"])
        title = pair["jaspar"].splitlines()[0][1:].split(" ", 1)
        assert pfm_to_jaspar(pfm, title[0], title[1]).strip() == pair["jaspar"].strip(), name


def test_metrics_surface_and_no_cpu_fallback():
    """get_metrics keeps the reference's keys (architectures/__init__.py:458-461); the ops refuse CPU tensors."""
    import numpy as np
    import torch
    from explainn_b200 import metrics, ops
    from oracle import metrics_oracle as MO
    assert list(metrics.get_metrics("binary")) == list(MO.get_metrics("binary"))
    assert list(metrics.get_metrics("other")) == list(MO.get_metrics("other"))
    with pytest.raises(RuntimeError):
        ops.metric_auc(torch.zeros((1, 4)), torch.zeros(4, dtype=torch.uint8))
    with pytest.raises(RuntimeError):
        ops.metric_pearson(torch.zeros((1, 4)), torch.zeros(4))
    with pytest.raises(RuntimeError):
        ops.metric_ranks(torch.zeros((1, 4)))
    lib = _lib.load()
    assert lib.xnn_metric_auc_work_bytes(1000, 3) == 12000
    # sklearn's own answers on the two-threshold example of its documentation pin the oracle wrapper
    y, s = np.array([0, 0, 1, 1]), np.array([0.1, 0.4, 0.35, 0.8])
    assert abs(MO.get_metrics()["aucROC"](y, s) - 0.75) < 1e-15
    assert abs(MO.get_metrics()["aucPR"](y, s) - 0.8333333333333333) < 1e-15


def test_pwm_scoring_command_line_errors(tmp_path):
    """a15: usage / exit codes / stderr of the command line (pwm_scoring.c:799-821, 893-899, 906-913); no GPU needed
    for the error paths"""
    import io
    from explainn_b200 import pwmscan
    err = io.StringIO()
    assert pwmscan.main(["pwm_scoring"], stderr=err) == 1
    assert err.getvalue().startswith("Usage: pwm_scoring [options] -m <matrix_file> [<] <fasta_file>")
    assert "--pwm                  Input matrix is a position weight matrix (PWM)" in err.getvalue()
    err = io.StringIO()
    assert pwmscan.main(["pwm_scoring", "-m", str(tmp_path / "missing.lpm"), "x.fa"], stderr=err) == 1
    assert "Unable to open" in err.getvalue()
    g = U.pwmscan_golden()
    lpm = tmp_path / "g.lpm"
    lpm.write_text(g["lpm"])
    err = io.StringIO()
    assert pwmscan.main(["pwm_scoring", "-p", "0.3,0.2,0.5", "-m", str(lpm), "x.fa"], stderr=err) == 1
    assert err.getvalue().splitlines()[0] == "Number of TOKENS: 3"
    err = io.StringIO()
    assert pwmscan.main(["pwm_scoring", "-m", str(lpm), str(tmp_path / "missing.fa")], stderr=err) == 1
    assert err.getvalue().startswith(f"Unable to open '{tmp_path / 'missing.fa'}': No such file or directory(2)")
    # an empty FASTA is not an error: no output, exit 0 (never reaches a kernel)
    empty = tmp_path / "e.fa"
    empty.write_text("")
    out = io.StringIO()
    assert pwmscan.main(["pwm_scoring", "-m", str(lpm), str(empty)], stdout=out, stderr=io.StringIO()) == 0
    assert out.getvalue() == ""


def test_interpret_postprocessing_matches_reference_functions():
    """f3: well-predicted selection (interpret.py:315-361) and per-filter importance filtering (:485-490) against the
    outputs of the reference's own functions (tests/golden/postproc.npz), the importance table against its
    definition, and the MEME writer's format (utils/jaspar2others.py:46-83)."""
    from explainn_b200 import interpret as I
    d = U.load_npz("postproc.npz")
    for tag, kind in (("bin", "binary"), ("lin", "linear")):
        for rc in (False, True):
            got = I.well_predicted_sequences(d[f"{tag}.preds"], d[f"{tag}.labels"], kind, rc)
            assert np.array_equal(np.sort(got), np.sort(d[f"{tag}.idxs.rc{int(rc)}"])), (tag, rc)
    out = I.filtered_filter_importances(d["imp.imps"], d["imp.idxs"], d["imp.acts"], d["imp.thr"])
    assert np.array_equal(out, d["imp.out"])
    # importance table: the same rows through the `fired` matrix instead of the (N, L') activations
    rng = np.random.default_rng(0)
    outs, W = rng.random((40, 5)), rng.standard_normal((3, 5))
    acts = (rng.random((40, 5, 30)) * 3).astype(np.float16)
    thr = np.full(5, 2.9, dtype=np.float16)
    thr[4] = np.float16(9.0)                                        # a filter that fires nowhere
    idxs = np.arange(0, 40, 2)
    fired = (acts > thr[None, :, None]).any(axis=2).T
    rows, med = I.filter_importance_table(outs, W, idxs, fired)
    imps = np.array([outs * W[f] for f in range(3)])                # interpret.py:180-181
    k = 0
    for u in range(5):
        ref = I.filtered_filter_importances(imps[:, :, u], idxs, acts[:, u, :], thr[u]).T
        for r in ref:
            assert rows[k][0] == u and np.allclose(rows[k][1], r, rtol=0, atol=1e-15)
            k += 1
        if len(ref):
            assert np.allclose(med[u], np.median(ref, axis=0))
        else:
            assert np.isnan(med[u]).all()
    assert k == len(rows)
    pfm = np.array([[[3, 0, 1], [0, 4, 1], [1, 0, 1], [0, 0, 1]], [[0, 0, 0], [0, 0, 0], [0, 0, 0], [0, 0, 0]]])
    text = I.pfms_to_meme(pfm)
    assert text.startswith("MEME version 4\n\nALPHABET= ACGT\n\nstrands: + -\n\n")
    assert "\nMOTIF filter0 filter0 ACA\nletter-probability matrix: alength= 4 w= 3 nsites= 4 E= 0\n" in text
    assert "       0.75         0.0        0.25         0.0\n" in text and "filter1" not in text


def test_interpret_chunks_are_plane_slices_or_shifted_index_slices():
    """interpret._chunk_of: without a selection a chunk is a contiguous slice of the planes (no index list); with one,
    the device-resident list sliced and shifted by the strand's base (interpret.py:375-421 walks fwd then rev halves)."""
    import torch
    from explainn_b200 import interpret as I
    planes = torch.arange(10 * 3 * 2, dtype=torch.int32).reshape(10, 3, 2)
    pl, idx = I._chunk_of(planes, None, 5, 2, 3)
    assert idx is None and torch.equal(pl, planes[7:10])
    sel = torch.tensor([4, 1, 3, 0], dtype=torch.int32)
    pl, idx = I._chunk_of(planes, sel, 0, 1, 2)
    assert pl is planes and idx.tolist() == [1, 3] and idx.is_contiguous()
    pl, idx = I._chunk_of(planes, sel, 5, 2, 2)
    assert pl is planes and idx.tolist() == [8, 5] and idx.dtype == torch.int32


def test_noalloc_limits_are_host_queries():
    """xnn_pwm_scan_tc_noalloc: the static scratch size and the work size it is compared with need no GPU"""
    from explainn_b200 import _lib
    lib = _lib.load()
    cap = lib.xnn_pwm_scan_tc_noalloc_max_bytes()
    assert cap == 1 << 20
    assert 0 < lib.xnn_pwm_scan_tc_work_bytes(750, 35) <= cap          # the cfg5 scan fits
    assert lib.xnn_pwm_scan_tc_work_bytes(1152, 35) <= cap < lib.xnn_pwm_scan_tc_work_bytes(1153, 35)
    assert lib.xnn_pwm_scan_tc_work_bytes(896, 48) <= cap < lib.xnn_pwm_scan_tc_work_bytes(897, 48)
