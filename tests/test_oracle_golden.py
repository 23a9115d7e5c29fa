"""CPU: the oracle (oracle/) against the golden vectors generated from the real reference, the
reference's own sites->JASPAR fixtures and the text output of the reference C tool."""
import io
import json
import os

import numpy as np
import pytest
import torch

import _util as U
from oracle import explainn_oracle as O

torch.set_num_threads(max(1, min(8, os.cpu_count() or 1)))


@pytest.mark.parametrize("name", ["eval_small.npz", "eval_cfg1.npz", "eval_k11.npz"])
def test_eval_forward_matches_reference(name):
    d = U.load_npz(name)
    out = O.explainn_forward(U.golden_sd(d), torch.tensor(d["x"]), train=False)
    # fp32 restatement vs fp32 reference: both within ~1e-6 of fp64
    U.assert_close(out["logits"].numpy(), d["logits"], rtol=5e-6, what="logits")
    U.assert_close(out["outs"].numpy(), d["outs"], rtol=5e-6, what="unit outputs")
    U.assert_close(out["acts"].numpy(), d["acts"], rtol=5e-6, what="activations")


@pytest.mark.parametrize("name,kind", [("train_mse.npz", "linear"), ("train_bce.npz", "binary")])
def test_train_step_matches_reference(name, kind):
    d = U.load_npz(name)
    sd = U.golden_sd(d)
    r = O.explainn_train_step(sd, torch.tensor(d["x"]), torch.tensor(d["y"]), kind)
    U.assert_close(r["loss"].numpy(), d["loss"], rtol=1e-5, what="loss")
    U.assert_close(r["logits"].numpy(), d["logits"], rtol=1e-5, what="logits")
    zero_class = {"linears.0.bias", "linears.1.bias", "linears.6.bias", "linears.10.bias"}
    for k in O.PARAM_KEYS:
        ref = d["grad." + k]
        got = r["grads"][k].numpy().reshape(ref.shape)
        if k in zero_class:
            assert np.abs(got - ref).max() < 1e-6, k  # round-off noise around an exactly-zero gradient
        else:
            U.assert_close(got, ref, rtol=2e-4, atol_scale=2e-5, what="grad " + k)
    for k, v in r["running"].items():
        U.assert_close(v.numpy(), d["after." + k], rtol=1e-5, what=k)
    # Adam (lr 0.003) on the golden gradients reproduces the reference's updated parameters
    for k in O.PARAM_KEYS:
        p0, g = torch.tensor(d["sd." + k]), torch.tensor(d["grad." + k])
        p1, _, _ = O.adam_step(p0, g, torch.zeros_like(p0), torch.zeros_like(p0), 1, lr=0.003)
        U.assert_close(p1.numpy(), d["after." + k], rtol=1e-6, what="adam " + k)


def test_train_step_cfg3_matches_reference():
    """the headline workload at its own shape: ExplaiNN(300,19,600,50), B=32 (reference-generated fixture)"""
    d, m, x, y = U.cfg3_golden()
    sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
    r = O.explainn_train_step(sd, x, y, "linear")
    U.assert_close(r["loss"].numpy(), d["loss"], rtol=1e-5, what="loss")
    U.assert_close(r["logits"].numpy(), d["logits"], rtol=1e-5, what="logits")
    U.check_cfg3_grads(d, {k: r["grads"][k].numpy() for k in O.PARAM_KEYS}, "oracle grad")
    for k, v in r["running"].items():
        U.assert_close(v.numpy(), d["after." + k], rtol=1e-5, what=k)


def test_interpret_thresholds_and_sites():
    d = U.load_npz("interpret_small.npz")
    e = U.load_npz("eval_small.npz")
    k = int(e["opts"][1])
    acts16 = d["acts16"]
    # the fp16 array itself: oracle forward -> fp16 must match the reference's stored array except at
    # fp16 rounding knife edges (counted)
    out = O.explainn_forward(U.golden_sd(e), torch.tensor(d["x2"]), train=False)
    mine16 = out["acts"].numpy().astype(np.float16)
    flips = int((mine16 != acts16).sum())
    assert flips <= max(3, acts16.size // 5000), flips
    for sfx, rc in (("", False), ("_rc", True)):
        idxs = d["idxs" + sfx]
        thr = O.act_thresholds(acts16, idxs, rc)
        assert thr.dtype == np.float16 and np.array_equal(thr, d["thr" + sfx])
        for u in range(acts16.shape[1]):
            got = O.sites(acts16[:, u, :], idxs, thr[u], rc)
            text = "".join(f">seq{i + (len(acts16) // 2 if (rc and s) else 0)}_{'-' if s else '+'}_from={p}_to={p + k}\n"
                           for i, s, p in got)
            ref_headers = "".join(ln + "\n" for ln in str(d["sites_fasta" + sfx][u]).splitlines() if ln.startswith(">"))
            assert text == ref_headers, (sfx, u)


def test_sites_to_pfm_fixtures():
    fx = json.load(open(os.path.join(U.GOLDEN, "sites_jaspar_fixtures.json")))
    assert len(fx) >= 2
    for name, pair in fx.items():
        kmers = O.parse_sites_fasta(pair["sites"])
        ref = O.parse_jaspar(pair["jaspar"])
        pfm = O.sites_to_pfm(kmers, ref.shape[1])
        assert np.array_equal(pfm, ref.astype(np.int64)), name


def test_sites_to_pfm_all_126_reference_fixtures():
    """every results/Human-Mouse-Project/CAM/sites/filter*.fa -> motifs/filter*.jaspar pair of the reference
    (k-mers stored as ASCII bytes): the restated per-column counts equal the Biopython output 126 / 126"""
    d = U.load_npz("sites_jaspar_all.npz")
    ids = d["ids"]
    assert len(ids) == 126
    for i in ids:
        kmers = ["".join(map(chr, row)) for row in d[f"kmers{i}"]]
        ref = O.parse_jaspar(str(d[f"jaspar{i}"]))
        assert np.array_equal(O.sites_to_pfm(kmers, ref.shape[1]), ref.astype(np.int64)), f"filter{i}"


def test_pwm_module_scores():
    d = U.load_npz("pwm_small.npz")
    for scoring in ("max", "sum"):
        got, _ = O.pwm_module_scores(d["pwms_log"], d["x"], scoring)
        U.assert_close(got, d[scoring], rtol=2e-6, what=scoring)


def _records(g):
    from explainn_b200 import pwmscan
    return pwmscan.read_fasta(io.StringIO(g["fasta"]))


def test_c_oracle_matches_reference_tool():
    """oracle/pwmscan_oracle.c vs stdout of the reference's pwm_scoring (oracle/_ref) on its own fixture"""
    g = U.pwmscan_golden()
    headers, codes, offsets = _records(g)
    lpm = U.parse_matrix_text(g["lpm"])
    ipwm = U.parse_matrix_text(g["pwm"], integer=True)
    w = lpm.shape[1]
    lpm5 = np.vstack([lpm, np.full((1, w), 0.25)])
    pwm5 = np.vstack([ipwm, np.full((1, w), -2 ** 31)]).astype(np.int32)
    base_bg = np.array([1, 1, 1, 1, 0.25])
    for name, run in g["runs"].items():
        opts = run["opts"]
        lines = run["stdout"].splitlines()
        assert len(lines) == len(headers)
        fwd = "-f" in opts
        if "--pwm" in opts:
            best, pos, strand = U.oracle_ipwm_best(codes, offsets, pwm5, fwd)
            for r, ln in enumerate(lines):
                f = ln.split("\t")
                if f[3] == "NOTAG":
                    assert pos[r] == -1
                    continue
                assert (int(f[1]), int(f[4]), f[5]) == (pos[r], best[r], "-" if strand[r] else "+"), (name, ln)
            continue
        m5, bg = lpm5.copy(), base_bg.copy()
        if "-w" in opts:
            pw = float(opts[opts.index("-w") + 1])
            m5[:4] = (m5[:4] + pw) / (m5[:4] + pw).sum(0)
        if "-u" in opts:
            bg[:4] = 0.25
        if "-p" in opts:
            bg[:4] = [float(v) for v in opts[opts.index("-p") + 1].split(",")]
        if "-q" in opts:
            from explainn_b200.pwmscan import PwmScoring
            ps = PwmScoring.__new__(PwmScoring)
            ps.bg, ps.forward = bg, fwd
            bg = ps._seq_bg(codes, offsets)
        if "-b" in opts:
            best, pos, strand, nties, ties = U.oracle_lpm_best(codes, offsets, m5, bg, fwd, tie_cap=64)
            for r, ln in enumerate(lines):
                f = ln.split("\t")
                f = f if "-r" in opts else f[1:]
                plist = ",".join([str(pos[r])] + [str(v) for v in ties[r][:nties[r]]])
                assert f == ["%g" % best[r], str(offsets[r + 1] - offsets[r]), plist, chr(strand[r])], (name, ln)
        else:
            s = U.oracle_lpm_sum(codes, offsets, m5, bg, fwd)
            for r, ln in enumerate(lines):
                assert ln.split("\t")[-1] == "%g" % s[r], (name, ln)
