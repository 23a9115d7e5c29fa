#!/usr/bin/env python
"""Generates tests/golden/*.npz / *.json by running the REAL reference (read-only checkout at
/root/reference) on CPU in the build container.  The reference cannot travel to the GPU box, so
the vectors are committed; re-run this script only here:

    python tests/golden/make_golden.py

Missing third-party modules of the reference (Bio, click_option_group, logomaker, matplotlib) are
stubbed with the minimum the imported code paths touch.
"""
import importlib.util
import io
import json
import os
import subprocess
import sys
import types

import numpy as np
import torch

REF = os.environ.get("EXPLAINN_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))


# ---- stubs -----------------------------------------------------------------------------------------
def _install_stubs():
    class Seq(str):
        def reverse_complement(self):
            return Seq(self.translate(str.maketrans("ACGTacgtNn", "TGCAtgcaNn"))[::-1])

    class SeqRecord:
        def __init__(self, seq, id="", name="", description=""):
            self.seq, self.id = seq, id

        def format(self, fmt):
            return f">{self.id}\n{self.seq}\n"

    for name in ["Bio", "Bio.motifs", "Bio.Seq", "Bio.SeqRecord", "Bio.SeqIO", "click_option_group", "logomaker",
                 "matplotlib", "matplotlib.pyplot", "matplotlib.font_manager"]:
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["Bio"].motifs = sys.modules["Bio.motifs"]
    sys.modules["Bio"].SeqIO = sys.modules["Bio.SeqIO"]
    sys.modules["Bio.Seq"].Seq = Seq
    sys.modules["Bio.SeqRecord"].SeqRecord = SeqRecord
    import click

    class _OptGroup:
        @staticmethod
        def group(*a, **k):
            return lambda f: f

        option = staticmethod(click.option)

    cog = sys.modules["click_option_group"]
    cog.optgroup = _OptGroup
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["matplotlib"].font_manager = sys.modules["matplotlib.font_manager"]
    sys.modules["matplotlib"].use = lambda *a, **k: None
    sys.modules["matplotlib.font_manager"].FontProperties = lambda *a, **k: None


def _load(name, relpath):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, "explainn", relpath))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def randomise_bn(model, seed):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for m in model.modules():
            if isinstance(m, torch.nn.BatchNorm1d):
                n = m.num_features
                m.running_mean.copy_(torch.randn(n, generator=g) * 0.1)
                m.running_var.copy_(torch.rand(n, generator=g) + 0.5)
                m.weight.copy_(torch.rand(n, generator=g) + 0.5)
                flip = torch.rand(n, generator=g) < 0.1
                m.weight[flip] *= -1
                m.bias.copy_(torch.randn(n, generator=g) * 0.1)


def random_onehot(B, L, seed, n_frac=0.0):
    g = torch.Generator().manual_seed(seed)
    codes = torch.randint(0, 4, (B, L), generator=g)
    x = torch.zeros(B, 4, L)
    x.scatter_(1, codes[:, None, :], 1.0)
    if n_frac > 0:
        mask = torch.rand(B, L, generator=g) < n_frac
        x = x * (~mask)[:, None, :].float()
    return x


def sd_np(model):
    return {k: v.detach().numpy().copy() for k, v in model.state_dict().items()}


def main():
    _install_stubs()
    sys.path.insert(0, os.path.join(REF, "explainn"))
    from architectures import ExplaiNN, PWM  # the reference's own modules
    import sequence as ref_sequence
    ref_test = _load("ref_test_py", "test.py")
    sys.modules["test"] = ref_test  # interpret.py does `from test import ...` (collides with the stdlib package)
    ref_interpret = _load("ref_interpret_py", "interpret.py")
    ref_pwm_scoring = _load("ref_pwm_scoring_py", "pwm-scoring.py")

    # ---- 1. eval forward + activations (a1-a5, a7) --------------------------------------------------
    for tag, (U, k, L, Fo, B, nfrac) in {"eval_small": (12, 19, 60, 3, 8, 0.05),
                                         "eval_cfg1": (100, 19, 200, 1, 4, 0.01),
                                         "eval_k11": (7, 11, 45, 2, 5, 0.0)}.items():
        torch.manual_seed(0)
        model = ExplaiNN(U, k, L, Fo)
        randomise_bn(model, 7)
        model.eval()
        x = random_onehot(B, L, 1234, nfrac)
        with torch.no_grad():
            xr = x.repeat(1, U, 1)
            outs = model.linears(xr)
            logits = model.final(outs)
            acts = model.linears[:3](xr)
            assert torch.equal(logits, model(x))
        out = {"sd." + k_: v for k_, v in sd_np(model).items()}
        out.update(x=x.numpy(), logits=logits.numpy(), outs=outs.numpy(), acts=acts.numpy(),
                   acts16=acts.numpy().astype(np.float16), opts=np.array([U, k, L, Fo, B]))
        np.savez_compressed(os.path.join(HERE, f"{tag}.npz"), **out)

    # ---- 2. training step (a6): loss, every gradient, running stats, Adam --------------------------
    for tag, (U, k, L, Fo, B, kind) in {"train_mse": (6, 19, 67, 2, 16, "linear"),
                                        "train_bce": (5, 19, 60, 1, 12, "binary")}.items():
        torch.manual_seed(0)
        model = ExplaiNN(U, k, L, Fo)
        randomise_bn(model, 11)
        model.train()
        model.linears[9].p = 0.0  # torch's CPU dropout stream cannot be reproduced on the device
        x = random_onehot(B, L, 4321, 0.02)
        g = torch.Generator().manual_seed(99)
        y = torch.randn(B, Fo, generator=g) if kind == "linear" else (torch.rand(B, Fo, generator=g) < 0.5).float()
        before = sd_np(model)
        import architectures as ref_arch
        crit = ref_arch.get_loss(kind)
        opt = ref_arch.get_optimizer(model.parameters(), 0.003)
        logits = model(x)
        loss = crit(logits, y)
        opt.zero_grad()
        loss.backward()
        grads = {k_: p.grad.detach().numpy().copy() for k_, p in model.named_parameters()}
        opt.step()
        after = sd_np(model)
        out = {"sd." + k_: v for k_, v in before.items()}
        out.update({"grad." + k_: v for k_, v in grads.items()})
        out.update({"after." + k_: v for k_, v in after.items()})
        out.update(x=x.numpy(), y=y.numpy(), loss=loss.detach().numpy(), logits=logits.detach().numpy(),
                   opts=np.array([U, k, L, Fo, B]))
        np.savez_compressed(os.path.join(HERE, f"{tag}.npz"), **out)

    # ---- 2b. the headline workload at its own shape (cfg3: U=300, L=600, F=50), B=32, p=0 -------------------
    # The state dict is 10.6 MB: instead of three copies of it the fixture keeps what regenerates it (seed 0
    # construction + randomise_bn(11), checked through per-key fp64 checksums), x as base codes, and for the 10 MB
    # tensor linears.6.weight the gradient / updated rows of SAMPLED units plus per-unit norms of all of them.
    U, k, L, Fo, B = 300, 19, 600, 50, 32
    torch.manual_seed(0)
    model = ExplaiNN(U, k, L, Fo)
    randomise_bn(model, 11)
    model.train()
    model.linears[9].p = 0.0
    x = random_onehot(B, L, 4321, 0.01)
    y = torch.randn(B, Fo, generator=torch.Generator().manual_seed(99))
    before = sd_np(model)
    import architectures as ref_arch
    crit, opt = ref_arch.get_loss("linear"), ref_arch.get_optimizer(model.parameters(), 0.003)
    logits = model(x)
    loss = crit(logits, y)
    opt.zero_grad()
    loss.backward()
    grads = {k_: p.grad.detach().numpy().copy() for k_, p in model.named_parameters()}
    opt.step()
    after = sd_np(model)
    units = np.array([0, 1, 57, 127, 128, 255, 256, 299])
    codes = np.where(x.numpy().sum(1) > 0, x.numpy().argmax(1), 4).astype(np.uint8)
    out = {"opts": np.array([U, k, L, Fo, B]), "codes": codes, "y": y.numpy(), "loss": loss.detach().numpy(),
           "logits": logits.detach().numpy(), "units": units,
           "sd_keys": np.array(list(before.keys())),
           "sd_checksum": np.array([np.asarray(v, dtype=np.float64).sum() for v in before.values()])}
    big = "linears.6.weight"
    n_ = before[big].shape[1]
    for tag_, dct in (("grad.", grads), ("after.", after)):
        for k_, v in dct.items():
            if k_ == big:
                v3 = v.reshape(U, 100, n_)
                out[tag_ + k_ + ".units"] = v3[units]
                out[tag_ + k_ + ".unit_norms"] = np.sqrt((v3.astype(np.float64) ** 2).sum(axis=(1, 2)))
            else:
                out[tag_ + k_] = v
    np.savez_compressed(os.path.join(HERE, "train_cfg3.npz"), **out)

    # ---- 3. interpret: thresholds + sites from the reference's own functions (a8, a9) -------------
    d = np.load(os.path.join(HERE, "eval_small.npz"))
    U, k, L, Fo, B = d["opts"]
    torch.manual_seed(0)
    model = ExplaiNN(int(U), int(k), int(L), int(Fo))
    model.load_state_dict({k_[3:]: torch.tensor(d[k_]) for k_ in d.files if k_.startswith("sd.")})
    model.eval()
    x = torch.tensor(d["x"])
    xx = torch.cat((x, torch.tensor(ref_sequence.rc_one_hot_encoding_many(x.numpy()).copy())))
    loader = torch.utils.data.DataLoader(torch.utils.data.TensorDataset(xx, torch.zeros(len(xx), 1)), 5)
    # force the CPU path of the reference helper regardless of the machine
    acts, outs, preds = ref_test._get_acts_outs_preds(model, loader)
    letters = np.array(list("ACGT"))
    seqs_str = []
    for s in xx.numpy():
        seqs_str.append("".join(letters[s.argmax(0)[i]] if s[:, i].sum() else "N" for i in range(s.shape[1])))
    ids = [f"seq{i}" for i in range(len(xx))]
    inter = {"acts16": acts, "outs16": outs, "preds16": preds, "x2": xx.numpy()}
    for rc_flag in (False, True):
        idxs = np.array([0, 2, 3, 5, 7]) if rc_flag else np.array([1, 2, 4, 5, 6, 7, 9, 12, 15])
        thr = ref_interpret._get_act_thresholds(acts, idxs, rc_flag)
        sites_all = []
        for u in range(int(U)):
            fh = io.StringIO()
            ref_interpret._get_sites(fh, idxs, ids, seqs_str, acts[:, u, :], thr[u], int(k), rc_flag)
            sites_all.append(fh.getvalue())
        sfx = "_rc" if rc_flag else ""
        inter["idxs" + sfx] = idxs
        inter["thr" + sfx] = thr
        inter["sites_fasta" + sfx] = np.array(sites_all)
    np.savez_compressed(os.path.join(HERE, "interpret_small.npz"), **inter)

    # ---- 4. PWM module + MEME parsing / resizing (a11, a12) ----------------------------------------
    rng = np.random.default_rng(5)
    meme = ["MEME version 4\n\nALPHABET= ACGT\n\n"]
    for i, w in enumerate([6, 9, 12, 9]):
        cols = rng.dirichlet([0.5] * 4, size=w)
        cols[0, 0] = 0.0  # exercises the 1e-4 floor
        meme.append(f"MOTIF M{i} name{i}\nletter-probability matrix: alength= 4 w= {w} nsites= 20 E= 0\n")
        meme += ["  %.6f  %.6f  %.6f  %.6f\n" % tuple(c) for c in cols]
        meme.append("\n")
    meme_text = "".join(meme)
    meme_path = os.path.join(HERE, "small.meme")
    with open(meme_path, "w") as fh:
        fh.write(meme_text)
    pwms_log, names = ref_pwm_scoring._get_PWMs(meme_path, resize_pwms=True, return_log=True)
    pwms_raw, _ = ref_pwm_scoring._get_PWMs(meme_path, resize_pwms=False, return_log=False)
    x = random_onehot(6, 40, 77, 0.03)
    res = {"pwms_log": np.asarray(pwms_log), "names": np.array(names), "x": x.numpy(),
           "raw_widths": np.array([len(p[0]) for p in pwms_raw])}
    for scoring in ("max", "sum"):
        m = PWM(np.asarray(pwms_log), 40, scoring)
        with torch.no_grad():
            res[scoring] = m(x).numpy()
    np.savez_compressed(os.path.join(HERE, "pwm_small.npz"), **res)

    # ---- 5. the C tool (a13-a15): outputs of oracle/_ref/pwm_scoring ---------------------------------
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=True, capture_output=True)
    tool = os.path.join(ROOT, "oracle", "_ref", "pwm_scoring")
    fasta = open(os.path.join(REF, "explainn", "test", "positives_short.fa")).read()
    # add records with N, lower case, a record shorter than the matrix and a homopolymer (ties)
    fasta += ">withN some description\nACGTNNACGTAGGCGCGCTTANACGATCGATTTTTTTTACGANN\n>lower\nacgtagctagctagctagcatcgatgcatgcatgcatgcta\n" \
             ">short\nACGTACG\n>ties\n" + "A" * 50 + "\n>multi\nACGTAGCTAGCATCGATCGATGC\nGGATCGATCGATTAGCTAGCTA\n"
    damo = open(os.path.join(REF, "explainn", "test", "filter4.damo")).read().splitlines()
    rows = np.array([[float(v) for v in ln.split("|")[1].split()] for ln in damo[1:5]])  # (4,19)
    lpm_text = ">letter-probability matrix filter4\n" + "".join(
        "%10.3f\t%10.3f\t%10.3f\t%10.3f\n" % tuple(rows[:, j]) for j in range(rows.shape[1]))
    ipwm = np.round(100 * np.log2(np.maximum(rows, 1e-3) / 0.25)).astype(int)
    pwm_text = "# integer PWM derived from filter4\n" + "".join(
        "%6d %6d %6d %6d\n" % tuple(ipwm[:, j]) for j in range(ipwm.shape[1]))
    tmpdir = os.path.join(ROOT, "scratch")
    os.makedirs(tmpdir, exist_ok=True)
    fa_p, lpm_p, pwm_p = (os.path.join(tmpdir, n) for n in ("g.fa", "g.lpm", "g.pwm"))
    open(fa_p, "w").write(fasta)
    open(lpm_p, "w").write(lpm_text)
    open(pwm_p, "w").write(pwm_text)
    runs = {}
    for name, opts in {"sum": [], "sum_f": ["-f"], "sum_u": ["-u"], "sum_w": ["-w", "0.0001"], "sum_q": ["-q"],
                       "sum_qf": ["-q", "-f"], "sum_p": ["-p", "0.29,0.21,0.21,0.29"], "best": ["-b"],
                       "best_f": ["-b", "-f"], "best_u_r": ["-b", "-u", "-r"], "pwm": ["--pwm"],
                       "pwm_f": ["--pwm", "-f"]}.items():
        mat = pwm_p if "--pwm" in opts else lpm_p
        r = subprocess.run([tool] + opts + ["-m", mat, fa_p], capture_output=True, text=True, check=True)
        runs[name] = {"opts": opts, "stdout": r.stdout}
    json.dump({"fasta": fasta, "lpm": lpm_text, "pwm": pwm_text, "runs": runs},
              open(os.path.join(HERE, "pwmscan_golden.json"), "w"), indent=1)

    # ---- 5b. interpret post-processing (f3): the reference's own selection / importance functions ----------
    rngp = np.random.default_rng(11)
    post = {}
    for tag_, (N_, F_) in {"bin": (40, 3), "lin": (60, 2)}.items():
        preds16 = (rngp.standard_normal((N_, F_)) * 2).astype(np.float16)
        labels = (rngp.random((N_, F_)) < 0.5).astype(np.float64) if tag_ == "bin" else rngp.standard_normal((N_, F_))
        if tag_ == "bin":                                  # make a good share of the rows well predicted
            labels[: N_ // 2] = (preds16[: N_ // 2] > 0).astype(np.float64)
        post[f"{tag_}.preds"], post[f"{tag_}.labels"] = preds16, labels
        for rc_ in (False, True):
            post[f"{tag_}.idxs.rc{int(rc_)}"] = np.asarray(ref_interpret._get_well_predicted_sequences(
                preds16, labels, "binary" if tag_ == "bin" else "linear", rc_), dtype=np.int64)
    acts_u = (rngp.random((40, 30)) * 3).astype(np.float16)
    imps_u = rngp.standard_normal((3, 40))
    post["imp.acts"], post["imp.imps"], post["imp.idxs"] = acts_u, imps_u, np.arange(0, 40, 3)
    post["imp.thr"] = np.float16(2.7)
    post["imp.out"] = ref_interpret._filter_filter_importances(imps_u, post["imp.idxs"], acts_u, post["imp.thr"])
    np.savez_compressed(os.path.join(HERE, "postproc.npz"), **post)

    # ---- 6. sites -> JASPAR fixtures of the reference (a10) -------------------------------------------
    cam = os.path.join(REF, "results", "Human-Mouse-Project", "CAM")
    fx = {}
    for i in (23, 39, 57):
        sp, mp = os.path.join(cam, "sites", f"filter{i}.fa"), os.path.join(cam, "motifs", f"filter{i}.jaspar")
        if os.path.exists(sp) and os.path.exists(mp) and os.path.getsize(sp) < 200000:
            fx[f"filter{i}"] = {"sites": open(sp).read(), "jaspar": open(mp).read()}
    json.dump(fx, open(os.path.join(HERE, "sites_jaspar_fixtures.json"), "w"))
    # all 126 pairs, compact: the emitted k-mers of every sites file as ASCII bytes (N, k) + the JASPAR text
    allfx, ids_ = {}, []
    for i in range(1000):
        sp, mp = os.path.join(cam, "sites", f"filter{i}.fa"), os.path.join(cam, "motifs", f"filter{i}.jaspar")
        if not (os.path.exists(sp) and os.path.exists(mp)):
            continue
        kmers = [ln.strip() for ln in open(sp) if ln.strip() and ln[0] != ">"]
        w = len(kmers[0]) if kmers else 0
        assert all(len(s_) == w for s_ in kmers)
        allfx[f"kmers{i}"] = np.frombuffer("".join(kmers).encode(), dtype=np.uint8).reshape(len(kmers), w)
        allfx[f"jaspar{i}"] = np.array(open(mp).read())
        ids_.append(i)
    allfx["ids"] = np.array(ids_)
    np.savez_compressed(os.path.join(HERE, "sites_jaspar_all.npz"), **allfx)
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
