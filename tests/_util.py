"""Shared helpers of the test-suite (oracle loading, golden parsing)."""
import ctypes
import json
import os
import subprocess

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def load_npz(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def golden_sd(d, prefix="sd."):
    return {k[len(prefix):]: torch.tensor(d[k]) for k in d.files if k.startswith(prefix)}


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def assert_close(a, b, rtol=1e-5, atol_scale=1e-5, what=""):
    """|a-b| <= rtol*|b| + atol_scale*max|b| elementwise: a relative tolerance plus an absolute floor scaled to the
    magnitude of the tensor (a relative error on entries near zero of a SIGNED tensor -- logits, gradients -- is
    ill-posed: SURVEY.md section 7).  Strictly positive tensors (activations) go through `assert_rel` instead."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, f"{what}: shape {a.shape} vs {b.shape}"
    scale = max(np.abs(b).max(), 1e-30) if b.size else 1.0
    err = np.abs(a - b)
    bad = err > rtol * np.abs(b) + atol_scale * scale
    if bad.any():
        i = np.unravel_index(np.argmax(err - rtol * np.abs(b)), err.shape) if err.ndim else ()
        raise AssertionError(f"{what}: {bad.sum()} of {bad.size} outside tolerance; max rel-to-scale "
                             f"{err.max() / scale:.3e}; worst element {i}: got {a[i]!r} want {b[i]!r}")


def assert_rel(a, b, rtol=1e-5, floor=1e-7, what=""):
    """Elementwise RELATIVE check for strictly positive tensors (unit activations = exp(...), pooled activations):
    |a-b| <= rtol * max(|b|, floor/rtol * max|b|), i.e. every element is held to `rtol` relative and only elements
    more than a factor rtol/floor (100x at the defaults) below the tensor's maximum get the absolute floor
    `floor * max|b|`.  BASELINE.json north_star: "unit activations within 1e-5 relative in fp32".
    Returns the worst elementwise relative error (printed by the callers' -s runs)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, f"{what}: shape {a.shape} vs {b.shape}"
    scale = max(np.abs(b).max(), 1e-300)
    denom = np.maximum(np.abs(b), floor / rtol * scale)
    rel = np.abs(a - b) / denom
    worst = float(rel.max()) if rel.size else 0.0
    if worst > rtol:
        i = np.unravel_index(np.argmax(rel), rel.shape)
        raise AssertionError(f"{what}: {(rel > rtol).sum()} of {rel.size} elements beyond {rtol:g} relative; worst "
                             f"{worst:.3e} at {i}: got {a[i]!r} want {b[i]!r} (tensor max {scale:.3e})")
    print(f"[assert_rel] {what}: worst elementwise relative error {worst:.3e} over {rel.size} elements")
    return worst


def random_model(Uu, k, L, F, seed=0, bn_seed=7):
    """explainn_b200 ExplaiNN with the reference's default initialisation under `seed` and randomised BatchNorm
    buffers / affine (about 10 % negative scales), the construction tests/golden/make_golden.py uses
    (`randomise_bn`): the same seeds regenerate the same state dict."""
    from explainn_b200.architectures import ExplaiNN
    torch.manual_seed(seed)
    m = ExplaiNN(Uu, k, L, F)
    g = torch.Generator().manual_seed(bn_seed)
    with torch.no_grad():
        for mod in m.modules():
            if isinstance(mod, torch.nn.BatchNorm1d):
                n = mod.num_features
                mod.running_mean.copy_(torch.randn(n, generator=g) * 0.1)
                mod.running_var.copy_(torch.rand(n, generator=g) + 0.5)
                mod.weight.copy_(torch.rand(n, generator=g) + 0.5)
                mod.weight[torch.rand(n, generator=g) < 0.1] *= -1
                mod.bias.copy_(torch.randn(n, generator=g) * 0.1)
    return m


def cfg3_golden():
    """tests/golden/train_cfg3.npz: (fixture, model in train mode with p = 0, x float one-hot, y).  The fixture
    stores what regenerates the 10.6 MB state dict instead of the state dict; the per-key checksums prove it."""
    d = load_npz("train_cfg3.npz")
    Uu, k, L, F, B = [int(v) for v in d["opts"]]
    m = random_model(Uu, k, L, F, seed=0, bn_seed=11)
    sd = m.state_dict()
    for key, chk in zip(d["sd_keys"], d["sd_checksum"]):
        got = float(sd[str(key)].double().sum())
        assert abs(got - float(chk)) <= 1e-9 * max(1.0, abs(float(chk))), f"regenerated {key} differs from the fixture"
    codes = torch.tensor(d["codes"].astype(np.int64))
    x = torch.zeros(B, 5, L)
    x.scatter_(1, codes[:, None, :], 1.0)
    m.train()
    m.linears[9].p = 0.0
    return d, m, x[:, :4].contiguous(), torch.tensor(d["y"])


ZERO_CLASS = {"linears.0.bias", "linears.1.bias", "linears.6.bias", "linears.10.bias"}


def check_cfg3_grads(d, grads, what, rtol=2e-4, atol_scale=2e-5, zero_tol=2e-6):
    """named gradients against the compact cfg3 fixture (linears.6.weight: sampled units + per-unit norms).
    Tolerance: 2e-4 relative + 2e-5 of the tensor's largest entry; 1e-4 for the two gradients downstream of dy (conv
    weight, first BatchNorm scale: sums of thousands of cancelling terms), the stated tolerance of the two-piece bf16
    backward (tests/test_gpu_parity.py FRONT_GRADS)."""
    Uu = int(d["opts"][0])
    for key, got in grads.items():
        got = np.asarray(got)
        if key == "linears.6.weight":
            g3 = got.reshape(Uu, 100, -1)
            assert_close(g3[d["units"]], d["grad." + key + ".units"], rtol=rtol, atol_scale=atol_scale,
                         what=f"{what} {key} (sampled units)")
            norms = np.sqrt((g3.astype(np.float64) ** 2).sum(axis=(1, 2)))
            assert_close(norms, d["grad." + key + ".unit_norms"], rtol=rtol, atol_scale=atol_scale,
                         what=f"{what} {key} (per-unit norms, all units)")
        elif key in ZERO_CLASS:
            assert np.abs(got - d["grad." + key].reshape(got.shape)).max() < zero_tol, key
        else:
            front = key in ("linears.0.weight", "linears.1.weight")
            assert_close(got, d["grad." + key].reshape(got.shape), rtol=rtol, atol_scale=1e-4 if front else atol_scale,
                         what=f"{what} {key}")


def check_cfg3_after(d, params, rtol=1e-4, atol_scale=1e-4):
    """parameters after one Adam step against the fixture.  Adam's first step is -lr * g / (|g| + eps): an element
    whose gradient is round-off noise moves by +-lr with the sign of the noise, and one whose gradient is not far
    above eps = 1e-8 moves by an amount that depends on its magnitude: only elements whose reference gradient is
    above 1e-3 of the tensor's largest AND above 1e-6 are compared (and none of the zero-class tensors)."""
    Uu = int(d["opts"][0])
    for key, got in params.items():
        if key in ZERO_CLASS:
            continue
        got = np.asarray(got)
        if key == "linears.6.weight":
            got = got.reshape(Uu, 100, -1)[d["units"]]
            ref, g = d["after." + key + ".units"], d["grad." + key + ".units"]
        else:
            ref, g = d["after." + key].reshape(got.shape), d["grad." + key].reshape(got.shape)
        keep = (np.abs(g) > 1e-3 * np.abs(g).max()) & (np.abs(g) > 1e-6)
        assert keep.mean() > 0.2, key
        assert_close(got[keep], ref[keep], rtol=rtol, atol_scale=atol_scale, what=f"after {key}")


_oracle_lib = None


def oracle_c():
    """ctypes handle of oracle/liboracle.so (built on demand with the committed Makefile)."""
    global _oracle_lib
    if _oracle_lib is None:
        path = os.path.join(ROOT, "oracle", "liboracle.so")
        if not os.path.exists(path):
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"], check=True,
                           capture_output=True)
        lib = ctypes.CDLL(path)
        lib.oracle_lpm_sum.restype = ctypes.c_double
        lib.oracle_lpm_sum.argtypes = [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_void_p,
                                       ctypes.c_void_p, ctypes.c_int]
        lib.oracle_lpm_best.restype = ctypes.c_int
        lib.oracle_lpm_best.argtypes = [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_void_p,
                                        ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        lib.oracle_ipwm_best.restype = None
        lib.oracle_ipwm_best.argtypes = [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_void_p,
                                         ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        _oracle_lib = lib
    return _oracle_lib


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def oracle_lpm_sum(codes, offsets, lpm5, bg, forward_only):
    lib = oracle_c()
    lpm5 = np.ascontiguousarray(lpm5, dtype=np.float64)
    w = lpm5.shape[1]
    out = np.zeros(len(offsets) - 1)
    for r in range(len(offsets) - 1):
        seg = np.ascontiguousarray(codes[offsets[r]:offsets[r + 1]], dtype=np.int8)
        b = np.ascontiguousarray(bg[r] if np.ndim(bg) == 2 else bg, dtype=np.float64)
        out[r] = lib.oracle_lpm_sum(_ptr(seg), len(seg), w, _ptr(lpm5), _ptr(b), int(forward_only))
    return out


def oracle_lpm_best(codes, offsets, lpm5, bg, forward_only, tie_cap=16):
    lib = oracle_c()
    lpm5 = np.ascontiguousarray(lpm5, dtype=np.float64)
    w = lpm5.shape[1]
    R = len(offsets) - 1
    best, pos, strand, nties = np.zeros(R), np.zeros(R, np.int32), np.zeros(R, np.int8), np.zeros(R, np.int32)
    ties = np.full((R, tie_cap), -1, np.int32)
    for r in range(R):
        seg = np.ascontiguousarray(codes[offsets[r]:offsets[r + 1]], dtype=np.int8)
        b = np.ascontiguousarray(bg[r] if np.ndim(bg) == 2 else bg, dtype=np.float64)
        bo, po, so = ctypes.c_double(), ctypes.c_int(), ctypes.c_char()
        t = np.full(tie_cap, -1, np.int32)
        nties[r] = lib.oracle_lpm_best(_ptr(seg), len(seg), w, _ptr(lpm5), _ptr(b), int(forward_only),
                                       ctypes.byref(bo), ctypes.byref(po), ctypes.byref(so), _ptr(t), tie_cap)
        best[r], pos[r], strand[r], ties[r] = bo.value, po.value, ord(so.value), t
    return best, pos, strand, nties, ties


def oracle_ipwm_best(codes, offsets, pwm5, forward_only):
    lib = oracle_c()
    pwm5 = np.ascontiguousarray(pwm5, dtype=np.int32)
    w = pwm5.shape[1]
    R = len(offsets) - 1
    best, pos, strand = np.zeros(R, np.int32), np.zeros(R, np.int32), np.zeros(R, np.int8)
    for r in range(R):
        seg = np.ascontiguousarray(codes[offsets[r]:offsets[r + 1]], dtype=np.int8)
        bo, po, so = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int8()
        lib.oracle_ipwm_best(_ptr(seg), len(seg), w, _ptr(pwm5), int(forward_only), ctypes.byref(bo),
                             ctypes.byref(po), ctypes.byref(so))
        best[r], pos[r], strand[r] = bo.value, po.value, so.value
    return best, pos, strand


def pwmscan_golden():
    return json.load(open(os.path.join(GOLDEN, "pwmscan_golden.json")))


def parse_matrix_text(text, integer=False):
    rows = [ln.split()[:4] for ln in text.splitlines() if ln.strip() and ln[0] not in "#>"]
    return np.array(rows, dtype=np.int64 if integer else np.float64).T.copy()


def random_codes(rng, lengths, n_frac=0.01):
    chunks, offsets = [], [0]
    for ln in lengths:
        c = rng.integers(0, 4, size=ln).astype(np.int8)
        c[rng.random(ln) < n_frac] = 4
        chunks.append(c)
        offsets.append(offsets[-1] + ln)
    return np.concatenate(chunks) if chunks else np.zeros(0, np.int8), np.array(offsets, np.int64)
