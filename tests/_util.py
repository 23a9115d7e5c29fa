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
    """|a-b| <= rtol*|b| + atol_scale*rtol*max|b|  -- relative tolerance with an absolute floor scaled to
    the magnitude of the tensor (relative error on entries near zero is ill-posed: SURVEY.md section 7)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, f"{what}: shape {a.shape} vs {b.shape}"
    scale = max(np.abs(b).max(), 1e-30)
    bad = np.abs(a - b) > rtol * np.abs(b) + atol_scale * scale
    assert not bad.any(), f"{what}: {bad.sum()} of {bad.size} outside tolerance; max rel-to-scale " \
                          f"{np.abs(a - b).max() / scale:.3e}"


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
