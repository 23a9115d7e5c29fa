"""Torch-tensor wrappers over the C-ABI entry points (include/explainn_b200.h).

PyTorch is plumbing here: it owns device memory and the current stream; every
computation is a hand-written sm_100a kernel in explainn_b200/csrc.  All ops raise
on non-CUDA tensors -- there is no CPU fallback.
"""
import ctypes

import torch

from . import _lib

HIDDEN = 100  # explainn/architectures/__init__.py:86
POOL = 7      # explainn/architectures/__init__.py:81
KERNEL_SIZE_MAX = 48        # xnn_kernel_size_limit(0): eval forward, interpret, PWM scans (tcgen05 Toeplitz kernels)
KERNEL_SIZE_MAX_TRAIN = 32  # xnn_kernel_size_limit(1): training and linears[:3] activations

_launches = 0  # C-ABI calls issued (bench.py reports kernel launches from the per-call table below)

# kernels launched by each entry point (for bench.py's gpu_launches claim)
KERNELS_PER_CALL = {
    "xnn_pack_onehot_f32": 1, "xnn_pack_ascii": 1, "xnn_unpack_onehot_f32": 1, "xnn_revcomp_planes": 1,
    "xnn_cooc_counts": 1, "xnn_bn1_train_prepare": 1, "xnn_bn_fold_eval": 1, "xnn_frontend_fwd": 2, "xnn_frontend_fwd_lut": 1,
    "xnn_frontend_acts": 1, "xnn_fc1_fwd": 1, "xnn_bn_train_prepare": 2, "xnn_fc2_fwd": 1, "xnn_head_fwd": 2,
    "xnn_linear_fwd": 1, "xnn_linear_bwd": 5, "xnn_bn_bwd": 3, "xnn_fc2_bwd": 2, "xnn_fc1_bwd": 5,
    "xnn_frontend_bwd": 2, "xnn_frontend_bwd_simt": 2, "xnn_adam_step": 1, "xnn_act_max": 1, "xnn_site_counts": 1, "xnn_site_emit": 1,
    "xnn_pwm_scan": 1, "xnn_pwm_scan_tc": 2, "xnn_act_max_tc": 2, "xnn_site_counts_tc": 2, "xnn_site_emit_tc": 2, "xnn_mlp_bn2_train_prepare": 2, "xnn_mlp_fwd": 1, "xnn_mlp_bwd": 4, "xnn_lpm_scan_sum": 1, "xnn_lpm_scan_sum_parts": 2, "xnn_lpm_scan_best_parts": 5, "xnn_ipwm_scan_best_parts": 2, "xnn_lpm_scan_best": 1, "xnn_ipwm_scan_best": 1,
    "xnn_adam_step_multi": 1, "xnn_counters_advance": 1, "xnn_counters_advance_list": 1, "xnn_bn_running_update": 1, "xnn_loss_fwd": 1, "xnn_loss_bwd": 1, "xnn_sort_columns": 1, "xnn_metric_auc": 2, "xnn_metric_ranks": 1, "xnn_metric_pearson_f32": 1, "xnn_metric_pearson_f64": 1,
}
kernel_launches = 0


_seen_devices = []      # devices of the tensors handed to the C call being assembled (filled by _p)
_STREAM = object()      # placeholder for "the current stream of those tensors' device", resolved in _call


def _call(name, *args):
    """One C-ABI call.  Every tensor of the call must live on ONE device; the call runs with that device
    current (the library sizes grids from the current device) on that device's current torch stream."""
    global _launches, kernel_launches
    lib = _lib.load()
    devs = set(_seen_devices)
    _seen_devices.clear()
    if len(devs) > 1:
        raise RuntimeError(f"{name}: tensors on different devices {sorted(map(str, devs))}")
    dev = devs.pop() if devs else torch.device("cuda", torch.cuda.current_device())
    _launches += 1
    kernel_launches += KERNELS_PER_CALL.get(name, 1)
    if dev.index is not None and dev.index != torch.cuda.current_device():
        with torch.cuda.device(dev):
            return _call_on(lib, name, dev, args)
    return _call_on(lib, name, dev, args)


def _call_on(lib, name, dev, args):
    st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(getattr(lib, name)(*[st if a is _STREAM else a for a in args]), name)


def _p(t):
    if t is None:
        return None
    _seen_devices.append(t.device)
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return _STREAM


def _req(t, dtype, name, contiguous=True):
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name}: expected a torch.Tensor")
    if not t.is_cuda:
        raise RuntimeError(f"{name}: explainn_b200 kernels run on CUDA tensors only (got {t.device}); "
                           "there is no CPU fallback")
    if t.dtype != dtype:
        raise TypeError(f"{name}: expected dtype {dtype}, got {t.dtype}")
    if contiguous and not t.is_contiguous():
        raise ValueError(f"{name}: expected a contiguous tensor")
    return t


def planes_words(L):
    return (L + 31) // 32


def kernel_size_limit(path):
    return int(_lib.load().xnn_kernel_size_limit(int(path)))


def set_sm_reserve(n):
    """leave n SMs out of the persistent grids (for a concurrent NCCL all-reduce); 0 = use every SM"""
    _lib.check(_lib.load().xnn_set_sm_reserve(int(n)), "xnn_set_sm_reserve")


def sm_count():
    out = ctypes.c_int(0)
    _lib.check(_lib.load().xnn_sm_count(ctypes.byref(out)), "xnn_sm_count")
    return out.value


# ---- a16: encoding -------------------------------------------------------------
def pack_onehot(x, check=False, bad=None, out=None):
    """float32 one-hot (B,C>=4,L) (channels 0..3 are read) -> planes int32 (B,3,ceil(L/32)).

    Mirrors the input contract of explainn/sequence/__init__.py:8-29 (all-zero column = N).  A column that is
    neither one-hot nor all-zero (soft encodings, 0.25 for N, perturbed inputs) cannot be expressed in the packed
    layout: `check=True` raises on it (one device->host read); `bad` (int32 (1,) device tensor) accumulates the
    number of such columns without a sync, for the caller to read later (UnitBank.check_inputs)."""
    _req(x, torch.float32, "x", contiguous=False)
    if x.dim() != 3 or x.shape[1] < 4:
        raise ValueError("x must be (B, 4*m, L)")
    if x.stride(2) != 1:
        x = x.contiguous()
    B, _, L = x.shape
    planes = _planes_out(out, B, L, x.device)
    if bad is None and check:
        bad = torch.zeros(1, dtype=torch.int32, device=x.device)
    _call("xnn_pack_onehot_f32", _p(x), x.stride(0), x.stride(1), B, L, _p(planes), _p(bad), _stream())
    if check and int(bad.item()) != 0:
        raise ValueError("input is not one-hot: it has columns that are neither one-hot nor all-zero "
                         "(explainn_b200 computes on the packed 2-bit layout; the reference's Conv1d accepts "
                         "arbitrary floats)")
    return planes


def _planes_out(out, B, L, dev):
    if out is None:
        return torch.empty((B, 3, planes_words(L)), dtype=torch.int32, device=dev)
    _req(out, torch.int32, "out")
    if tuple(out.shape) != (B, 3, planes_words(L)):
        raise ValueError(f"out: expected planes of shape {(B, 3, planes_words(L))}, got {tuple(out.shape)}")
    return out


def pack_ascii(seqs, out=None):
    """uint8 ASCII (B,L) -> planes."""
    _req(seqs, torch.uint8, "seqs")
    B, L = seqs.shape
    planes = _planes_out(out, B, L, seqs.device)
    _call("xnn_pack_ascii", _p(seqs), seqs.stride(0), B, L, _p(planes), _stream())
    return planes


def unpack_onehot(planes, L):
    _req(planes, torch.int32, "planes")
    B = planes.shape[0]
    x = torch.empty((B, 4, L), dtype=torch.float32, device=planes.device)
    _call("xnn_unpack_onehot_f32", _p(planes), B, L, _p(x), _stream())
    return x


def revcomp_planes(planes, L):
    _req(planes, torch.int32, "planes")
    out = torch.empty_like(planes)
    _call("xnn_revcomp_planes", _p(planes), planes.shape[0], L, _p(out), _stream())
    return out


# ---- a3: front end -----------------------------------------------------------------
def cooc_counts(planes, L, k):
    _req(planes, torch.int32, "planes")
    cnt1 = torch.empty(4 * k, dtype=torch.int64, device=planes.device)      # cleared by the call itself
    cnt2 = torch.empty((4 * k, 4 * k), dtype=torch.int64, device=planes.device)
    _call("xnn_cooc_counts", _p(planes), planes.shape[0], L, k, _p(cnt1), _p(cnt2), _stream())
    return cnt1, cnt2


def bn1_train_prepare(cnt1, cnt2, n_pos, conv_w, conv_b, gamma, beta, eps, momentum, running_mean, running_var):
    U, _, k = conv_w.shape
    dev = conv_w.device
    scale = torch.empty(U, dtype=torch.float32, device=dev)
    shift = torch.empty(U, dtype=torch.float32, device=dev)
    stats = torch.empty((U, 4), dtype=torch.float64, device=dev)
    cov_w = torch.empty((U, 4 * k), dtype=torch.float64, device=dev)
    _call("xnn_bn1_train_prepare", _p(cnt1), _p(cnt2), int(n_pos), _p(_req(conv_w, torch.float32, "conv_w")),
          _p(conv_b), _p(gamma), _p(beta), float(eps), float(momentum), U, k, _p(running_mean), _p(running_var),
          _p(scale), _p(shift), _p(stats), _p(cov_w), _stream())
    return scale, shift, stats, cov_w


def bn_fold_eval(gamma, beta, running_mean, running_var, pre_bias, eps):
    C = gamma.numel()
    scale = torch.empty(C, dtype=torch.float32, device=gamma.device)
    shift = torch.empty(C, dtype=torch.float32, device=gamma.device)
    _call("xnn_bn_fold_eval", _p(_req(gamma, torch.float32, "gamma")), _p(beta), _p(running_mean),
          _p(running_var), _p(pre_bias), float(eps), C, _p(scale), _p(shift), _stream())
    return scale, shift


def frontend_fwd(planes, L, conv_w, scale, shift, need_argmax, lut=False):
    _req(planes, torch.int32, "planes")
    _req(conv_w, torch.float32, "conv_w")
    B = planes.shape[0]
    U, _, k = conv_w.shape
    n = (L - k + 1) // POOL
    dev = planes.device
    pool = torch.empty((B, U, n), dtype=torch.float32, device=dev)
    zmax = torch.empty((B, U, n), dtype=torch.float32, device=dev) if need_argmax else None
    arg = torch.empty((B, U, n), dtype=torch.uint8, device=dev) if need_argmax else None
    if lut:
        _call("xnn_frontend_fwd_lut", _p(planes), B, L, U, k, _p(conv_w), _p(scale), _p(shift), _p(pool), _p(zmax),
              _p(arg), _stream())
        return pool, zmax, arg
    work = _work(_lib.load().xnn_frontend_work_bytes(U, k), dev)
    _call("xnn_frontend_fwd", _p(planes), B, L, U, k, _p(conv_w), _p(scale), _p(shift), _p(pool), _p(zmax),
          _p(arg), _p(work), _stream())
    return pool, zmax, arg


def frontend_acts(planes, L, conv_w, scale, shift, f32=True, f16=False):
    _req(planes, torch.int32, "planes")
    _req(conv_w, torch.float32, "conv_w")
    B = planes.shape[0]
    U, _, k = conv_w.shape
    Lc = L - k + 1
    dev = planes.device
    a32 = torch.empty((B, U, Lc), dtype=torch.float32, device=dev) if f32 else None
    a16 = torch.empty((B, U, Lc), dtype=torch.float16, device=dev) if f16 else None
    _call("xnn_frontend_acts", _p(planes), B, L, U, k, _p(conv_w), _p(scale), _p(shift), _p(a32), _p(a16),
          _stream())
    return a32, a16


# ---- a4/a5: MLP + final -------------------------------------------------------------
def fc1_fwd(pool, w1, b1):
    B, U, n = pool.shape
    hpre = torch.empty((B, U, HIDDEN), dtype=torch.float32, device=pool.device)
    _call("xnn_fc1_fwd", _p(_req(pool, torch.float32, "pool")), _p(_req(w1, torch.float32, "w1")), _p(b1), B, U, n,
          _p(hpre), _stream())
    return hpre


def _work(nbytes, dev):
    return torch.empty((int(nbytes) + 7) // 8, dtype=torch.float64, device=dev)


def _partial_view(work, nbytes):
    """the leading `nbytes` of a work buffer as float64 (the partial sums a sync-BatchNorm exchange all-reduces)"""
    return work.view(-1).view(torch.float64)[: int(nbytes) // 8] if work.dtype != torch.float64 else \
        work.view(-1)[: int(nbytes) // 8]


def bn_train_prepare(x2d, gamma, beta, eps, momentum, running_mean, running_var, sync=None, world=1):
    """x2d: (B, C) view of the BatchNorm input; returns scale, shift, stats.
    sync: callable all-reducing (SUM) a tensor in place over the data-parallel ranks -> statistics of the GLOBAL
    batch (B * world values per channel)."""
    B, C = x2d.shape
    if B * world < 2:
        raise ValueError("Expected more than 1 value per channel when training")
    lib = _lib.load()
    dev = x2d.device
    scale = torch.empty(C, dtype=torch.float32, device=dev)
    shift = torch.empty(C, dtype=torch.float32, device=dev)
    stats = torch.empty((C, 4), dtype=torch.float64, device=dev)
    work = _work(lib.xnn_bn_work_bytes(B, C), dev)

    def call(phase):
        _call("xnn_bn_train_prepare", _p(_req(x2d, torch.float32, "x")), B, C, _p(gamma), _p(beta), float(eps),
              float(momentum), _p(running_mean), _p(running_var), _p(scale), _p(shift), _p(stats), _p(work), phase,
              B * world, _stream())
    if sync is None:
        call(0)
    else:
        call(1)
        sync(_partial_view(work, lib.xnn_bn_partial_bytes(B, C)))
        call(2)
    return scale, shift, stats


def fc2_fwd(hpre, scale2, shift2, w2, b2, drop_p=0.0, seed=0, mask=None):
    B, U, _ = hpre.shape
    opre = torch.empty((B, U), dtype=torch.float32, device=hpre.device)
    if mask is not None:
        _req(mask, torch.uint8, "mask")
    _call("xnn_fc2_fwd", _p(_req(hpre, torch.float32, "hpre")), _p(scale2), _p(shift2),
          _p(_req(w2, torch.float32, "w2")), _p(b2), B, U, float(drop_p), int(seed), _p(mask), _p(opre), _stream())
    return opre


def mlp_supported(n):
    return bool(_lib.load().xnn_mlp_supported(int(n)))


def mlp_image_buffer(B, U, n, dev):
    """buffer for the tile images of one training step (xnn_mlp_image_bytes), 256-byte aligned"""
    nbytes = int(_lib.load().xnn_mlp_image_bytes(int(B), int(U), int(n)))
    buf = torch.empty(nbytes + 256, dtype=torch.uint8, device=dev)
    off = (-buf.data_ptr()) % 256
    return buf[off:off + nbytes]


def mlp_bn2_train_prepare(pool, w1, b1, gamma, beta, eps, momentum, running_mean, running_var, sync=None, world=1,
                          image=None):
    """BatchNorm (linears.7) batch statistics of hpre = FC1(pool) without materialising hpre (tcgen05).
    image: optional mlp_image_buffer: the pass also writes its operand tiles for mlp_fwd / mlp_bwd of this step."""
    B, U, n = pool.shape
    if B * world < 2:
        raise ValueError("Expected more than 1 value per channel when training")
    lib = _lib.load()
    dev = pool.device
    C = U * HIDDEN
    scale = torch.empty(C, dtype=torch.float32, device=dev)
    shift = torch.empty(C, dtype=torch.float32, device=dev)
    stats = torch.empty((C, 4), dtype=torch.float64, device=dev)
    work = _work(lib.xnn_mlp_bn2_work_bytes(B, U), dev)

    def call(phase):
        _call("xnn_mlp_bn2_train_prepare", _p(_req(pool, torch.float32, "pool")), _p(_req(w1, torch.float32, "w1")),
              _p(b1), B, U, n, _p(gamma), _p(beta), float(eps), float(momentum), _p(running_mean), _p(running_var),
              _p(scale), _p(shift), _p(stats), _p(work), _p(image), phase, B * world, _stream())
    if sync is None:
        call(0)
    else:
        call(1)
        sync(work)
        call(2)
    return scale, shift, stats


def mlp_fwd(pool, w1, b1, scale2, shift2, w2, b2, drop_p=0.0, seed=0, mask=None, want_bits=False, seed_dev=None,
            image=None):
    """opre (B,U) = FC2(drop(relu(BN2(FC1(pool))))) in one tcgen05 pass; optional keep-bits (B,U,4) int32.
    seed_dev: int64 (1,) device draw counter mixed into the seed (CUDA-graph replays)."""
    B, U, n = pool.shape
    dev = pool.device
    opre = torch.empty((B, U), dtype=torch.float32, device=dev)
    bits = torch.empty((B, U, 4), dtype=torch.int32, device=dev) if want_bits else None
    if mask is not None:
        _req(mask, torch.uint8, "mask")
    _call("xnn_mlp_fwd", _p(_req(pool, torch.float32, "pool")), _p(_req(w1, torch.float32, "w1")), _p(b1),
          _p(scale2), _p(shift2), _p(_req(w2, torch.float32, "w2")), _p(b2), B, U, n, float(drop_p), int(seed),
          _p(seed_dev), _p(mask), _p(bits), _p(opre), _p(image), _stream())
    return opre, bits


def _outs(out, shapes, dev):
    """`out`: None (allocate) or a list of preallocated contiguous fp32 tensors with these element counts (views
    of the flat gradient arena: the backward kernels write the gradients where the all-reduce and Adam read them)"""
    if out is None:
        return [torch.empty(sh, dtype=torch.float32, device=dev) for sh in shapes]
    if len(out) != len(shapes):
        raise ValueError("out: wrong number of tensors")
    for t, sh in zip(out, shapes):
        _req(t, torch.float32, "out")
        numel = 1
        for d in sh:
            numel *= int(d)
        if t.numel() != numel:
            raise ValueError(f"out: expected {numel} elements, got {t.numel()}")
    return list(out)


def mlp_bwd(dopre, pool, w1, b1, scale2, shift2, stats2, gamma2, w2, bits, drop_p, out=None, sync=None, world=1,
            image=None):
    """Backward of FC1 -> BN2 -> ReLU -> dropout -> FC2 from dopre (B,U); see xnn_mlp_bwd.
    out: optional [dw1, db1, dg2, dbe2, dw2, db2] destinations.
    The last result is dy = dpool * pool -- or, with `image`, dpool alone (frontend_bwd applies the factor)."""
    B, U, n = pool.shape
    dev = pool.device
    lib = _lib.load()
    C = U * HIDDEN
    dw1, db1, dg2, dbe2, dw2, db2 = _outs(out, [(C, n), (C,), (C,), (C,), (U, HIDDEN), (U,)], dev)
    dy = torch.empty_like(pool)
    nbytes = int(lib.xnn_mlp_bwd_work_bytes(B, U, n))
    work = torch.empty(nbytes + 256, dtype=torch.uint8, device=dev)
    off = (-work.data_ptr()) % 256
    wk = work[off:off + nbytes]

    def call(phase):
        _call("xnn_mlp_bwd", _p(_req(dopre, torch.float32, "dopre")), _p(_req(pool, torch.float32, "pool")),
              _p(_req(w1, torch.float32, "w1")), _p(b1), _p(scale2), _p(shift2),
              _p(_req(stats2, torch.float64, "stats2")), _p(gamma2), _p(_req(w2, torch.float32, "w2")), _p(bits),
              float(drop_p), B, U, n, _p(dw1), _p(db1), _p(dg2), _p(dbe2), _p(dw2), _p(db2), _p(dy), _p(wk), _p(image),
              phase, B * world, 1.0 / world, _stream())
    if sync is None:
        call(0)
    else:
        call(1)
        sync(_partial_view(wk, lib.xnn_mlp_bwd_partial_bytes(B, U)))
        call(2)
    return dw1, db1, dg2, dbe2, dw2, db2, dy


def head_fwd(opre, scale3, shift3, wf=None, bf=None):
    B, U = opre.shape
    o = torch.empty((B, U), dtype=torch.float32, device=opre.device)
    logits = None
    F = 1
    if wf is not None:
        F = wf.shape[0]
        logits = torch.empty((B, F), dtype=torch.float32, device=opre.device)
    _call("xnn_head_fwd", _p(_req(opre, torch.float32, "opre")), _p(scale3), _p(shift3), _p(wf), _p(bf), B, U, F,
          _p(o), _p(logits), _stream())
    return o, logits


def linear_fwd(x, w, b):
    B, U = x.shape
    F = w.shape[0]
    y = torch.empty((B, F), dtype=torch.float32, device=x.device)
    _call("xnn_linear_fwd", _p(_req(x, torch.float32, "x")), _p(_req(w, torch.float32, "w")),
          _p(_req(b, torch.float32, "b")), B, U, F, _p(y), _stream())
    return y


# ---- a6: backward ----------------------------------------------------------------------
def linear_bwd(dy, w, x, out=None):
    B, F = dy.shape
    U = w.shape[1]
    dev = dy.device
    lib = _lib.load()
    dx = torch.empty((B, U), dtype=torch.float32, device=dev)
    dw, db = _outs(out, [(F, U), (F,)], dev)
    work = _work(lib.xnn_linear_bwd_work_bytes(B, U, F), dev)
    _call("xnn_linear_bwd", _p(_req(dy, torch.float32, "dy")), _p(w), _p(_req(x, torch.float32, "x")), B, U, F,
          _p(dx), _p(dw), _p(db), _p(work), _stream())
    return dx, dw, db


def bn_bwd(g2d, relu_out, xpre2d, stats, gamma, inplace=True, out=None, sync=None, world=1):
    B, C = g2d.shape
    dev = g2d.device
    lib = _lib.load()
    dgamma, dbeta = _outs(out, [(C,), (C,)], dev)
    dx = g2d if inplace else torch.empty_like(g2d)
    work = _work(lib.xnn_bn_work_bytes(B, C), dev)

    def call(phase):
        _call("xnn_bn_bwd", _p(_req(g2d, torch.float32, "g")), _p(relu_out), _p(_req(xpre2d, torch.float32, "xpre")),
              _p(stats), _p(gamma), B, C, _p(dgamma), _p(dbeta), _p(dx), _p(work), phase, B * world, 1.0 / world,
              _stream())
    if sync is None:
        call(0)
    else:
        call(1)
        sync(_partial_view(work, lib.xnn_bn_partial_bytes(B, C)))
        call(2)
    return dgamma, dbeta, dx


def fc2_bwd(dopre, hpre, scale2, shift2, w2, drop_p=0.0, seed=0, mask=None):
    B, U, _ = hpre.shape
    dev = hpre.device
    lib = _lib.load()
    dw2 = torch.empty((U, HIDDEN), dtype=torch.float32, device=dev)
    db2 = torch.empty(U, dtype=torch.float32, device=dev)
    g2 = torch.empty_like(hpre)
    work = _work(lib.xnn_fc2_bwd_work_bytes(B, U), dev)
    _call("xnn_fc2_bwd", _p(_req(dopre, torch.float32, "dopre")), _p(hpre), _p(scale2), _p(shift2), _p(w2), B, U,
          float(drop_p), int(seed), _p(mask), _p(dw2), _p(db2), _p(g2), _p(work), _stream())
    return dw2, db2, g2


def fc1_bwd(dhpre, pool, w1):
    B, U, n = pool.shape
    dev = pool.device
    lib = _lib.load()
    dw1 = torch.empty((U * HIDDEN, n), dtype=torch.float32, device=dev)
    db1 = torch.empty(U * HIDDEN, dtype=torch.float32, device=dev)
    dy = torch.empty_like(pool)
    work = _work(lib.xnn_fc1_bwd_work_bytes(B, U, n), dev)
    _call("xnn_fc1_bwd", _p(_req(dhpre, torch.float32, "dhpre")), _p(pool), _p(w1), B, U, n, _p(dw1), _p(db1),
          _p(dy), _p(work), _stream())
    return dw1, db1, dy


def frontend_bwd(planes, L, k, dy, zmax, argmax, gamma, stats, cov_w, cnt1, n_pos, simt=False, out=None,
                 act_scale=None, act_shift=None):
    """act_scale / act_shift: `dy` is dL/dpool (what mlp_bwd returns on tile images); the kernel multiplies by
    pool = exp(act_scale * zmax + act_shift)."""
    B, U, n = dy.shape
    dev = dy.device
    lib = _lib.load()
    dconv_w, dconv_b, dgamma, dbeta = _outs(out, [(U, 4, k), (U,), (U,), (U,)], dev)
    work = _work(lib.xnn_frontend_bwd_work_bytes(B, U, k), dev)
    if simt:
        if act_scale is not None:
            raise ValueError("the FP64 cross-check kernel takes the true dy")
        _call("xnn_frontend_bwd_simt", _p(planes), B, L, U, k, _p(_req(dy, torch.float32, "dy")), _p(zmax), _p(argmax),
              _p(gamma), _p(stats), _p(cov_w), _p(cnt1), int(n_pos), _p(dconv_w), _p(dconv_b), _p(dgamma), _p(dbeta),
              _p(work), _stream())
    else:
        _call("xnn_frontend_bwd", _p(planes), B, L, U, k, _p(_req(dy, torch.float32, "dy")), _p(zmax), _p(argmax),
              _p(gamma), _p(stats), _p(cov_w), _p(cnt1), int(n_pos), _p(act_scale), _p(act_shift), _p(dconv_w),
              _p(dconv_b), _p(dgamma), _p(dbeta), _p(work), _stream())
    return dconv_w, dconv_b, dgamma, dbeta


def adam_step(param, grad, exp_avg, exp_avg_sq, lr, beta1, beta2, eps, weight_decay, step, grad_scale=1.0):
    for t, nme in ((param, "param"), (grad, "grad"), (exp_avg, "exp_avg"), (exp_avg_sq, "exp_avg_sq")):
        _req(t, torch.float32, nme)
    _call("xnn_adam_step", _p(param), _p(grad), _p(exp_avg), _p(exp_avg_sq), param.numel(), float(lr), float(beta1),
          float(beta2), float(eps), float(weight_decay), int(step), float(grad_scale), _stream())


# ---- a7-a10: interpret ----------------------------------------------------------------------
def _scan_tc_work(L, rows, width, dev, tensor_cores):
    """work buffer of the tcgen05 interpret kernels, or None when the shape is outside that path (the three
    passes switch together: counts and emission must classify every start identically)"""
    lib = _lib.load()
    if not tensor_cores or not lib.xnn_interpret_tc_supported(int(L), int(width)):
        return None
    return _work(lib.xnn_pwm_scan_tc_work_bytes(int(rows), int(width)), dev)


def act_max(planes, L, conv_w, scale, shift, idx=None, tensor_cores=True):
    """fp16-rounded per-unit maximum activation over the listed sequences -> float32 (U,)."""
    B = planes.shape[0]
    U, _, k = conv_w.shape
    n_idx = B if idx is None else idx.numel()
    bits = torch.zeros(U, dtype=torch.int32, device=planes.device)
    if idx is not None:
        _req(idx, torch.int32, "idx")
    work = _scan_tc_work(L, U, k, planes.device, tensor_cores)
    if work is not None:
        _call("xnn_act_max_tc", _p(_req(planes, torch.int32, "planes")), B, L, U, k, _p(conv_w), _p(scale),
              _p(shift), _p(idx), n_idx, _p(bits), _p(work), _stream())
    else:
        _call("xnn_act_max", _p(_req(planes, torch.int32, "planes")), B, L, U, k, _p(conv_w), _p(scale), _p(shift),
              _p(idx), n_idx, _p(bits), _stream())
    return bits.view(torch.float32)


def site_counts(planes, L, conv_w, scale, shift, thr, idx=None, tensor_cores=True):
    B = planes.shape[0]
    U, _, k = conv_w.shape
    n_idx = B if idx is None else idx.numel()
    counts = torch.zeros((U, n_idx), dtype=torch.int32, device=planes.device)
    work = _scan_tc_work(L, U, k, planes.device, tensor_cores)
    if work is not None:
        _call("xnn_site_counts_tc", _p(planes), B, L, U, k, _p(conv_w), _p(scale), _p(shift),
              _p(_req(thr, torch.float32, "thr")), _p(idx), n_idx, _p(counts), _p(work), _stream())
    else:
        _call("xnn_site_counts", _p(planes), B, L, U, k, _p(conv_w), _p(scale), _p(shift),
              _p(_req(thr, torch.float32, "thr")), _p(idx), n_idx, _p(counts), _stream())
    return counts


def site_offsets(counts, carry, out=None):
    """offsets[u, i] = carry[u] + counts[u, :i].sum() (int64), then carry[u] += counts[u].sum(), in one launch.
    counts (U, n) int32 from site_counts; carry (U,) int64, updated in place."""
    U, n = counts.shape
    _req(counts, torch.int32, "counts")
    _req(carry, torch.int64, "carry")
    if out is None:
        out = torch.empty((U, n), dtype=torch.int64, device=counts.device)
    _call("xnn_site_offsets", _p(counts), U, n, _p(carry), _p(_req(out, torch.int64, "offsets")), _stream())
    return out


def site_emit(planes, L, conv_w, scale, shift, thr, offsets, cap, idx=None, pfm=None, sites=None, tensor_cores=True):
    B = planes.shape[0]
    U, _, k = conv_w.shape
    n_idx = B if idx is None else idx.numel()
    if pfm is None:
        pfm = torch.zeros((U, 4, k), dtype=torch.int32, device=planes.device)
    site_cap = 0 if sites is None else sites.shape[1]
    work = _scan_tc_work(L, U, k, planes.device, tensor_cores)
    if work is not None:
        _call("xnn_site_emit_tc", _p(planes), B, L, U, k, _p(conv_w), _p(scale), _p(shift), _p(thr), _p(idx), n_idx,
              _p(_req(offsets, torch.int64, "offsets")), int(cap), _p(pfm), _p(sites), int(site_cap), _p(work),
              _stream())
        return pfm
    _call("xnn_site_emit", _p(planes), B, L, U, k, _p(conv_w), _p(scale), _p(shift), _p(thr), _p(idx), n_idx,
          _p(_req(offsets, torch.int64, "offsets")), int(cap), _p(pfm), _p(sites), int(site_cap), _stream())
    return pfm


# ---- a11: PWM module ---------------------------------------------------------------------------
def pwm_scan(planes, L, pwm_w, mode, want_argmax=False, tensor_cores=True):
    """The PWM module's scan.  tensor_cores: the tcgen05 Toeplitz kernel (xnn_pwm_scan_tc) whenever the shape
    fits it; False (or a width above 48 / a sequence too long for shared memory) runs the FP32 gather-add scan."""
    B = planes.shape[0]
    G, _, w = pwm_w.shape
    out = torch.empty((B, G), dtype=torch.float32, device=planes.device)
    arg = torch.empty((B, G), dtype=torch.int32, device=planes.device) if want_argmax else None
    if tensor_cores and _lib.load().xnn_pwm_scan_tc_supported(int(L), int(w)):
        work = _work(_lib.load().xnn_pwm_scan_tc_work_bytes(int(G), int(w)), planes.device)
        _call("xnn_pwm_scan_tc", _p(_req(planes, torch.int32, "planes")), B, L, G, w,
              _p(_req(pwm_w, torch.float32, "pwm_w")), 1 if mode == "max" else 0, _p(out), _p(arg), _p(work), _stream())
        return out, arg
    _call("xnn_pwm_scan", _p(_req(planes, torch.int32, "planes")), B, L, G, w,
          _p(_req(pwm_w, torch.float32, "pwm_w")), 1 if mode == "max" else 0, _p(out), _p(arg), _stream())
    return out, arg


# ---- a13/a14: PWMScan semantics -------------------------------------------------------------------
def lpm_scan_sum(codes, offsets, lpm, bg, forward_only=False):
    R = offsets.numel() - 1
    per_rec = int(bg.dim() == 2)
    w = lpm.shape[1]
    out = torch.empty(R, dtype=torch.float64, device=codes.device)
    work = _work(_lib.load().xnn_scan_work_bytes(int(R)), codes.device)
    _call("xnn_lpm_scan_sum_parts", _p(_req(codes, torch.int8, "codes")), _p(_req(offsets, torch.int64, "offsets")), R,
          w, _p(_req(lpm, torch.float64, "lpm")), _p(_req(bg, torch.float64, "bg")), per_rec, int(forward_only),
          _p(out), _p(work), _stream())
    return out


def lpm_scan_best(codes, offsets, lpm, bg, forward_only=False, tie_cap=16):
    R = offsets.numel() - 1
    per_rec = int(bg.dim() == 2)
    w = lpm.shape[1]
    dev = codes.device
    best = torch.empty(R, dtype=torch.float64, device=dev)
    pos = torch.empty(R, dtype=torch.int32, device=dev)
    strand = torch.empty(R, dtype=torch.int8, device=dev)
    nties = torch.empty(R, dtype=torch.int32, device=dev)
    ties = torch.full((R, max(tie_cap, 1)), -1, dtype=torch.int32, device=dev)
    work = _work(_lib.load().xnn_scan_work_bytes(int(R)), dev)
    _call("xnn_lpm_scan_best_parts", _p(_req(codes, torch.int8, "codes")), _p(_req(offsets, torch.int64, "offsets")),
          R, w, _p(_req(lpm, torch.float64, "lpm")), _p(_req(bg, torch.float64, "bg")), per_rec, int(forward_only),
          _p(best), _p(pos), _p(strand), _p(nties), _p(ties), int(tie_cap), _p(work), _stream())
    return best, pos, strand, nties, ties


def ipwm_scan_best(codes, offsets, pwm, forward_only=False):
    R = offsets.numel() - 1
    w = pwm.shape[1]
    dev = codes.device
    best = torch.empty(R, dtype=torch.int32, device=dev)
    pos = torch.empty(R, dtype=torch.int32, device=dev)
    strand = torch.empty(R, dtype=torch.int8, device=dev)
    work = _work(_lib.load().xnn_scan_work_bytes(int(R)), dev)
    _call("xnn_ipwm_scan_best_parts", _p(_req(codes, torch.int8, "codes")), _p(_req(offsets, torch.int64, "offsets")),
          R, w, _p(_req(pwm, torch.int32, "pwm")), int(forward_only), _p(best), _p(pos), _p(strand), _p(work),
          _stream())
    return best, pos, strand


def counters_advance(counters):
    """counters (int64 device tensor, <= 32 values) += 1: the device-resident Adam step / dropout draw counters"""
    _req(counters, torch.int64, "counters")
    _call("xnn_counters_advance", _p(counters), counters.numel(), _stream())


def counters_advance_list(tensors):
    """every int64 scalar tensor of the list += 1 in one launch (BatchNorm's num_batches_tracked buffers)"""
    ts = [t for t in tensors if t is not None]
    if not ts:
        return
    for t in ts:
        _req(t, torch.int64, "counter")
        _seen_devices.append(t.device)
    arr = (ctypes.c_void_p * len(ts))(*[t.data_ptr() for t in ts])
    _call("xnn_counters_advance_list", len(ts), arr, _stream())


def bn_running_update(stats, running_means, running_vars, nbts, momenta, counts):
    """running statistics (+ num_batches_tracked) of up to 4 train-mode BatchNorm layers in one launch"""
    n = len(stats)
    for t in stats:
        _req(t, torch.float64, "stats")
        _seen_devices.append(t.device)
    for t in list(running_means) + list(running_vars):
        _req(t, torch.float32, "running statistics")
    P = ctypes.c_void_p * n
    nb = P(*[(None if t is None else t.data_ptr()) for t in nbts]) if nbts is not None else None
    _call("xnn_bn_running_update", n, P(*[t.data_ptr() for t in stats]), P(*[t.data_ptr() for t in running_means]),
          P(*[t.data_ptr() for t in running_vars]), nb, (ctypes.c_int * n)(*[t.shape[0] for t in stats]),
          (ctypes.c_float * n)(*[float(m) for m in momenta]), (ctypes.c_double * n)(*[float(c) for c in counts]),
          _stream())


_loss_work = {}


def _loss_workspace(dev):
    w = _loss_work.get(dev)
    if w is None:
        w = _loss_work[dev] = torch.zeros((int(_lib.load().xnn_loss_work_bytes()) + 7) // 8, dtype=torch.float64,
                                          device=dev)
    return w


LOSS_KINDS = {"mse": 0, "bce_logits": 1}


def loss_fwd(logits, target, kind):
    """mean MSE / BCE-with-logits over all elements -> float32 scalar tensor (architectures/__init__.py:446-456)"""
    _req(logits, torch.float32, "logits")
    _req(target, torch.float32, "target")
    if logits.shape != target.shape:
        raise ValueError(f"loss: input {tuple(logits.shape)} and target {tuple(target.shape)} differ in shape")
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    _call("xnn_loss_fwd", _p(logits), _p(target), logits.numel(), LOSS_KINDS[kind], _p(loss),
          _p(_loss_workspace(logits.device)), _stream())
    return loss


def loss_bwd(logits, target, grad_out, kind):
    dl = torch.empty_like(logits)
    _call("xnn_loss_bwd", _p(logits), _p(target), _p(grad_out), logits.numel(), LOSS_KINDS[kind], _p(dl), _stream())
    return dl


def adam_step_multi(params, grads, exp_avgs, exp_avg_sqs, lr, beta1, beta2, eps, weight_decay, step, grad_scale=1.0,
                    step_dev=None):
    """One launch for a list of parameter tensors that share `step` (xnn_adam_step_multi); bit-identical to
    adam_step per tensor.  step_dev: int64 (1,) device tensor holding the step count (CUDA-graph replays)."""
    n = len(params)
    if not (len(grads) == len(exp_avgs) == len(exp_avg_sqs) == n):
        raise ValueError("adam_step_multi: lists of different lengths")
    for p_, g_, m_, v_ in zip(params, grads, exp_avgs, exp_avg_sqs):
        _req(p_, torch.float32, "param"); _req(g_, torch.float32, "grad")
        _req(m_, torch.float32, "exp_avg"); _req(v_, torch.float32, "exp_avg_sq")
        if not (g_.numel() == m_.numel() == v_.numel() == p_.numel()):
            raise ValueError("adam_step_multi: tensor sizes differ")
    arr = ctypes.c_void_p * n
    ptrs = [arr(*[t.data_ptr() for t in lst]) for lst in (params, grads, exp_avgs, exp_avg_sqs)]
    numels = (ctypes.c_longlong * n)(*[t.numel() for t in params])
    for t in params:
        _seen_devices.append(t.device)
    _call("xnn_adam_step_multi", n, ptrs[0], ptrs[1], ptrs[2], ptrs[3], numels, float(lr), float(beta1), float(beta2),
          float(eps), float(weight_decay), int(step), _p(step_dev), float(grad_scale), _stream())


# ---- validation metrics (explainn_b200/metrics.py is the reference-facing surface) ----
def sort_columns(vals, descending=False):
    """every row of vals float32 (G,N) sorted on its own -> (sorted values, int64 permutation); xnn_sort_columns
    (stable radix sort, one block per row) instead of torch.sort"""
    _req(vals, torch.float32, "vals")
    G, N = vals.shape
    sv = torch.empty_like(vals)
    perm = torch.empty((G, N), dtype=torch.int64, device=vals.device)
    if G == 0 or N == 0:
        return sv, perm
    work = torch.empty(int(_lib.load().xnn_sort_columns_work_bytes(N, G)) // 4, dtype=torch.int32, device=vals.device)
    _call("xnn_sort_columns", _p(vals), N, G, int(bool(descending)), _p(sv), _p(perm), _p(work), _stream())
    return sv, perm


def metric_auc(scores, labels):
    """scores float32 (G,N) columns contiguous; labels uint8 (N,) shared or (G,N).  Returns (auroc, aupr) float64 (G,).
    The descending sort is xnn_sort_columns; ranks over ties, the cumulative positives and both reductions are
    xnn_metric_auc."""
    _req(scores, torch.float32, "scores")
    _req(labels, torch.uint8, "labels")
    G, N = scores.shape
    if labels.dim() == 1:
        if labels.shape[0] != N:
            raise ValueError("labels: expected N entries")
        stride = 0
    else:
        if tuple(labels.shape) != (G, N):
            raise ValueError("labels: expected shape (N,) or (G,N)")
        stride = N
    sv, perm = sort_columns(scores, descending=True)
    auroc = torch.empty(G, dtype=torch.float64, device=scores.device)
    aupr = torch.empty(G, dtype=torch.float64, device=scores.device)
    lib = _lib.load()
    work = torch.empty(int(lib.xnn_metric_auc_work_bytes(N, G)), dtype=torch.uint8, device=scores.device)
    _call("xnn_metric_auc", _p(sv), _p(perm), _p(labels), stride, N, G, _p(auroc), _p(aupr), _p(work), _stream())
    return auroc, aupr


def metric_ranks(vals):
    """Average 1-based ranks (ties share their mean position) of every row of vals float32 (G,N) -> float64 (G,N)."""
    _req(vals, torch.float32, "vals")
    G, N = vals.shape
    sv, perm = sort_columns(vals)
    ranks = torch.empty((G, N), dtype=torch.float64, device=vals.device)
    _call("xnn_metric_ranks", _p(sv), _p(perm), N, G, _p(ranks), _stream())
    return ranks


def metric_pearson(x, y):
    """Pearson r of row g of x (G,N) against y (N,) or row g of y (G,N); float32 or float64 inputs -> float64 (G,)."""
    if x.dtype not in (torch.float32, torch.float64):
        raise TypeError("x: expected float32 or float64")
    _req(x, x.dtype, "x")
    _req(y, x.dtype, "y")
    G, N = x.shape
    if y.dim() == 1:
        if y.shape[0] != N:
            raise ValueError("y: expected N entries")
        stride = 0
    else:
        if tuple(y.shape) != (G, N):
            raise ValueError("y: expected shape (N,) or (G,N)")
        stride = N
    r = torch.empty(G, dtype=torch.float64, device=x.device)
    _call("xnn_metric_pearson_f32" if x.dtype == torch.float32 else "xnn_metric_pearson_f64", _p(x), _p(y), stride, N, G,
          _p(r), _stream())
    return r
