"""xnn_pwm_scan_tc_noalloc (static device scratch, no allocation) against xnn_pwm_scan_tc, through the C-ABI.
Collected last on purpose."""
import ctypes

import numpy as np
import pytest
import torch

import _util as U
from oracle import explainn_oracle as O

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _scan(lib, name, planes, B, L, G, w, pw, mode, want_arg, work=None):
    out = torch.full((B, G), float("nan"), dtype=torch.float32, device=planes.device)
    arg = torch.full((B, G), -7, dtype=torch.int32, device=planes.device) if want_arg else None
    st = ctypes.c_void_p(torch.cuda.current_stream(planes.device).cuda_stream)
    args = [_ptr(planes), B, L, G, w, _ptr(pw), mode, _ptr(out), _ptr(arg)]
    if work is not None:
        args.append(_ptr(work))
    rc = getattr(lib, name)(*args, st)
    torch.cuda.synchronize()
    return rc, out, arg


@pytest.mark.parametrize("cfg", [(70, 35, 300, 9), (300, 19, 1000, 5), (5, 6, 40, 3)])
def test_pwm_scan_noalloc_matches_the_work_buffer_entry_point(cfg):
    from explainn_b200 import _lib, ops
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    G, w, L, B = cfg
    assert lib.xnn_pwm_scan_tc_supported(L, w) == 1
    assert lib.xnn_pwm_scan_tc_work_bytes(G, w) <= lib.xnn_pwm_scan_tc_noalloc_max_bytes() == 1 << 20
    # two calls in a row on DIFFERENT inputs: the second finds the first call's pieces in the shared scratch
    for call, seed in enumerate((G + w, G + w + 1)):
        rng = np.random.default_rng(seed)
        pw_h = np.ascontiguousarray(np.log(np.maximum(rng.dirichlet([0.5] * 4, size=(G, w)).transpose(0, 2, 1), 1e-4)))
        x = O.random_onehot(B, L, seed=seed, n_frac=0.01)
        planes = ops.pack_onehot(x.to(dev))
        pw = torch.tensor(pw_h, dtype=torch.float32, device=dev)
        work = torch.empty(lib.xnn_pwm_scan_tc_work_bytes(G, w), dtype=torch.uint8, device=dev)
        for mode, want_arg in ((1, True), (1, False), (0, False)):
            rc0, ref, ref_i = _scan(lib, "xnn_pwm_scan_tc", planes, B, L, G, w, pw, mode, want_arg, work)
            rc1, got, got_i = _scan(lib, "xnn_pwm_scan_tc_noalloc", planes, B, L, G, w, pw, mode, want_arg)
            assert rc0 == 0 and rc1 == 0, lib.xnn_last_error()
            U.assert_close(got.cpu().numpy(), ref.cpu().numpy(), what=f"noalloc call {call} mode {mode}")
            assert torch.equal(got, ref)                      # the same kernels on the same pieces: bit-identical
            if want_arg:
                assert torch.equal(got_i, ref_i)
        ora, ora_i = O.pwm_module_scores(pw_h, x.numpy(), "max")
        _, got, got_i = _scan(lib, "xnn_pwm_scan_tc_noalloc", planes, B, L, G, w, pw, 1, True)
        U.assert_close(got.cpu().numpy(), ora, what="noalloc vs oracle")
        assert np.array_equal(got_i.cpu().numpy(), ora_i)


def test_pwm_scan_noalloc_rejects_what_does_not_fit():
    from explainn_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    G, w, L, B = 1200, 48, 100, 2                             # 10 M tiles x 147456 B > 1 MiB
    assert lib.xnn_pwm_scan_tc_work_bytes(G, w) > lib.xnn_pwm_scan_tc_noalloc_max_bytes()
    planes = torch.zeros((B, 3, (L + 31) // 32), dtype=torch.int32, device=dev)
    pw = torch.zeros((G, 4, w), dtype=torch.float32, device=dev)
    rc, out, _ = _scan(lib, "xnn_pwm_scan_tc_noalloc", planes, B, L, G, w, pw, 1, False)
    assert rc != 0 and b"static array" in lib.xnn_last_error()
    assert bool(torch.isnan(out).all())                       # nothing was launched
    # a width outside the tensor-core path is refused as well
    rc, _, _ = _scan(lib, "xnn_pwm_scan_tc_noalloc", planes, B, L, 4, 49, torch.zeros((4, 4, 49), device=dev), 1, False)
    assert rc != 0
