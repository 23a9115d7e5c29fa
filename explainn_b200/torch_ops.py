"""`torch.library` custom ops over the C ABI (BASELINE.json: "thin C-ABI torch custom op").

Every op here is registered under `torch.ops.explainn_b200.*` with a schema, a fake (meta) implementation and --
where it is differentiable -- `register_autograd`, so the path is visible to `torch.export`, `opcheck` and CUDA-graph
capture.  The bodies call explainn_b200/ops.py (ctypes -> include/explainn_b200.h); there is no CPU implementation:
the ops are registered for the "cuda" device type only, so a CPU tensor raises NotImplementedError from the
dispatcher (no fallback).

    pack_onehot           a16   float one-hot -> planes
    frontend_acts/_pool   a3    linears[:3] / linears[:4]           (explainn/architectures/__init__.py:73-83)
    unit_bank_eval        a3-a4 linears(x) with running statistics  (:73-101)
    unit_bank_train       a3-a4 + a6: train-mode forward; backward = unit_bank_train_bwd
    linear                a5    final Linear (:104); backward = linear_bwd
    loss                  K12   nn.MSELoss / nn.BCEWithLogitsLoss (:446-456); backward = loss_bwd
    pwm_scan              a11   PWM module (:116-170)

Gradient arena: the training ops take an optional flat fp32 `arena` and element offsets; their backward ops write
the parameter gradients INTO it (mutated argument) and autograd receives views of it, so p.grad of every parameter
is a slice of one buffer -- the data-parallel all-reduce and Adam read it in place (no bucket pack/unpack copies).
"""
from typing import List, Optional, Sequence

import torch
from torch.library import custom_op, register_autograd

from . import ops

HIDDEN = ops.HIDDEN

# order of the 12 parameters of the unit bank in every `params` list
BANK_PARAMS = ("conv_w", "conv_b", "g1", "be1", "w1", "b1", "g2", "be2", "w2", "b2", "g3", "be3")

# called with the arena right after the MLP-segment gradients (final, BN3, FC2, BN2, FC1) are complete and before the
# conv weight gradient starts: data-parallel training hooks its segment all-reduce here (parallel.GradBucket)
_after_mlp_grads = []
# sync-BatchNorm exchange (parallel.SyncBN.install): callables all-reducing a tensor in place across ranks
_sync = {"sum": None, "world": 1}


def _n_pool(L, k):
    return (L - k + 1) // ops.POOL


def _image_bytes(B, U, n):
    """xnn_mlp_image_bytes restated for the fake implementations: [U][ceil(B/128)] pool stages + [U] W1 images"""
    chunks = ((n + 15) // 16) * 2
    return U * ((B + 127) // 128) * 3 * chunks * 2048 + U * 3 * chunks * 112 * 16


# ------------------------------------------------------------------------------------------------------------
# a16 / a3
# ------------------------------------------------------------------------------------------------------------
@custom_op("explainn_b200::pack_onehot", mutates_args=("bad",), device_types="cuda")
def pack_onehot(x: torch.Tensor, bad: Optional[torch.Tensor]) -> torch.Tensor:
    return ops.pack_onehot(x, bad=bad)


@pack_onehot.register_fake
def _(x, bad):
    return x.new_empty((x.shape[0], 3, (x.shape[2] + 31) // 32), dtype=torch.int32)


@custom_op("explainn_b200::frontend_acts", mutates_args=(), device_types="cuda")
def frontend_acts(planes: torch.Tensor, L: int, conv_w: torch.Tensor, scale: torch.Tensor,
                  shift: torch.Tensor) -> torch.Tensor:
    return ops.frontend_acts(planes, L, conv_w, scale, shift)[0]


@frontend_acts.register_fake
def _(planes, L, conv_w, scale, shift):
    U, _, k = conv_w.shape
    return conv_w.new_empty((planes.shape[0], U, L - k + 1))


@custom_op("explainn_b200::frontend_pool", mutates_args=(), device_types="cuda")
def frontend_pool(planes: torch.Tensor, L: int, conv_w: torch.Tensor, scale: torch.Tensor,
                  shift: torch.Tensor) -> torch.Tensor:
    return ops.frontend_fwd(planes, L, conv_w, scale, shift, need_argmax=False)[0]


@frontend_pool.register_fake
def _(planes, L, conv_w, scale, shift):
    U, _, k = conv_w.shape
    return conv_w.new_empty((planes.shape[0], U, _n_pool(L, k)))


# ------------------------------------------------------------------------------------------------------------
# unit bank, eval mode (running statistics folded)
# ------------------------------------------------------------------------------------------------------------
@custom_op("explainn_b200::unit_bank_eval", mutates_args=(), device_types="cuda")
def unit_bank_eval(planes: torch.Tensor, L: int, params: Sequence[torch.Tensor], running: Sequence[torch.Tensor],
                   eps: Sequence[float], drop_p: float, seed: int, drop_mask: Optional[torch.Tensor]) -> torch.Tensor:
    cw, cb, g1, be1, w1, b1, g2, be2, w2, b2, g3, be3 = params
    rm1, rv1, rm2, rv2, rm3, rv3 = running
    scale1, shift1 = ops.bn_fold_eval(g1, be1, rm1, rv1, cb, eps[0])
    pool, _, _ = ops.frontend_fwd(planes, L, cw, scale1, shift1, need_argmax=False)
    scale2, shift2 = ops.bn_fold_eval(g2, be2, rm2, rv2, None, eps[1])
    if ops.mlp_supported(pool.shape[2]):
        opre, _ = ops.mlp_fwd(pool, w1, b1, scale2, shift2, w2, b2, drop_p, seed, drop_mask)
    else:
        hpre = ops.fc1_fwd(pool, w1, b1)
        opre = ops.fc2_fwd(hpre, scale2, shift2, w2, b2, drop_p, seed, drop_mask)
    scale3, shift3 = ops.bn_fold_eval(g3, be3, rm3, rv3, None, eps[2])
    return ops.head_fwd(opre, scale3, shift3)[0]


@unit_bank_eval.register_fake
def _(planes, L, params, running, eps, drop_p, seed, drop_mask):
    return params[0].new_empty((planes.shape[0], params[0].shape[0]))


# ------------------------------------------------------------------------------------------------------------
# unit bank, train mode
# outputs: o, pool, zmax, arg, opre, bits, scale2, shift2, stats1, stats2, stats3, cov_w, cnt1, hpre, image,
#          scale1, shift1
# ------------------------------------------------------------------------------------------------------------
N_TRAIN_OUT = 17


@custom_op("explainn_b200::unit_bank_train", mutates_args=(), device_types="cuda")
def unit_bank_train(planes: torch.Tensor, L: int, params: Sequence[torch.Tensor], eps: Sequence[float],
                    drop_p: float, seed: int, seed_dev: Optional[torch.Tensor], drop_mask: Optional[torch.Tensor],
                    arena: Optional[torch.Tensor], offs: Sequence[int]) -> List[torch.Tensor]:
    """FUNCTIONAL train-mode forward (register_autograd needs one): batch statistics come back in stats1/2/3 and
    the running statistics are updated from them by `bn_running_update` afterwards."""
    cw, cb, g1, be1, w1, b1, g2, be2, w2, b2, g3, be3 = params
    B = planes.shape[0]
    U, _, k = cw.shape
    sync, world = _sync["sum"], (_sync["world"] if _sync["sum"] else 1)
    n_pos = B * (L - k + 1) * world
    cnt1, cnt2 = ops.cooc_counts(planes, L, k)
    if sync:
        # synchronised BatchNorm, layer 1: the batch statistics are closed forms of INTEGER co-occurrence counts,
        # so one int64 all-reduce makes them exactly the single-process statistics of the global batch
        cnt = torch.cat((cnt1, cnt2.reshape(-1)))
        sync(cnt)
        cnt1, cnt2 = cnt[:4 * k].contiguous(), cnt[4 * k:].view(4 * k, 4 * k).contiguous()
    scale1, shift1, stats1, cov_w = ops.bn1_train_prepare(cnt1, cnt2, n_pos, cw, cb, g1, be1, eps[0], 0.0, None, None)
    pool, zmax, arg = ops.frontend_fwd(planes, L, cw, scale1, shift1, need_argmax=True)
    empty = cw.new_empty(0)
    if ops.mlp_supported(pool.shape[2]):
        # tensor-core path: the (B,100U) hidden activations are never materialised
        # the statistics pass also writes the tile images the later passes of this step load their operands from
        image = ops.mlp_image_buffer(B, U, pool.shape[2], cw.device)
        scale2, shift2, stats2 = ops.mlp_bn2_train_prepare(pool, w1, b1, g2, be2, eps[1], 0.0, None, None,
                                                           sync=sync, world=world, image=image)
        opre, bits = ops.mlp_fwd(pool, w1, b1, scale2, shift2, w2, b2, drop_p, seed, drop_mask, want_bits=True,
                                 seed_dev=seed_dev, image=image)
        hpre = empty
    else:
        if seed_dev is not None:
            raise NotImplementedError("device-resident dropout draws need the tensor-core MLP path (n <= 96)")
        hpre = ops.fc1_fwd(pool, w1, b1)
        scale2, shift2, stats2 = ops.bn_train_prepare(hpre.view(B, U * HIDDEN), g2, be2, eps[1], 0.0, None, None,
                                                      sync=sync, world=world)
        opre = ops.fc2_fwd(hpre, scale2, shift2, w2, b2, drop_p, seed, drop_mask)
        bits = torch.empty(0, dtype=torch.int32, device=cw.device)
        image = torch.empty(0, dtype=torch.uint8, device=cw.device)
    scale3, shift3, stats3 = ops.bn_train_prepare(opre, g3, be3, eps[2], 0.0, None, None, sync=sync, world=world)
    o, _ = ops.head_fwd(opre, scale3, shift3)
    return [o, pool, zmax, arg, opre, bits, scale2, shift2, stats1, stats2, stats3, cov_w, cnt1, hpre, image,
            scale1, shift1]


@custom_op("explainn_b200::bn_running_update", mutates_args=("running_means", "running_vars", "nbt"),
           device_types="cuda")
def bn_running_update(stats: Sequence[torch.Tensor], running_means: Sequence[torch.Tensor],
                      running_vars: Sequence[torch.Tensor], nbt: Sequence[torch.Tensor], momentum: Sequence[float],
                      counts: Sequence[float]) -> None:
    ops.bn_running_update(list(stats), list(running_means), list(running_vars), list(nbt) if len(nbt) else None,
                          momentum, counts)


@bn_running_update.register_fake
def _(stats, running_means, running_vars, nbt, momentum, counts):
    return None


@unit_bank_train.register_fake
def _(planes, L, params, eps, drop_p, seed, seed_dev, drop_mask, arena, offs):
    cw = params[0]
    B = planes.shape[0]
    U, _, k = cw.shape
    n = _n_pool(L, k)
    C = U * HIDDEN
    f32 = lambda *sh: cw.new_empty(sh)
    f64 = lambda *sh: cw.new_empty(sh, dtype=torch.float64)
    tc = 1 <= n <= 96
    return [f32(B, U), f32(B, U, n), f32(B, U, n), cw.new_empty((B, U, n), dtype=torch.uint8), f32(B, U),
            cw.new_empty((B, U, 4) if tc else (0,), dtype=torch.int32), f32(C), f32(C), f64(U, 4), f64(C, 4),
            f64(U, 4), f64(U, 4 * k), cw.new_empty((4 * k,), dtype=torch.int64),
            f32(0) if tc else f32(B, U, HIDDEN),
            cw.new_empty((_image_bytes(B, U, n) if tc else 0,), dtype=torch.uint8), f32(U), f32(U)]


def _grad_views(arena, offs, shapes):
    out, res = [], []
    for off, sh in zip(offs, shapes):
        numel = 1
        for d in sh:
            numel *= d
        out.append(arena[off:off + numel])
        res.append(arena[off:off + numel].view(sh))
    return out, res


def _bank_grad_shapes(cw, n):
    U, _, k = cw.shape
    C = U * HIDDEN
    return [(U, 4, k), (U,), (U,), (U,), (C, n), (C,), (C,), (C,), (U, HIDDEN), (U,), (U,), (U,)]


def _bank_bwd_into(flat, do, planes, L, params, saved, drop_p, seed, drop_mask, arena):
    """the backward of the unit bank, every parameter gradient written into `flat` (BANK_PARAMS order)"""
    cw, cb, g1, be1, w1, b1, g2, be2, w2, b2, g3, be3 = params
    (o, pool, zmax, arg, opre, bits, scale2, shift2, stats1, stats2, stats3, cov_w, cnt1, hpre, image, scale1,
     shift1) = saved
    B, U = o.shape
    k = cw.shape[2]
    C = U * HIDDEN
    sync, world = _sync["sum"], (_sync["world"] if _sync["sum"] else 1)
    n_pos = B * (L - k + 1) * world
    d_cw, d_cb, d_g1, d_be1, d_w1, d_b1, d_g2, d_be2, d_w2, d_b2, d_g3, d_be3 = flat
    _, _, dopre = ops.bn_bwd(do.contiguous(), o, opre, stats3, g3, inplace=False, out=[d_g3, d_be3], sync=sync,
                             world=world)
    act = {}
    if bits.numel():
        dy = ops.mlp_bwd(dopre, pool, w1, b1, scale2, shift2, stats2, g2, w2, bits, drop_p,
                         out=[d_w1, d_b1, d_g2, d_be2, d_w2, d_b2], sync=sync, world=world,
                         image=image if image.numel() else None)[6]
        if image.numel():        # on tile images the pass returns dL/dpool: the conv weight gradient applies exp'
            act = dict(act_scale=scale1, act_shift=shift1)
    else:
        dw2_, db2_, gh = ops.fc2_bwd(dopre, hpre, scale2, shift2, w2, drop_p, seed, drop_mask)
        d_w2.view(-1).copy_(dw2_.view(-1)); d_b2.copy_(db2_)
        _, _, dhpre = ops.bn_bwd(gh.view(B, C), None, hpre.view(B, C), stats2, g2, out=[d_g2, d_be2], sync=sync,
                                 world=world)
        dw1_, db1_, dy = ops.fc1_bwd(dhpre.view(B, U, HIDDEN), pool, w1)
        d_w1.view(-1).copy_(dw1_.view(-1)); d_b1.copy_(db1_)
    reserve = 0
    if arena is not None:
        for hook in _after_mlp_grads:
            reserve = max(reserve, int(hook(arena) or 0))
    # a hook that started a collective asks for a few SMs to be left out of the NEXT kernel's persistent grid (only
    # that one: every other kernel of the step keeps the whole chip)
    if reserve:
        ops.set_sm_reserve(reserve)
    try:
        ops.frontend_bwd(planes, L, k, dy, zmax, arg, g1, stats1, cov_w, cnt1, n_pos, out=[d_cw, d_cb, d_g1, d_be1],
                         **act)
    finally:
        if reserve:
            ops.set_sm_reserve(0)


@custom_op("explainn_b200::unit_bank_train_bwd", mutates_args=(), device_types="cuda")
def unit_bank_train_bwd(do: torch.Tensor, planes: torch.Tensor, L: int, params: Sequence[torch.Tensor],
                        saved: Sequence[torch.Tensor], drop_p: float, seed: int,
                        drop_mask: Optional[torch.Tensor]) -> List[torch.Tensor]:
    """the 12 parameter gradients in BANK_PARAMS order (functional form)"""
    shapes = _bank_grad_shapes(params[0], saved[1].shape[2])
    flat = [params[0].new_empty(sh) for sh in shapes]
    _bank_bwd_into(flat, do, planes, L, params, saved, drop_p, seed, drop_mask, None)
    return flat


@unit_bank_train_bwd.register_fake
def _(do, planes, L, params, saved, drop_p, seed, drop_mask):
    return [params[0].new_empty(sh) for sh in _bank_grad_shapes(params[0], saved[1].shape[2])]


@custom_op("explainn_b200::unit_bank_train_bwd_arena", mutates_args=("arena",), device_types="cuda")
def unit_bank_train_bwd_arena(do: torch.Tensor, planes: torch.Tensor, L: int, params: Sequence[torch.Tensor],
                              saved: Sequence[torch.Tensor], drop_p: float, seed: int,
                              drop_mask: Optional[torch.Tensor], arena: torch.Tensor, offs: Sequence[int]) -> None:
    """the same with the gradients written at arena[offs[i]:] (the autograd formula hands out views of them)"""
    flat, _ = _grad_views(arena, offs, _bank_grad_shapes(params[0], saved[1].shape[2]))
    _bank_bwd_into(flat, do, planes, L, params, saved, drop_p, seed, drop_mask, arena)


@unit_bank_train_bwd_arena.register_fake
def _(do, planes, L, params, saved, drop_p, seed, drop_mask, arena, offs):
    return None


def _bank_setup(ctx, inputs, output):
    planes, L, params, eps, drop_p, seed, seed_dev, drop_mask, arena, offs = inputs
    ctx.L, ctx.drop_p, ctx.seed, ctx.drop_mask, ctx.arena, ctx.offs = L, drop_p, seed, drop_mask, arena, list(offs)
    ctx.planes = planes
    ctx.n_params = len(params)
    # only `o` carries a gradient: without this autograd would ZERO-FILL a gradient for every other output of the
    # list (pool, zmax, the tile images ...: 1.5 GB of library fill kernels per step)
    ctx.set_materialize_grads(False)
    ctx.mark_non_differentiable(*output[1:])
    ctx.save_for_backward(*params, *output)


def _bank_backward(ctx, *grads):
    saved = ctx.saved_tensors
    params, outs = list(saved[:ctx.n_params]), list(saved[ctx.n_params:])
    do = grads[0][0] if isinstance(grads[0], (list, tuple)) else grads[0]
    if ctx.arena is not None:
        torch.ops.explainn_b200.unit_bank_train_bwd_arena(do, ctx.planes, ctx.L, params, outs, ctx.drop_p, ctx.seed,
                                                          ctx.drop_mask, ctx.arena, ctx.offs)
        _, res = _grad_views(ctx.arena, ctx.offs, _bank_grad_shapes(params[0], outs[1].shape[2]))
    else:
        res = torch.ops.explainn_b200.unit_bank_train_bwd(do, ctx.planes, ctx.L, params, outs, ctx.drop_p, ctx.seed,
                                                          ctx.drop_mask)
    pg = [g.view_as(p) for g, p in zip(res, params)]
    return None, None, pg, None, None, None, None, None, None, None


register_autograd("explainn_b200::unit_bank_train", _bank_backward, setup_context=_bank_setup)


# ------------------------------------------------------------------------------------------------------------
# a5: final Linear
# ------------------------------------------------------------------------------------------------------------
@custom_op("explainn_b200::linear", mutates_args=(), device_types="cuda")
def linear(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, arena: Optional[torch.Tensor],
           offs: Sequence[int]) -> torch.Tensor:
    return ops.linear_fwd(x, w, b)


@linear.register_fake
def _(x, w, b, arena, offs):
    return x.new_empty((x.shape[0], w.shape[0]))


@custom_op("explainn_b200::linear_bwd", mutates_args=(), device_types="cuda")
def linear_bwd(dy: torch.Tensor, w: torch.Tensor, x: torch.Tensor) -> List[torch.Tensor]:
    """[dx, dw, db] (functional form)"""
    return list(ops.linear_bwd(dy, w, x))


@linear_bwd.register_fake
def _(dy, w, x):
    return [x.new_empty(x.shape), w.new_empty(w.shape), w.new_empty((w.shape[0],))]


@custom_op("explainn_b200::linear_bwd_arena", mutates_args=("arena",), device_types="cuda")
def linear_bwd_arena(dy: torch.Tensor, w: torch.Tensor, x: torch.Tensor, arena: torch.Tensor,
                     offs: Sequence[int]) -> torch.Tensor:
    """dx; the weight / bias gradients are written at arena[offs[0]:], arena[offs[1]:]"""
    F, U = w.shape
    out = [arena[offs[0]:offs[0] + F * U], arena[offs[1]:offs[1] + F]]
    return ops.linear_bwd(dy, w, x, out=out)[0]


@linear_bwd_arena.register_fake
def _(dy, w, x, arena, offs):
    return x.new_empty(x.shape)


def _linear_setup(ctx, inputs, output):
    x, w, b, arena, offs = inputs
    ctx.arena, ctx.offs = arena, list(offs)
    ctx.save_for_backward(x, w)


def _linear_backward(ctx, dy):
    x, w = ctx.saved_tensors
    F, U = w.shape
    if ctx.arena is not None:
        a, o = ctx.arena, ctx.offs
        dx = torch.ops.explainn_b200.linear_bwd_arena(dy.contiguous(), w, x, a, o)
        return dx, a[o[0]:o[0] + F * U].view(F, U), a[o[1]:o[1] + F], None, None
    res = torch.ops.explainn_b200.linear_bwd(dy.contiguous(), w, x)
    return res[0], res[1], res[2], None, None


register_autograd("explainn_b200::linear", _linear_backward, setup_context=_linear_setup)


# ------------------------------------------------------------------------------------------------------------
# K12: loss
# ------------------------------------------------------------------------------------------------------------
@custom_op("explainn_b200::loss", mutates_args=(), device_types="cuda")
def loss(logits: torch.Tensor, target: torch.Tensor, kind: str) -> torch.Tensor:
    return ops.loss_fwd(logits, target, kind)


@loss.register_fake
def _(logits, target, kind):
    return logits.new_empty(())


@custom_op("explainn_b200::loss_bwd", mutates_args=(), device_types="cuda")
def loss_bwd(grad_out: torch.Tensor, logits: torch.Tensor, target: torch.Tensor, kind: str) -> torch.Tensor:
    return ops.loss_bwd(logits, target, grad_out, kind)


@loss_bwd.register_fake
def _(grad_out, logits, target, kind):
    return logits.new_empty(logits.shape)


def _loss_setup(ctx, inputs, output):
    logits, target, kind = inputs
    ctx.kind = kind
    ctx.save_for_backward(logits, target)


def _loss_backward(ctx, g):
    logits, target = ctx.saved_tensors
    return torch.ops.explainn_b200.loss_bwd(g.contiguous(), logits, target, ctx.kind), None, None


register_autograd("explainn_b200::loss", _loss_backward, setup_context=_loss_setup)


# ------------------------------------------------------------------------------------------------------------
# a11: PWM module
# ------------------------------------------------------------------------------------------------------------
@custom_op("explainn_b200::pwm_scan", mutates_args=(), device_types="cuda")
def pwm_scan(planes: torch.Tensor, L: int, pwm_w: torch.Tensor, mode: str, want_argmax: bool) -> List[torch.Tensor]:
    out, arg = ops.pwm_scan(planes, L, pwm_w, mode, want_argmax=want_argmax)
    return [out, arg] if want_argmax else [out]


@pwm_scan.register_fake
def _(planes, L, pwm_w, mode, want_argmax):
    out = pwm_w.new_empty((planes.shape[0], pwm_w.shape[0]))
    return [out, out.new_empty(out.shape, dtype=torch.int32)] if want_argmax else [out]
