"""Data-parallel plumbing: one process per GPU over torch.distributed.

Training shards the batch across ranks; the only exchange is ONE all-reduce of a flat fp32 gradient
bucket per step (10.6 MB at cfg3 -- latency-bound on NVSwitch).  BatchNorm statistics are per rank
(the DDP default); inference, interpret and scans shard sequences with no collective beyond a final
gather.  The reference is single-device (selene/__init__.py:176-180): this module has no counterpart.
"""
import os

import torch
import torch.distributed as dist


def init_from_env(backend=None):
    """Initialise the default process group from RANK / WORLD_SIZE / MASTER_* (torchrun)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return rank, local, world


def shard_range(n, rank, world):
    """Contiguous ascending shard [lo, hi) of n items (interpret's site order needs contiguity)."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class GradBucket:
    """The gradient exchange of data-parallel training over the model's FLAT gradient arena
    (ExplaiNN.use_grad_arena): the backward kernels write every parameter gradient into one fp32 buffer, so the
    exchange is all-reduce(SUM) on that buffer in place -- no pack / unpack copies -- and the 1/world factor is folded
    into Adam's `grad_scale` (`scale`).

    overlap: the arena keeps the conv / BN1 gradients (the LAST ones the backward produces, 23 K values at cfg3)
    behind `model.arena_split`; everything before it (W1 = 9.96 of the 10.6 MB) is complete when xnn_mlp_bwd
    finishes.  That segment's all-reduce is issued from inside the backward (torch_ops._after_mlp_grads) on NCCL's
    stream, under xnn_frontend_bwd; `sm_reserve` SMs are left out of THAT kernel's persistent grid (and of no other
    kernel's) so that the NCCL kernel has somewhere to run.  Gradients that an autograd hook replaced (selene/__init__.py:254-257 freezes filters that way)
    do not live in the arena: they are copied in before their segment is reduced, and a hook on an MLP-segment
    parameter switches the early all-reduce off.

    Works for any list of parameters too (`GradBucket(params)`): then it is the plain pack -> all-reduce -> unpack."""

    def __init__(self, model_or_params, group=None, overlap=True, sm_reserve=4):
        self.group = group
        self.model = model_or_params if hasattr(model_or_params, "use_grad_arena") else None
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.scale = 1.0 / self.world
        self._early = None
        self.flat = None
        if self.model is None:
            self.params = [p for p in model_or_params if p.requires_grad]
            self.numel = sum(p.numel() for p in self.params)
            return
        m = self.model
        self.arena = m.use_grad_arena()
        names, offs, total, split = m._arena_plan()
        pd = dict(m.named_parameters())
        self.slots = [(pd[nm], off) for nm, off in zip(names, offs)]
        self.params = [p for p, _ in self.slots]
        self.split, self.numel = split, total
        self.overlap = bool(overlap) and self.world > 1 and split > 0
        self.sm_reserve = int(sm_reserve)
        if self.overlap:
            from . import torch_ops
            torch_ops._after_mlp_grads.append(self._after_mlp_grads)

    def close(self):
        if self.model is not None and self.overlap:
            from . import torch_ops
            if self._after_mlp_grads in torch_ops._after_mlp_grads:
                torch_ops._after_mlp_grads.remove(self._after_mlp_grads)

    # -- arena path ----------------------------------------------------------------------------------------
    def _after_mlp_grads(self, arena):
        """returns the number of SMs the conv weight gradient (the kernel this all-reduce runs under) should leave
        to the NCCL kernel"""
        if arena is not self.arena:
            return 0
        if any(p._backward_hooks for p, off in self.slots if off < self.split):
            return 0                                # a hook may still replace one of these gradients
        self._early = dist.all_reduce(self.arena[:self.split], op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        return self.sm_reserve

    def _adopt(self, lo, hi):
        """gradients of [lo, hi) that do not live in the arena (hooks, accumulation) are copied into it"""
        for p, off in self.slots:
            if not (lo <= off < hi):
                continue
            n = p.numel()
            view = self.arena[off:off + n]
            if p.grad is None:
                view.zero_()
            elif p.grad.data_ptr() != view.data_ptr():
                view.copy_(p.grad.reshape(-1))
            else:
                continue
            p.grad = view.view_as(p)

    def allreduce_sum(self):
        """After backward: all gradients summed over the ranks, in the arena; multiply by `scale` (Adam's
        grad_scale) for the mean."""
        if self.world == 1:
            return
        if self.model is None:
            return self._allreduce_list(mean=False)
        early, self._early = self._early, None
        if early is None:
            self._adopt(0, self.numel)
            dist.all_reduce(self.arena, op=dist.ReduceOp.SUM, group=self.group)
            return
        self._adopt(self.split, self.numel)
        dist.all_reduce(self.arena[self.split:], op=dist.ReduceOp.SUM, group=self.group)
        early.wait()

    def allreduce_mean(self, group=None):
        """gradients averaged over the ranks (the reference-style call: p.grad holds the mean afterwards)"""
        if self.world == 1:
            return
        if self.model is None:
            return self._allreduce_list(mean=True)
        self.allreduce_sum()
        self.arena.mul_(self.scale)

    # -- list path -----------------------------------------------------------------------------------------
    def _allreduce_list(self, mean):
        p0 = self.params[0]
        if self.flat is None or self.flat.device != p0.device:
            self.flat = torch.empty(self.numel, dtype=torch.float32, device=p0.device)
        off = 0
        for p in self.params:
            n = p.numel()
            if p.grad is None:
                self.flat[off:off + n].zero_()
            else:
                self.flat[off:off + n].copy_(p.grad.reshape(-1))
            off += n
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
        if mean:
            self.flat.mul_(1.0 / self.world)
        off = 0
        for p in self.params:
            n = p.numel()
            if p.grad is None:
                p.grad = self.flat[off:off + n].view_as(p).clone()
            else:
                p.grad.copy_(self.flat[off:off + n].view_as(p))
            off += n


def enable_sync_batchnorm(group=None):
    """Synchronised BatchNorm statistics (SURVEY.md 8e option (ii)): the three train-mode BatchNorm layers then
    normalise with the statistics of the GLOBAL batch -- the reference's single-process numerics at the same global
    batch (explainn/architectures/__init__.py:79,90,99; single device, selene/__init__.py:176-180).
      linears.1 : ONE int64 all-reduce of the co-occurrence counts (exact);
      linears.7 / .11 : fp64 (sum, sum of squares) partials all-reduced between the statistics kernels and their
                  finalize; the two BatchNorm-backward reductions likewise (torch_ops.unit_bank_train / _bwd).
    Every rank must run the same number of steps with at least one sample each."""
    from . import torch_ops
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        torch_ops._sync.update(sum=None, world=1)
        return
    torch_ops._sync.update(sum=lambda t: dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group), world=world)


def disable_sync_batchnorm():
    from . import torch_ops
    torch_ops._sync.update(sum=None, world=1)


def broadcast_parameters(model, src=0):
    if dist.is_initialized() and dist.get_world_size() > 1:
        for t in list(model.parameters()) + list(model.buffers()):
            dist.broadcast(t.data, src)


def gather_rows(local, n_total, rank, world):
    """Final gather of per-shard result rows (inference logits, scan scores) onto every rank."""
    if world == 1:
        return local
    sizes = [shard_range(n_total, r, world) for r in range(world)]
    mx = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad)
    return torch.cat([o[:hi - lo] for o, (lo, hi) in zip(out, sizes)])


def exclusive_carry(local_totals, group=None):
    """Exchange step of the sharded interpret pass 2.  local_totals int64 (S,U): this
    shard's uncapped site totals per strand and unit.  The reference counts sites strand by strand, then
    by ascending sequence (interpret.py:388-417), and shards are contiguous ascending ranges, so the
    global rank of this shard's first site on strand s is
        sum over all shards of strands < s  +  sum over lower ranks of strand s.
    Returns (carry0 (S,U), grand total (U,)), both int64 on local_totals' device."""
    if not (dist.is_initialized() and dist.get_world_size(group) > 1):
        before = torch.cumsum(local_totals, dim=0) - local_totals
        return before, local_totals.sum(dim=0)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    allt = [torch.empty_like(local_totals) for _ in range(world)]
    dist.all_gather(allt, local_totals.contiguous(), group=group)
    allt = torch.stack(allt)                                   # (world, S, U)
    per_strand = allt.sum(dim=0)                               # (S, U)
    strands_before = torch.cumsum(per_strand, dim=0) - per_strand
    ranks_before = allt[:rank].sum(dim=0) if rank > 0 else torch.zeros_like(local_totals)
    return strands_before + ranks_before, per_strand.sum(dim=0)


def allreduce_max_(t, group=None):
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return t


def allreduce_sum_(t, group=None):
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


def interpret_sharded(model, local_seqs, rev_complement=False, cap=None, group=None):
    """The interpret path (a7-a10) over a contiguous, ascending shard of the sequences per rank.
    local_seqs: this rank's PackedDNA / float one-hot (both halves when rev_complement).
    pass 1: local per-unit maxima -> all-reduce(max) -> the same fp16 thresholds on every rank;
    pass 2: local totals -> exclusive_carry -> capped local emission -> all-reduce(sum) of the PFMs.
    Returns dict(thresholds (U,) float16, counts (U,) int64, pfms (U,4,k) int64), identical on all ranks
    and identical to the single-process result."""
    import numpy as np
    from . import interpret as I
    cap = I.SITE_CAP if cap is None else cap
    packed = I._as_packed(local_seqs)
    cw, scale, shift = I._front_eval_affine(model)
    from . import ops
    N = len(packed)
    half = N // 2 if rev_complement else N
    amax = ops.act_max(packed.planes[:half], packed.length, cw, scale, shift)
    if rev_complement:
        amax = torch.maximum(amax, ops.act_max(packed.planes[half:2 * half], packed.length, cw, scale, shift))
    allreduce_max_(amax, group)
    thr = (0.5 * amax.cpu().numpy().astype(np.float16)).astype(np.float16)
    cache = {}                                                  # pass 2a's counts, reused by pass 2b
    tot = I.site_totals(model, packed, thr, None, rev_complement, keep_counts=cache)
    carry0, grand = exclusive_carry(tot, group)
    res = I.sites_and_pfms(model, packed, thr, None, rev_complement, cap=cap, carry0=carry0, counts_cache=cache)
    pfm = torch.as_tensor(res["pfms"], device=packed.planes.device)
    allreduce_sum_(pfm, group)
    return {"thresholds": thr, "counts": torch.clamp(grand, max=cap).cpu().numpy(), "pfms": pfm.cpu().numpy()}
