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
    """Flat fp32 bucket over the gradients of `params`: pack -> one all-reduce -> unpack (mean)."""

    def __init__(self, params):
        self.params = [p for p in params if p.requires_grad]
        self.numel = sum(p.numel() for p in self.params)
        self.flat = None

    def allreduce_mean(self, group=None):
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        if world == 1:
            return
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
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
        self.flat.mul_(1.0 / world)
        off = 0
        for p in self.params:
            n = p.numel()
            if p.grad is None:
                p.grad = self.flat[off:off + n].view_as(p).clone()
            else:
                p.grad.copy_(self.flat[off:off + n].view_as(p))
            off += n


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
    idx = torch.arange(half, dtype=torch.int32, device=packed.planes.device)
    amax = ops.act_max(packed.planes, packed.length, cw, scale, shift, idx)
    if rev_complement:
        amax = torch.maximum(amax, ops.act_max(packed.planes, packed.length, cw, scale, shift, idx + half))
    allreduce_max_(amax, group)
    thr = (0.5 * amax.cpu().numpy().astype(np.float16)).astype(np.float16)
    cache = {}                                                  # pass 2a's counts, reused by pass 2b
    tot = I.site_totals(model, packed, thr, None, rev_complement, keep_counts=cache)
    carry0, grand = exclusive_carry(tot, group)
    res = I.sites_and_pfms(model, packed, thr, None, rev_complement, cap=cap, carry0=carry0, counts_cache=cache)
    pfm = torch.as_tensor(res["pfms"], device=packed.planes.device)
    allreduce_sum_(pfm, group)
    return {"thresholds": thr, "counts": torch.clamp(grand, max=cap).cpu().numpy(), "pfms": pfm.cpu().numpy()}
