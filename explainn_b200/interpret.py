"""The interpret path (a7-a10) without the (N,U,Lc) fp16 activation array.

Reference flow (explainn/interpret.py:164-219): `_get_acts_outs_preds` (test.py:128-166)
materialises every unit activation as np.float16 on the host, `_get_act_thresholds`
(:363-373) takes half of each unit's maximum over the well-predicted sequences,
`_get_sites` (:375-421) lists every (sequence, start) above it -- at most 1e6 per unit, in
(strand, listed sequence, ascending start) order -- and `_sites_to_motif` (:423-454) counts
A/C/G/T per column of the listed k-mers.

Here the sequences stay packed on the device and three kernel passes recompute the
activations (rounded to fp16 exactly like the reference's stored array):
  pass 1  xnn_act_max      per-unit maximum            -> thresholds (fp16)
  pass 2a xnn_site_counts  per (unit, sequence) counts -> exclusive offsets (order + cap)
  pass 2b xnn_site_emit    PFM counts (+ optional site index list)
With `rev_complement` the data set is [forward ; reverse-complement] as built by
train.py:277-280 and `idxs` index the forward half (interpret.py:366-369,388-391).
"""
import numpy as np
import torch

from . import ops
from .architectures import PackedDNA, _check_param

SITE_CAP = 1000000  # interpret.py:416


def _front_eval_affine(model):
    conv, bn1 = model.linears[0], model.linears[1]
    cw = _check_param(conv.weight, "linears.0.weight").detach()
    scale, shift = ops.bn_fold_eval(bn1.weight.detach(), bn1.bias.detach(), bn1.running_mean, bn1.running_var,
                                    conv.bias.detach(), bn1.eps)
    return cw, scale, shift


def _as_packed(seqs):
    if isinstance(seqs, PackedDNA):
        return seqs
    return PackedDNA.from_onehot(seqs)


def act_thresholds(model, seqs, idxs=None, rev_complement=False):
    """interpret.py:363-373 -> np.float16 (U,).  `seqs` holds all sequences (both halves when
    rev_complement) as PackedDNA or a float one-hot CUDA tensor."""
    packed = _as_packed(seqs)
    cw, scale, shift = _front_eval_affine(model)
    N = len(packed)
    dev = packed.planes.device
    if idxs is None:
        idx_f = torch.arange(N // 2 if rev_complement else N, dtype=torch.int32, device=dev)
    else:
        idx_f = torch.as_tensor(np.asarray(idxs), dtype=torch.int32, device=dev)
    if idx_f.numel() == 0:
        raise ValueError("zero-size selection: no sequences to take the maximum over")
    amax = ops.act_max(packed.planes, packed.length, cw, scale, shift, idx_f)
    if rev_complement:
        amax = torch.maximum(amax, ops.act_max(packed.planes, packed.length, cw, scale, shift, idx_f + N // 2))
    # np.float16 arithmetic exactly as `0.5 * np.amax(acts_idxs, axis=(0, 2))` on a float16 array
    return (0.5 * amax.cpu().numpy().astype(np.float16)).astype(np.float16)


def site_totals(model, seqs, thresholds, idxs=None, rev_complement=False, chunk=65536, keep_counts=None):
    """Pass 2a alone: uncapped number of above-threshold positions per (strand, unit) -> int64 (S,U)
    device tensor.  A sequence shard needs the totals of the shards before it (parallel.exclusive_carry)
    to place the reference's 1e6 cap (interpret.py:414-417) on the same global prefix.  keep_counts: a dict that
    receives the per-chunk count matrices, to be handed to sites_and_pfms(counts_cache=...) with the same chunk."""
    packed = _as_packed(seqs)
    cw, scale, shift = _front_eval_affine(model)
    U = cw.shape[0]
    N = len(packed)
    dev = packed.planes.device
    half = N // 2 if rev_complement else N
    idx_all = np.arange(half) if idxs is None else np.asarray(idxs)
    thr = torch.as_tensor(np.asarray(thresholds, dtype=np.float16).astype(np.float32), device=dev)
    strands = (0, 1) if rev_complement else (0,)
    tot = torch.zeros((len(strands), U), dtype=torch.int64, device=dev)
    for strand in strands:
        for c0 in range(0, len(idx_all), chunk):
            idx_dev = torch.as_tensor(idx_all[c0:c0 + chunk] + strand * half, dtype=torch.int32, device=dev)
            counts = ops.site_counts(packed.planes, packed.length, cw, scale, shift, thr, idx_dev)
            if keep_counts is not None:
                keep_counts[(strand, c0)] = counts         # (U, chunk) int32: 4*U bytes per listed sequence
            tot[strand] += counts.sum(dim=1, dtype=torch.int64)
    return tot


def sites_and_pfms(model, seqs, thresholds, idxs=None, rev_complement=False, cap=SITE_CAP,
                   return_sites=False, site_capacity=None, chunk=65536, carry0=None, counts_cache=None):
    """interpret.py:375-454.  Returns dict(counts (U,) int64 sites emitted per unit,
    pfms (U,4,k) int64, sites: list per unit of int32 arrays (m,3) = (sequence index, strand 0/1, start)
    when return_sites).  carry0 (S,U) int64: global rank of this shard's first site per strand and unit
    (sequence-sharded runs, parallel.interpret_sharded); counts are then the ranks reached, capped."""
    packed = _as_packed(seqs)
    cw, scale, shift = _front_eval_affine(model)
    U, _, k = cw.shape
    N = len(packed)
    dev = packed.planes.device
    half = N // 2 if rev_complement else N
    idx_all = np.arange(half) if idxs is None else np.asarray(idxs)
    thr = torch.as_tensor(np.asarray(thresholds, dtype=np.float16).astype(np.float32), device=dev)
    pfm = torch.zeros((U, 4, k), dtype=torch.int32, device=dev)
    carry = torch.zeros(U, dtype=torch.int64, device=dev)
    site_chunks = [[] for _ in range(U)] if return_sites else None
    if return_sites and site_capacity is None:
        site_capacity = cap
    for strand in ((0, 1) if rev_complement else (0,)):
        if carry0 is not None:
            carry = torch.as_tensor(carry0[strand], dtype=torch.int64, device=dev).clone()
        for c0 in range(0, len(idx_all), chunk):
            ids = idx_all[c0:c0 + chunk]
            idx_dev = torch.as_tensor(ids + strand * half, dtype=torch.int32, device=dev)
            counts = counts_cache.get((strand, c0)) if counts_cache is not None else None
            if counts is None:
                counts = ops.site_counts(packed.planes, packed.length, cw, scale, shift, thr, idx_dev)
            csum = torch.cumsum(counts.to(torch.int64), dim=1)
            offsets = (csum - counts + carry[:, None]).contiguous()
            sites = None
            if return_sites:
                # rows are addressed by global rank: give the kernel a window starting at `carry`
                width = int(min(site_capacity, int(csum[:, -1].max().item())))
                sites = torch.full((U, max(width, 1), 2), -1, dtype=torch.int32, device=dev)
                local_off = (csum - counts).contiguous()
                local_cap = torch.clamp(cap - carry, min=0)
                # per-unit caps differ once carries differ: emit with offsets shifted so that rank < cap
                # is equivalent to local rank < cap - carry; done by passing global offsets and a
                # global cap for the PFM, and a second call layout for the site window.
                ops.site_emit(packed.planes, packed.length, cw, scale, shift, thr, offsets, cap, idx_dev, pfm,
                              None)
                tmp_pfm = torch.zeros_like(pfm)
                ops.site_emit(packed.planes, packed.length, cw, scale, shift, thr, local_off, 1 << 62, idx_dev,
                              tmp_pfm, sites)
                sites_h = sites.cpu().numpy()
                lc = local_cap.cpu().numpy()
                tot = csum[:, -1].cpu().numpy()
                for u in range(U):
                    m = int(min(tot[u], lc[u], sites_h.shape[1]))
                    if m > 0:
                        s = sites_h[u, :m]
                        rec = np.empty((m, 3), dtype=np.int64)
                        rec[:, 0] = ids[s[:, 0]]
                        rec[:, 1] = strand
                        rec[:, 2] = s[:, 1]
                        site_chunks[u].append(rec)
            else:
                ops.site_emit(packed.planes, packed.length, cw, scale, shift, thr, offsets, cap, idx_dev, pfm, None)
            carry = carry + csum[:, -1]
            if carry0 is None and bool((carry >= cap).all()):
                break
    out = {
        "counts": torch.clamp(carry, max=cap).cpu().numpy(),
        "pfms": pfm.cpu().numpy().astype(np.int64),
    }
    if return_sites:
        out["sites"] = [np.concatenate(c) if c else np.zeros((0, 3), dtype=np.int64) for c in site_chunks]
    return out


def acts_outs_preds(model, seqs, batch_size=1024):
    """Drop-in for test.py:128-166: np.float16 activations (N,U,Lc), unit outputs (N,U) and
    predictions (N,F), computed by the fused kernels batch by batch (eval mode)."""
    packed = _as_packed(seqs)
    o = model._options
    N, U, Lc, F = len(packed), o["cnn_units"], o["sequence_length"] - o["kernel_size"] + 1, o["n_features"]
    acts = np.zeros((N, U, Lc), dtype=np.float16)
    outs = np.zeros((N, U), dtype=np.float16)
    preds = np.zeros((N, F), dtype=np.float16)
    cw, scale, shift = _front_eval_affine(model)
    was_training = model.training
    model.eval()
    with torch.no_grad():
        for i in range(0, N, batch_size):
            pb = PackedDNA(packed.planes[i:i + batch_size].contiguous(), packed.length)
            uo = model.linears(pb)
            outs[i:i + len(pb)] = uo.cpu().numpy()
            preds[i:i + len(pb)] = model.final(uo).cpu().numpy()
            acts[i:i + len(pb)] = ops.frontend_acts(pb.planes, pb.length, cw, scale, shift, f32=False,
                                                    f16=True)[1].cpu().numpy()
    model.train(was_training)
    return acts, outs, preds


def filter_importances(outs, final_weight):
    """interpret.py:180-181: importance[f, i, u] = unit output * final-layer weight."""
    return np.asarray(outs)[None, :, :] * np.asarray(final_weight)[:, None, :]


def sites_to_fasta(ids, seq_strings, sites, kernel_size):
    """interpret.py:404-412 site records: `>{id}_{strand}_from={start}_to={end}` + the k-mer."""
    lines = []
    half = len(seq_strings) // 2
    for i, strand, start in sites:
        s = seq_strings[i + (half if strand else 0)] if strand else seq_strings[i]
        lines.append(f">{ids[i]}_{'-' if strand else '+'}_from={start}_to={start + kernel_size}\n"
                     f"{s[start:start + kernel_size]}\n")
    return "".join(lines)


def pfm_to_jaspar(pfm, matrix_id, name):
    """JASPAR text as Biopython writes it (interpret.py:209-216; fixtures under
    results/Human-Mouse-Project/CAM/motifs)."""
    out = [f">{matrix_id} {name}\n"]
    for letter, row in zip("ACGT", np.asarray(pfm)):
        out.append("%s [%s]\n" % (letter, " ".join("%6.2f" % v for v in row)))
    return "".join(out)
