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


def _chunk_of(planes, idx_dev_all, base, c0, n):
    """One chunk of the listed sequences for the chunked passes: without a list, a contiguous slice of the planes
    (no index array at all); with one, the device-resident list sliced and shifted by the strand's base.
    Returns (planes, idx or None)."""
    if idx_dev_all is None:
        return planes[base + c0:base + c0 + n], None
    return planes, (idx_dev_all[c0:c0 + n] + base) if base else idx_dev_all[c0:c0 + n].contiguous()


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
    idx_dev_all = None if idxs is None else torch.as_tensor(idx_all, dtype=torch.int32, device=dev)
    for strand in strands:
        for c0 in range(0, len(idx_all), chunk):
            pl, idx_dev = _chunk_of(packed.planes, idx_dev_all, strand * half, c0, min(chunk, len(idx_all) - c0))
            counts = ops.site_counts(pl, packed.length, cw, scale, shift, thr, idx_dev)
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
    idx_dev_all = None if idxs is None else torch.as_tensor(idx_all, dtype=torch.int32, device=dev)
    for strand in ((0, 1) if rev_complement else (0,)):
        if carry0 is not None:
            carry = torch.as_tensor(carry0[strand], dtype=torch.int64, device=dev).clone()
        for c0 in range(0, len(idx_all), chunk):
            ids = idx_all[c0:c0 + chunk]
            planes_c, idx_dev = _chunk_of(packed.planes, idx_dev_all, strand * half, c0, len(ids))
            counts = counts_cache.get((strand, c0)) if counts_cache is not None else None
            if counts is None:
                counts = ops.site_counts(planes_c, packed.length, cw, scale, shift, thr, idx_dev)
            sites = None
            if not return_sites:
                offsets = ops.site_offsets(counts, carry)              # carry advanced in place
                ops.site_emit(planes_c, packed.length, cw, scale, shift, thr, offsets, cap, idx_dev, pfm, None)
            else:
                csum = torch.cumsum(counts.to(torch.int64), dim=1)
                offsets = (csum - counts + carry[:, None]).contiguous()
                # rows are addressed by global rank: give the kernel a window starting at `carry`
                width = int(min(site_capacity, int(csum[:, -1].max().item())))
                sites = torch.full((U, max(width, 1), 2), -1, dtype=torch.int32, device=dev)
                local_off = (csum - counts).contiguous()
                local_cap = torch.clamp(cap - carry, min=0)
                # per-unit caps differ once carries differ: emit with offsets shifted so that rank < cap
                # is equivalent to local rank < cap - carry; done by passing global offsets and a
                # global cap for the PFM, and a second call layout for the site window.
                ops.site_emit(planes_c, packed.length, cw, scale, shift, thr, offsets, cap, idx_dev, pfm,
                              None)
                tmp_pfm = torch.zeros_like(pfm)
                ops.site_emit(planes_c, packed.length, cw, scale, shift, thr, local_off, 1 << 62, idx_dev,
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


# ---------------------------------------------------------------------------------------------------------
# row f3: what interpret.py does around the activations -- well-predicted selection (interpret.py:315-361), filter
# importances restricted to the sequences a filter fires on (:175-192, :485-490), the per-filter median table, and
# the MEME file of the filters (:259-274 -> utils/jaspar2others.py:46-83).
# ---------------------------------------------------------------------------------------------------------
def _fwd_rev(a, strand):
    """first / second half along axis 0 (test.py `__get_fwd_rev`: the data set is [forward ; reverse complement])"""
    half = len(a) // 2
    return a[:half] if strand == "fwd" else a[half:]


def well_predicted_sequences(preds, labels, input_data="binary", rev_complement=False):
    """interpret.py:315-361.  Binary: sequences whose every (sigmoid) prediction is on the label's side of 0.5.
    Otherwise: intersection of the top 5 % of the labels and of the predictions (flattened indices, as the reference
    computes them).  preds are the float16 arrays of acts_outs_preds; the arithmetic keeps the reference's dtypes
    (float16 sigmoid for single-strand input, float64 mean of the two strands otherwise)."""
    n = .05
    preds, labels = np.asarray(preds), np.asarray(labels)
    if rev_complement:
        fwd, rev = _fwd_rev(preds, "fwd"), _fwd_rev(preds, "rev")
        ys = _fwd_rev(labels, "fwd")
        p = np.empty(fwd.shape)
        for i in range(p.shape[1]):
            p[:, i] = np.mean([fwd[:, i], rev[:, i]], axis=0)
            if input_data == "binary":
                p[:, i] = torch.sigmoid(torch.from_numpy(p[:, i])).numpy()
    else:
        p = torch.sigmoid(torch.from_numpy(preds)).numpy() if input_data == "binary" else preds
        ys = labels
    if input_data == "binary":
        ok = (ys == (p > .5).astype(int))
        # the reference counts matches per row in first-seen order of np.where(...)[0] (ascending rows)
        return np.flatnonzero(ok.sum(axis=1) == ys.shape[1])
    k_y, k_p = int(max(ys.shape) * n), int(max(p.shape) * n)
    idxs_ys = np.argsort(-ys.flatten())[:k_y]
    idxs_p = np.argsort(-p.flatten())[:k_p]
    return np.intersect1d(idxs_ys, idxs_p)


def filtered_filter_importances(imps_u, idxs, acts_u, threshold):
    """interpret.py:485-490: importances (F, N) of ONE filter over the well-predicted sequences on which its
    activation exceeds the threshold anywhere ((N, L') fp16 activations; np.where(acts > thr)[0] may repeat a row,
    intersect1d makes the rows unique and sorted)"""
    rows = np.intersect1d(idxs, np.where(acts_u > threshold)[0])
    return imps_u[:, rows]


def filter_importance_table(outs, final_weight, idxs, fired):
    """The two tables of interpret.py:175-192 without the (N, U, L') activation array: `fired` (U, N) bool says on
    which sequences a filter exceeds its threshold (site counts > 0 of interpret.site_totals / sites_and_pfms).
    Returns (rows, medians): rows = [(filter index, importances (F,)) ...] in the reference's order (filter, then
    ascending sequence), medians (U, F) with NaN for filters that fire nowhere (pandas drops them)."""
    outs, W = np.asarray(outs), np.asarray(final_weight)
    U = W.shape[1]
    rows, med = [], np.full((U, W.shape[0]), np.nan)
    for u in range(U):
        sel = np.intersect1d(idxs, np.flatnonzero(fired[u]))
        if sel.size:
            imp = outs[sel, u][:, None] * W[:, u][None, :]          # (n_sel, F) = np.multiply(outs, weights[f])[:, u]
            rows.extend((u, r) for r in imp)
            med[u] = np.median(imp, axis=0)
    return rows, med


def pfms_to_meme(pfms, names=None, matrix_ids=None):
    """utils/jaspar2others.py:46-83 (`reformat_motif(..., "meme")`) for count matrices (U, 4, k): header, then per
    motif `MOTIF id name consensus`, `letter-probability matrix: alength= 4 w= k nsites= N E= 0` and the
    column-normalised counts printed as str(round(p, 8)).rjust(11).  Motifs without sites are skipped (the
    reference reads only the JASPAR files that exist).  The consensus and the probabilities come from Biopython's
    `pssm.consensus` / `pwm` there (counts / column sum, no pseudocounts; arg-max letter, first of A,C,G,T on
    ties): Biopython is absent here, so this writer is PARITY-UNPINNED beyond that published behaviour."""
    s = ["MEME version 4\n\n", "ALPHABET= ACGT\n\n", "strands: + -\n\n",
         "Background letter frequencies (from uniform background):\n", "A 0.25000 C 0.25000 G 0.25000 T 0.25000\n"]
    for u, pfm in enumerate(np.asarray(pfms, dtype=np.float64)):
        tot = pfm.sum(axis=0)
        if not (tot > 0).all():
            continue
        pwm = pfm / tot
        cons = "".join("ACGT"[int(np.argmax(col))] for col in pwm.T)
        mid = matrix_ids[u] if matrix_ids is not None else f"filter{u}"
        name = names[u] if names is not None else mid
        s.append("\nMOTIF %s %s %s\n" % (mid, name, cons))
        s.append("letter-probability matrix: alength= 4 w= %s nsites= %s E= 0\n" % (pwm.shape[1], int(pfm[:, 0].sum())))
        for row in pwm.T:
            s.append(" ".join(str(round(float(r), 8)).rjust(11) for r in row) + "\n")
    return "".join(s)
