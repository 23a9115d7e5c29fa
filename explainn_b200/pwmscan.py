"""Host side of the PWMScan `pwm_scoring` tool (explainn/pwmscan/pwm_scoring.c) over the CUDA scan
kernels: matrix and FASTA readers, option handling (-b -f -u -p -q -r -w --lpm --pwm, :752-798,
:859-903) and the tool's text output (:441-444, :470-473, :561-564)."""
import numpy as np
import torch

from . import ops

INT_MIN = -2 ** 31
_NUC = "ACGTN"


def read_matrix(path, integer=False):
    """pwm_scoring.c:91-338: one matrix row per line (A C G T); lines starting with # or > are skipped.
    Returns (4, w) float64 or int32."""
    rows = []
    with open(path) as fh:
        for line in fh:
            if line.startswith("#") or line.startswith(">") or not line.strip():
                continue
            vals = line.split()[:4]
            if len(vals) < 4:
                raise ValueError(f"matrix row {len(rows)} has fewer than 4 values")
            rows.append([int(v) for v in vals] if integer else [float(v) for v in vals])
    if not rows:
        raise ValueError("empty matrix")
    return np.array(rows, dtype=np.int32 if integer else np.float64).T.copy()


def read_fasta(path_or_lines):
    """pwm_scoring.c:575-673: header = first word after '>', letters upper-cased, A0 C1 G2 T3, any other
    letter 4, non-letters dropped, empty records skipped.  Returns (headers, codes int8, offsets int64)."""
    lines = open(path_or_lines) if isinstance(path_or_lines, str) else path_or_lines
    lut = np.full(256, -1, dtype=np.int8)
    for ch in range(256):
        if chr(ch).isalpha() and ch < 128:
            lut[ch] = 4
    for i, ch in enumerate("ACGT"):
        lut[ord(ch)] = i
        lut[ord(ch.lower())] = i
    headers, chunks, offsets = [], [], [0]
    cur, hdr = None, None

    def flush():
        if hdr is not None and cur:
            rec = np.concatenate(cur) if len(cur) > 1 else cur[0]
            if rec.size:
                headers.append(hdr)
                chunks.append(rec)
                offsets.append(offsets[-1] + rec.size)

    for line in lines:
        if line.startswith(">"):
            flush()
            parts = line[1:].split()
            hdr = parts[0] if parts else ""
            cur = []
        elif hdr is not None:
            b = lut[np.frombuffer(line.encode("latin-1", "replace"), dtype=np.uint8)]
            cur.append(b[b >= 0])
    flush()
    if hasattr(lines, "close") and isinstance(path_or_lines, str):
        lines.close()
    codes = np.concatenate(chunks) if chunks else np.zeros(0, dtype=np.int8)
    return headers, codes.astype(np.int8), np.array(offsets, dtype=np.int64)


class PwmScoring:
    """One matrix, many records.  Options as the CLI: best (-b), forward (-f), unorm (-u),
    prob (-p "a,c,g,t"), seqnorm (-q), pweight (-w), mode 'lpm' (default) or 'pwm'."""

    def __init__(self, matrix, mode="lpm", best=False, forward=False, unorm=False, prob=None, seqnorm=False,
                 pweight=0.0, device="cuda", tie_cap=64):
        self.mode, self.best, self.forward, self.seqnorm = mode, best, forward, seqnorm
        self.tie_cap = tie_cap  # the C tool keeps tied positions in a 256-byte buffer (BEST_HIT_POS)
        self.device = device
        m = np.asarray(matrix)
        self.w = m.shape[1]
        if mode == "lpm":
            lpm = np.empty((5, self.w), dtype=np.float64)
            lpm[:4] = m
            lpm[4] = 0.25                                            # :862-865
            if pweight != 0.0:                                       # :866-875
                tot = np.zeros(self.w)
                for i in range(4):                                   # same accumulation order as the C loop
                    tot = tot + (lpm[i] + pweight)
                lpm[:4] = (lpm[:4] + pweight) / tot
            bg = np.array([1.0, 1.0, 1.0, 1.0, 0.25])                # :74
            if unorm:
                bg[:4] = 0.25                                        # :881-886
            elif prob is not None:
                vals = [float(v) for v in (prob.split(",") if isinstance(prob, str) else prob)]
                if len(vals) != 4:
                    raise ValueError("specify four comma-separated background frequencies")
                bg[:4] = vals
            self.lpm = torch.tensor(lpm, dtype=torch.float64, device=device)
            self.bg = bg
        else:
            pwm = np.empty((5, self.w), dtype=np.int64)
            pwm[:4] = m
            pwm[4] = INT_MIN                                         # :877-879
            self.pwm = torch.tensor(pwm.astype(np.int32), dtype=torch.int32, device=device)

    def _seq_bg(self, codes, offsets):
        """-q: per-record background (:355-372); integer division of (A+T)/2 as in the C expression."""
        R = len(offsets) - 1
        bg = np.tile(self.bg, (R, 1))
        for r in range(R):
            seg = codes[offsets[r]:offsets[r + 1]]
            cnt = np.bincount(seg, minlength=5)
            n = float(seg.size)
            if self.forward:
                bg[r, :4] = cnt[:4] / n
            else:
                at = (float((cnt[0] + cnt[3]) // 2) + cnt[4] / 4.0) / n
                bg[r, 0] = bg[r, 3] = at
                bg[r, 1] = bg[r, 2] = 0.5 - at
        return bg

    def score(self, codes, offsets):
        """codes int8 (numpy or CUDA tensor), offsets int64 -> dict of numpy arrays."""
        codes_h = codes if isinstance(codes, np.ndarray) else None
        cd = torch.as_tensor(codes, dtype=torch.int8).to(self.device)
        od = torch.as_tensor(offsets, dtype=torch.int64).to(self.device)
        if self.mode == "pwm":
            best, pos, strand = ops.ipwm_scan_best(cd, od, self.pwm, self.forward)
            return {"best": best.cpu().numpy(), "pos": pos.cpu().numpy(), "strand": strand.cpu().numpy()}
        if self.seqnorm:
            if codes_h is None:
                codes_h = cd.cpu().numpy()
            bg = torch.tensor(self._seq_bg(codes_h, np.asarray(offsets)), dtype=torch.float64, device=self.device)
        else:
            bg = torch.tensor(self.bg, dtype=torch.float64, device=self.device)
        if not self.best:
            return {"sum": ops.lpm_scan_sum(cd, od, self.lpm, bg, self.forward).cpu().numpy()}
        best, pos, strand, nties, ties = ops.lpm_scan_best(cd, od, self.lpm, bg, self.forward, self.tie_cap)
        return {"best": best.cpu().numpy(), "pos": pos.cpu().numpy(), "strand": strand.cpu().numpy(),
                "n_ties": nties.cpu().numpy(), "ties": ties.cpu().numpy()}

    def format(self, headers, codes, offsets, result, nohdr=False):
        """The tool's stdout, one line per record."""
        out = []
        lens = np.diff(np.asarray(offsets))
        for r, h in enumerate(headers):
            pre = "" if nohdr else f"{h}\t"
            if self.mode == "pwm":
                if lens[r] < self.w:
                    out.append(f"{pre}0\t0\tNOTAG\t{INT_MIN}\t0")
                    continue
                p, st = int(result["pos"][r]), int(result["strand"][r])
                word = np.asarray(codes[offsets[r] + p: offsets[r] + p + self.w])
                if st:
                    word = np.where(word == 4, 4, 3 - word)[::-1]
                tag = "".join(_NUC[c] for c in word)
                out.append(f"{pre}{p}\t{p + self.w}\t{tag}\t{int(result['best'][r])}\t{'-' if st else '+'}")
            elif self.best:
                plist = [str(int(result["pos"][r]))]
                nt = int(result["n_ties"][r])
                plist += [str(int(v)) for v in result["ties"][r][:nt]]
                out.append("%s%g\t%d\t%s\t%s" % (pre, result["best"][r], lens[r], ",".join(plist),
                                                  chr(int(result["strand"][r]))))
            else:
                out.append("%s%g" % (pre, result["sum"][r]))
        return out
