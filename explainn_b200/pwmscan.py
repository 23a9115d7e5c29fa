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


def _fasta_lut():
    lut = np.full(256, -1, dtype=np.int8)
    for ch in range(128):
        if chr(ch).isalpha():
            lut[ch] = 4
    for i, ch in enumerate("ACGT"):
        lut[ord(ch)] = i
        lut[ord(ch.lower())] = i
    return lut


def read_fasta(path_or_lines):
    """pwm_scoring.c:575-673: header = first word after '>', letters upper-cased, A0 C1 G2 T3, any other
    letter 4, non-letters dropped, empty records skipped.  Returns (headers, codes int8, offsets int64).
    Vectorised over the whole file (a chromosome-sized FASTA is one LUT gather and two prefix sums, not a Python
    loop per line)."""
    if isinstance(path_or_lines, (str, bytes)) and not isinstance(path_or_lines, bytes):
        with open(path_or_lines, "rb") as fh:
            data = fh.read()
    elif hasattr(path_or_lines, "read"):
        data = path_or_lines.read()
    else:
        data = "".join(path_or_lines)
    if isinstance(data, str):
        data = data.encode("latin-1", "replace")
    buf = np.frombuffer(data, dtype=np.uint8)
    if buf.size == 0:
        return [], np.zeros(0, dtype=np.int8), np.zeros(1, dtype=np.int64)
    nl = np.flatnonzero(buf == 10)
    starts = np.concatenate(([0], nl + 1))
    starts = starts[starts < buf.size]
    ends = np.concatenate((nl, [buf.size]))[:starts.size]
    is_hdr = buf[starts] == ord(">")
    hs, he = starts[is_hdr], ends[is_hdr]
    if hs.size == 0:
        return [], np.zeros(0, dtype=np.int8), np.zeros(1, dtype=np.int64)
    mapped = _fasta_lut()[buf]
    delta = np.zeros(buf.size + 1, dtype=np.int8)
    delta[hs] += 1
    delta[he] -= 1
    mapped[np.cumsum(delta[:-1], dtype=np.int8) > 0] = -1        # header lines carry no sequence
    mapped[:hs[0]] = -1                                          # anything before the first header is ignored
    keep = mapped >= 0
    csum = np.concatenate(([0], np.cumsum(keep, dtype=np.int64)))
    bounds = csum[np.append(hs, buf.size)]
    lens = np.diff(bounds)
    headers = []
    for a, b, ln in zip(hs, he, lens):
        if ln > 0:
            parts = data[a + 1:b].decode("latin-1").split()
            headers.append(parts[0] if parts else "")
    offsets = np.concatenate(([0], np.cumsum(lens[lens > 0]))).astype(np.int64)
    return headers, mapped[keep].astype(np.int8), offsets


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


# ---------------------------------------------------------------------------------------------------------
# command line: the interface of the reference tool (pwm_scoring.c:715-1000), the only native boundary the
# reference has: argv in, one text line per FASTA record on stdout, exit code 0 / 1,
# errors on stderr.
# ---------------------------------------------------------------------------------------------------------
USAGE = """Usage: %s [options] -m <matrix_file> [<] <fasta_file>
   where options are:
     -b[--best]             Compute best single match scores
     -d[--debug]            Produce debugging output
     -h[--help]             Show this stuff
     -f[--forward]          Scan sequences in forward direction [def=bidirectional]
     -u[--unorm]            Normalize pwm scores by a uniform background base composition (Default=0.25)
     -p[--prob] <bg freq>   Normalize pwm scores by library-dependent nucleotide frequencies <bg freq>: 0.29,0.21,0.21,0.29
                            Note that nucleotide frequencies (<bg freq>) MUST BE comma-separated
     -q[--seqnorm]          Normalize pwm scores by sequence-based nucleotide composition
     -r[--nohdr]            Output raw scores (with no FASTA header)
     --lpm                  Input matrix is a letter probability matrix (LPM) [Default]
     --pwm                  Input matrix is a position weight matrix (PWM)
     -w[--pweight]          Set a pseudo-weight to re-normalize the frequencies of the letter-probability matrix (LPM)
                            Recommended value is 0.0001 [Default=0.0]

   Score a set of nucleotide sequences in FASTA format (<fasta_file>), based on matches to a sequence motif
   represented by an INTEGER position weight matrix [--pwm] or a base probability matrix [--lpm] (<matrix_file>).
   Note that the background normalization options (-u, -p, -q) are only valid for base probability matrices.
   For integer PWMs, only the best single match scores are reported, along with the position, strand, and sequence match.

"""


def main(argv=None, stdin=None, stdout=None, stderr=None):
    """`pwm_scoring [options] -m <matrix_file> [<] <fasta_file>` -> exit code.  Options, their precedence
    (--pwm switches the background normalisations off, :840-843; -u wins over -p, :881-903), messages and exit
    codes follow pwm_scoring.c:752-821,881-913,986-987; the scoring runs in the CUDA scan kernels."""
    import getopt
    import os
    import sys
    argv = list(sys.argv if argv is None else argv)
    stdin = stdin or sys.stdin
    stdout = stdout or sys.stdout
    stderr = stderr or sys.stderr
    prog, args = argv[0], argv[1:]
    longopts = ["best", "debug", "help", "forward", "matrix=", "prob=", "seqnorm", "unorm", "nohdr", "pweight=",
                "lpm", "pwm"]
    opts = []
    while True:                                        # getopt_long reports an unknown option and carries on
        try:
            opts, rest = getopt.gnu_getopt(args, "bdhfm:p:uqrw:", longopts)
            break
        except getopt.GetoptError as e:
            print(f"{prog}: {e.msg}", file=stderr)
            bad = [a for a in args if a.lstrip("-").split("=")[0] == e.opt and a.startswith("-")]
            if not bad:
                opts, rest = [], []
                break
            args.remove(bad[0])
    o = dict(best=False, debug=False, forward=False, matrix=None, prob=None, seqnorm=False, unorm=False, nohdr=False,
             pweight=0.0, pwm=False)
    for k, v in opts:
        if k in ("-b", "--best"): o["best"] = True
        elif k in ("-d", "--debug"): o["debug"] = True
        elif k in ("-f", "--forward"): o["forward"] = True
        elif k in ("-m", "--matrix"): o["matrix"] = v
        elif k in ("-p", "--prob"): o["prob"] = v
        elif k in ("-q", "--seqnorm"): o["seqnorm"] = True
        elif k in ("-u", "--unorm"): o["unorm"] = True
        elif k in ("-r", "--nohdr"): o["nohdr"] = True
        elif k in ("-w", "--pweight"):
            try:
                o["pweight"] = float(v)                # atof
            except ValueError:
                o["pweight"] = 0.0
        elif k == "--pwm": o["pwm"] = True
    if o["matrix"] is None:
        stderr.write(USAGE % prog)
        return 1
    try:
        matrix = read_matrix(o["matrix"], integer=o["pwm"])
    except OSError as e:
        print(f"Unable to open '{o['matrix']}': {e.strerror}({e.errno})", file=stderr)
        return 1
    except ValueError as e:
        print(f"Error: {e}", file=stderr)
        return 1
    prob = None
    if not o["pwm"] and not o["unorm"] and o["prob"] is not None:
        toks = [t for t in o["prob"].split(",") if t != ""]
        if len(toks) != 4:
            print(f"Number of TOKENS: {len(toks)}", file=stderr)
            print("Please, specify correct library-dependent nucleotide frequencies <bg freq>: they MUST BE "
                  "comma-separated!", file=stderr)
            return 1
        prob = toks
    if rest and rest[0] != "-":
        try:
            fh = open(rest[0], "rb")
        except OSError as e:
            print(f"Unable to open '{rest[0]}': {e.strerror}({e.errno})", file=stderr)
            return 1
        with fh:
            headers, codes, offsets = read_fasta(fh)
        if o["debug"]:
            print(f"Processing file {rest[0]}", file=stderr)
    else:
        headers, codes, offsets = read_fasta(getattr(stdin, "buffer", stdin))
    if o["debug"]:
        print(f"Motif length: {matrix.shape[1]}", file=stderr)
    if not headers:
        return 0
    if o["pwm"]:
        ps = PwmScoring(matrix, mode="pwm", forward=o["forward"])
    else:
        ps = PwmScoring(matrix, mode="lpm", best=o["best"], forward=o["forward"], unorm=o["unorm"], prob=prob,
                        seqnorm=o["seqnorm"], pweight=o["pweight"])
    if headers:
        res = ps.score(codes, offsets)
        for line in ps.format(headers, codes, offsets, res, nohdr=o["nohdr"]):
            stdout.write(line + "\n")
    return 0


if __name__ == "__main__":
    import sys
    sys.exit(main())
