"""Host side of explainn/pwm-scoring.py (a12): MEME parsing, PWM resizing, log transform, and the
batch scoring loop over the fused PWM scan kernel."""
import gzip
import re

import numpy as np
import torch

from .architectures import PWM

_ROW = re.compile(r"^\s*(\S+)\s+(\S+)\s+(\S+)\s+(\S+)$")


def _open(path):
    return gzip.open(path, "rt") if str(path).endswith(".gz") else open(path, "rt")


def resize_pwm(columns, size):
    """pwm-scoring.py:190-213.  Pads `size` uniform columns on both sides, then trims one column at a
    time from the end whose maximum probability is lower (ties: the side trimmed less often so far,
    left first) until `size` columns remain."""
    pad = [[.25, .25, .25, .25]]
    cols = pad * size + [list(c) for c in columns] + pad * size
    lo, hi = 0, len(cols)
    lpop = rpop = 0
    while hi - lo > size:
        left, right = max(cols[lo]), max(cols[hi - 1])
        if left < right or (left == right and lpop <= rpop):
            lo += 1
            lpop += 1
        else:
            hi -= 1
            rpop += 1
    return cols[lo:hi]


def get_pwms(meme_file, resize_pwms=False, return_log=False):
    """pwm-scoring.py:137-188: every probability is floored at 1e-4 (no renormalisation);
    returns (array (G,4,w) or nested lists, names)."""
    motifs, names = [], []
    with _open(meme_file) as fh:
        for line in fh:
            line = line.strip("\n")
            if line.startswith("MOTIF"):
                names.append(line.split(" ")[1])
                motifs.append([])
            elif not motifs or line.startswith("letter-probability matrix:"):
                continue
            else:
                m = _ROW.search(line)
                if m:
                    motifs[-1].append([max(1e-4, float(m.group(i + 1))) for i in range(4)])
    max_size = max((len(m) for m in motifs), default=0)
    pwms = []
    for cols in motifs:
        if resize_pwms:
            cols = resize_pwm(cols, max_size)
        pwms.append([list(r) for r in zip(*cols)])  # (4, w)
    if return_log:
        return np.log(pwms), names
    return pwms, names


# reference-compatible aliases (pwm-scoring.py:137,190)
_get_PWMs = get_pwms


def score_sequences(pwms, x, scoring="max", batch_size=4096, device="cuda"):
    """pwm-scoring.py:81-105: (N,4,L) one-hot -> (N,G) float64 score matrix."""
    x = torch.as_tensor(x, dtype=torch.float32)
    model = PWM(np.asarray(pwms), x.shape[2], scoring).to(device)
    scores = np.zeros((x.shape[0], model._options["groups"]))
    with torch.no_grad():
        for i in range(0, x.shape[0], batch_size):
            xb = x[i:i + batch_size].to(device, non_blocking=True)
            scores[i:i + xb.shape[0]] = model(xb).cpu().numpy()
    return scores
