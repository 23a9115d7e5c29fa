"""Synthetic inputs of bench.py / bench_paths.py: iid uniform bases.  The product arms import their
generators from here, nothing from oracle/."""
import numpy as np
import torch

_LETTERS = np.frombuffer(b"ACGT", dtype=np.uint8)


def random_codes(B, L, seed=1234):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 4, (B, L), generator=g)


def onehot_from_codes(codes):
    """float32 one-hot (B,4,L), rows A,C,G,T -- what one_hot_encode (explainn/sequence/__init__.py:8-29) yields"""
    B, L = codes.shape
    x = torch.zeros(B, 4, L)
    x.scatter_(1, codes[:, None, :], 1.0)
    return x


def ascii_from_codes(codes):
    """uint8 (B,L) upper-case bases: the raw sequence column of the reference's TSV input (train.py:266-284)"""
    return torch.from_numpy(_LETTERS[codes.numpy()])


def random_onehot(B, L, seed=1234, n_frac=0.0):
    x = onehot_from_codes(random_codes(B, L, seed))
    if n_frac > 0:
        g = torch.Generator().manual_seed(seed + 77)
        x = x * (~(torch.rand(B, L, generator=g) < n_frac))[:, None, :].float()
    return x
