"""Host-side mirror of explainn/sequence/__init__.py (the input contract, a16) plus the
device packers that replace its per-character Python loop on the hot path."""
import numpy as np
import torch

from . import ops

_COMPLEMENT = bytes.maketrans(b"ACGTacgtNn", b"TGCAtgcaNn")


def one_hot_encode(seq):
    """(4, len) float array, rows A,C,G,T; any other letter is an all-zero column
    (sequence/__init__.py:8-29).  Vectorised host version."""
    b = np.frombuffer(seq.upper().encode("ascii", "replace"), dtype=np.uint8)
    out = np.zeros((4, b.size), dtype=float)
    for row, letter in enumerate(b"ACGT"):
        out[row, b == letter] = 1.0
    return out


def one_hot_encode_many(seqs):
    return np.array([one_hot_encode(s) for s in seqs])


def one_hot_decode(encoded_seq):
    """sequence/__init__.py:35-49: columns that are not one-hot decode to N."""
    enc = np.asarray(encoded_seq)
    letters = np.array(list("ACGT"))
    idx = enc.argmax(axis=0)
    ok = (enc == 1).sum(axis=0) == 1
    return "".join(np.where(ok, letters[idx], "N"))


def one_hot_decode_many(seqs):
    return np.array([one_hot_decode(s) for s in seqs])


def rc_one_hot_encoding(encoded_seq):
    """sequence/__init__.py:58-60"""
    return encoded_seq[::-1, ::-1]


def rc_one_hot_encoding_many(arr):
    return np.array([rc_one_hot_encoding(e) for e in arr])


def rc(seq):
    """sequence/__init__.py:66-68 (Biopython reverse_complement for plain DNA + N)"""
    return seq.encode("ascii").translate(_COMPLEMENT)[::-1].decode("ascii")


def rc_many(arr):
    return np.array([rc(s) for s in arr])


def pack_strings(seqs, device="cuda"):
    """list of equal-length DNA strings -> PackedDNA on the device (ASCII -> planes kernel)."""
    from .architectures import PackedDNA
    L = len(seqs[0])
    if any(len(s) != L for s in seqs):
        raise ValueError("all sequences must have the same length")
    host = np.frombuffer("".join(seqs).encode("ascii", "replace"), dtype=np.uint8).reshape(len(seqs), L)
    dev = torch.from_numpy(host.copy()).to(device, non_blocking=True)
    return PackedDNA(ops.pack_ascii(dev), L)
