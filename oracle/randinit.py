"""ORACLE (test infrastructure): numpy restatement of the deterministic random initialiser.

Checkpoints and Ollama pulls are unavailable offline (BASELINE.json north_star), so every model is
random-init.  The initialiser is integer-exact so that the host C++ generator, the CUDA generator
and this numpy restatement produce identical bytes:

    key      = splitmix64(seed * 0x100000001B3 + tensor_id)            (mod 2^64)
    h(i)     = splitmix64(key + i)                    i = row * cols + col
    S        = sum of the four 16-bit fields of h     (Irwin-Hall, integer)
    g        = float32(S - 131070) * float32(GSCALE)  ~ N(0, 1)
    value    = g * sigma (+ offset)                   (float32, one rounding per op)

Matrices use sigma 0.02 and are then quantised with the pinned ggml quantiser (oracle/quants.py).
"""
from __future__ import annotations

import numpy as np

from . import quants
from .archs import Arch, tensor_id, tensor_table

GSCALE = np.float32(2.6429157e-05)  # 1 / sqrt(4 * (65536^2 - 1) / 12)
MASK = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(z: np.ndarray) -> np.ndarray:
    z = np.asarray(z, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = z + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def tensor_key(seed: int, tid: int) -> np.uint64:
    v = (seed * 0x100000001B3 + tid) & 0xFFFFFFFFFFFFFFFF
    return splitmix64(np.array(v, dtype=np.uint64))


def gauss(seed: int, tid: int, n: int, start: int = 0) -> np.ndarray:
    """float32[n] of approximately N(0,1) values for elements start..start+n-1 of tensor `tid`."""
    key = tensor_key(seed, tid)
    with np.errstate(over="ignore"):
        h = splitmix64(key + np.arange(start, start + n, dtype=np.uint64))
    s = ((h & np.uint64(0xFFFF)) + ((h >> np.uint64(16)) & np.uint64(0xFFFF))
         + ((h >> np.uint64(32)) & np.uint64(0xFFFF)) + (h >> np.uint64(48))).astype(np.int64)
    return (s - 131070).astype(np.float32) * GSCALE


# initialisation recipe per tensor kind: (sigma, offset)
def recipe(a: Arch, kind: str):
    if kind in ("attn_norm", "ffn_norm", "output_norm"):
        return (np.float32(0.1), np.float32(0.0 if a.norm_plus_one else 1.0))
    if kind.endswith("_bias"):
        return (np.float32(0.1), np.float32(0.0))
    return (np.float32(0.02), np.float32(0.0))


def make_tensor(a: Arch, seed: int, kind: str, layer: int, rows: int, cols: int, qtype: str,
                row0: int = 0, nrows: int | None = None):
    """Returns float32[rows] for vectors, else the quantised blob uint8[nrows, row_bytes]."""
    sigma, off = recipe(a, kind)
    tid = tensor_id(kind, layer)
    if cols == 0:
        return (gauss(seed, tid, rows) * sigma + off).astype(np.float32)
    nrows = rows - row0 if nrows is None else nrows
    x = (gauss(seed, tid, nrows * cols, row0 * cols) * sigma).astype(np.float32).reshape(nrows, cols)
    if qtype == "f32":
        return x
    return quants.quantize(x, qtype)


def qtype_for(kind: str, qtype: str, head_qtype: str | None = None) -> str:
    if kind in ("output", "token_embd") and head_qtype:
        return head_qtype
    return qtype


def make_model(a: Arch, seed: int, qtype: str = "q4_0", head_qtype: str | None = None) -> dict:
    """name -> (qtype, ndarray).  Vectors are ('f32', float32[n])."""
    out = {}
    for name, kind, layer, rows, cols in tensor_table(a):
        qt = "f32" if cols == 0 else qtype_for(kind, qtype, head_qtype)
        out[name] = (qt, make_tensor(a, seed, kind, layer, rows, cols, qt))
    return out
