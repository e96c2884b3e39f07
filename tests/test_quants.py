"""Oracle quant formats vs the committed gguf-py golden vectors (tests/golden/make_quant_golden.py)."""
import os

import numpy as np
import pytest

from oracle import quants as Q

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "quant_golden.npz"))


def test_q4_0_quantize_matches_gguf():
    assert np.array_equal(Q.quantize_q4_0(G["x"]), G["q4_0"])


def test_q4_0_dequantize_matches_gguf():
    assert np.array_equal(Q.dequantize_q4_0(G["q4_0"]), G["dq4_0"])
    assert np.array_equal(Q.dequantize_q4_0(G["b40"]), G["dqb40"])


def test_q8_0_matches_gguf():
    assert np.array_equal(Q.quantize_q8_0(G["x"]), G["q8_0"])
    assert np.array_equal(Q.dequantize_q8_0(G["q8_0"]), G["dq8_0"])


def test_k_quant_dequantize_matches_gguf():
    assert np.array_equal(Q.dequantize_q4_k(G["b4k"]), G["dq4_k"])
    assert np.array_equal(Q.dequantize_q6_k(G["b6k"]), G["dq6_k"])


@pytest.mark.parametrize("qt,bound", [("q4_0", 0.12), ("q4_k", 0.10), ("q6_k", 0.03), ("q8_0", 0.01)])
def test_round_trip_error_bounds(qt, bound):
    x = np.random.default_rng(3).standard_normal((16, 1024)).astype(np.float32) * 0.02
    blob = Q.quantize(x, qt)
    assert blob.shape == (16, Q.row_bytes(1024, qt))
    y = Q.dequantize(blob, qt)
    assert np.sqrt(((y - x) ** 2).mean()) / x.std() < bound
    # idempotence: re-quantising the dequantised values of the affine-free formats reproduces the bytes
    if qt in ("q4_0", "q8_0"):
        assert np.array_equal(Q.dequantize(Q.quantize(y, qt), qt), y)


def test_k_quant_scale_packing_round_trip():
    rng = np.random.default_rng(5)
    s = rng.integers(0, 64, (9, 8), dtype=np.uint8)
    m = rng.integers(0, 64, (9, 8), dtype=np.uint8)
    s2, m2 = Q._unpack_scale_min_k4(Q._pack_scale_min_k4(s, m))
    assert np.array_equal(s, s2) and np.array_equal(m, m2)


def test_zero_and_edge_blocks():
    z = np.zeros((1, 256), dtype=np.float32)
    for qt in ("q4_0", "q8_0", "q4_k", "q6_k"):
        assert np.all(Q.dequantize(Q.quantize(z, qt), qt) == 0)
    with pytest.raises(AssertionError):
        Q.row_bytes(100, "q4_0")
