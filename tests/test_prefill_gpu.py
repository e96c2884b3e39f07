"""Prompt-evaluation GEMM (tcgen05 tensor cores, csrc/prefill.cu) against the decode GEMV kernels and the numpy oracle:
the same block fixed-point arithmetic, so results agree to fp32 accumulation order (the integer block dot products are exact)."""
import numpy as np
import pytest
import torch

from oracle import quants as Q
from test_ops_gpu import chk, pack, ref_gemv   # noqa: F401  (helpers of the op-level tests)
from test_ops_gpu import D  # noqa: F401  (fixture)

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,cols,M", [(128, 128, 1), (256, 288, 5), (384, 4096, 16), (1000, 512, 17), (4096, 4096, 33),
                                         (128, 14336, 64), (6144, 4096, 64), (2048, 1536, 70),
                                         (4096, 4096, 128), (128, 14336, 100), (1000, 1536, 200)])   # > 64 tokens: 128 per weight pass, one drain
def test_prefill_gemm_matches_decode_gemv(lib, D, rows, cols, M):
    rng = np.random.default_rng(rows * 3 + cols + M)
    blob = Q.quantize_q4_0((rng.standard_normal((rows, cols)) * 0.02).astype(np.float32))
    x = (rng.standard_normal((M, cols)) * np.exp(rng.standard_normal((M, cols)))).astype(np.float32)
    bias = rng.standard_normal(rows).astype(np.float32)
    dW, dx, db = pack(lib, D, blob, rows, cols), D.from_numpy(x), D.from_numpy(bias)
    dy = D(M * rows * 4)
    chk(lib, lib.b2_op_prefill_gemm(2, dW.ptr, rows, cols, dx.ptr, M, None, 0.0, 0, db.ptr, 0, 0, dy.ptr))
    y = dy.to_numpy(np.float32, (M, rows))
    dy1, dx1 = D(rows * 4), D(cols * 4)
    for m in sorted({0, M // 2, M - 1}):
        ref, mag = ref_gemv(blob, x[m])
        assert np.all(np.abs(y[m] - (ref + bias)) <= 6e-6 * mag + 1e-6 * np.abs(bias) + 1e-30), f"token {m} vs oracle"
        chk(lib, lib.b2_dev_upload(dx1.ptr, x[m].ctypes.data, cols * 4))
        chk(lib, lib.b2_op_gemv_f32(2, dW.ptr, rows, cols, dx1.ptr, db.ptr, 0, dy1.ptr))
        g = dy1.to_numpy(np.float32, (rows,))
        assert np.all(np.abs(y[m] - g) <= 2e-6 * mag + 1e-30), f"token {m} vs decode kernel"


@pytest.mark.parametrize("H,M,plus_one", [(288, 3, 0), (2048, 20, 1), (4096, 40, 0), (4096, 128, 0)])
def test_prefill_gemm_norm_and_residual(lib, D, H, M, plus_one):
    rng = np.random.default_rng(H + M)
    rows = 256
    blob = Q.quantize_q4_0((rng.standard_normal((rows, H)) * 0.02).astype(np.float32))
    x = (rng.standard_normal((M, H)) * 2).astype(np.float32)
    w = (rng.standard_normal(H) * 0.1 + (0 if plus_one else 1)).astype(np.float32)
    res = rng.standard_normal((M, rows)).astype(np.float32)
    eps = np.float32(1e-5)
    dW, dx, dw, dres = pack(lib, D, blob, rows, H), D.from_numpy(x), D.from_numpy(w), D.from_numpy(res)
    chk(lib, lib.b2_op_prefill_gemm(2, dW.ptr, rows, H, dx.ptr, M, dw.ptr, eps, plus_one, dres.ptr, 1, 0, dres.ptr))   # in place
    y = dres.to_numpy(np.float32, (M, rows))
    for m in range(M):
        tx = torch.from_numpy(x[m])
        xn = (tx * (1.0 / torch.sqrt((tx * tx).mean() + eps))) * ((1.0 + torch.from_numpy(w)) if plus_one else torch.from_numpy(w))
        ref, mag = ref_gemv(blob, xn.numpy())
        assert np.all(np.abs(y[m] - (ref + res[m])) <= 8e-6 * mag + 1e-6 * np.abs(res[m]) + 1e-30)


@pytest.mark.parametrize("act_id,I,H,M", [(0, 416, 288, 7), (1, 512, 256, 33), (0, 1024, 4096, 64), (0, 1024, 4096, 97)])
def test_prefill_gemm_gate_up(lib, D, act_id, I, H, M):
    rng = np.random.default_rng(I + H + M)
    g = Q.quantize_q4_0((rng.standard_normal((I, H)) * 0.05).astype(np.float32))
    u = Q.quantize_q4_0((rng.standard_normal((I, H)) * 0.05).astype(np.float32))
    x = rng.standard_normal((M, H)).astype(np.float32)
    dW, dx = pack(lib, D, g, I, H, u), D.from_numpy(x)
    dy = D(M * I * 4)
    chk(lib, lib.b2_op_prefill_gemm(2, dW.ptr, 2 * I, H, dx.ptr, M, None, 0.0, 0, None, 2, act_id, dy.ptr))
    y = dy.to_numpy(np.float32, (M, I))
    for m in (0, M - 1):
        gv, _ = ref_gemv(g, x[m])
        uv, _ = ref_gemv(u, x[m])
        gt = torch.from_numpy(gv)
        a = gt / (1 + torch.exp(-gt)) if act_id == 0 else 0.5 * gt * (1 + torch.tanh(0.7978845608028654 * (gt + 0.044715 * gt ** 3)))
        ref = a.numpy() * uv
        assert np.abs(y[m] - ref).max() <= 2e-5 * np.abs(ref).max()
