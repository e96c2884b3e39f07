"""Q4_K matrices (GGML super-blocks) through the streaming GEMV: the device re-pack keeps every bit of
the blocks, and y = W x agrees with the pinned dequantisation (oracle/quants.py, checked against gguf-py) to fp32 summation order."""
import numpy as np
import pytest
import torch

from oracle import quants as Q
from oracle.model import fxp_round
from test_ops_gpu import chk, D  # noqa: F401

pytestmark = pytest.mark.gpu
Q4K = 12


def pack_k(lib, D, blob, rows, cols, blob2=None):
    total = rows * (2 if blob2 is not None else 1)
    d = D(lib.b2_packed_bytes(Q4K, total, cols))
    b1 = np.ascontiguousarray(blob)
    b2 = np.ascontiguousarray(blob2) if blob2 is not None else None
    chk(lib, lib.b2_op_pack_weights(Q4K, rows, cols, b1.ctypes.data, b2.ctypes.data if b2 is not None else None, d.ptr))
    return d


def ref(blob, x):
    W = Q.dequantize_q4_k(blob).astype(np.float64)
    xq = fxp_round(torch.from_numpy(x)).numpy().astype(np.float64)
    return W @ xq, np.abs(W) @ np.abs(xq)


def rand_blob(rng, rows, cols, scale=0.02):
    return Q.quantize_q4_k((rng.standard_normal((rows, cols)) * scale + 0.003).astype(np.float32))


@pytest.mark.parametrize("rows,cols", [(16, 256), (100, 512), (256, 4096), (64, 14336), (1000, 1536)])
def test_q4k_gemv_f32_input(lib, D, rows, cols):
    rng = np.random.default_rng(rows + cols)
    blob = rand_blob(rng, rows, cols)
    x = (rng.standard_normal(cols) * np.exp(rng.standard_normal(cols))).astype(np.float32)
    res = rng.standard_normal(rows).astype(np.float32)
    dW, dx, dres, dy = pack_k(lib, D, blob, rows, cols), D.from_numpy(x), D.from_numpy(res), D(rows * 4)
    want, mag = ref(blob, x)
    chk(lib, lib.b2_op_gemv_f32(Q4K, dW.ptr, rows, cols, dx.ptr, None, 0, dy.ptr))
    y = dy.to_numpy(np.float32, (rows,))
    assert np.all(np.abs(y - want) <= 8e-6 * mag + 1e-30)
    chk(lib, lib.b2_op_gemv_f32(Q4K, dW.ptr, rows, cols, dx.ptr, dres.ptr, 1, dres.ptr))
    assert np.all(np.abs(dres.to_numpy(np.float32, (rows,)) - (want + res)) <= 8e-6 * mag + 1e-6 * np.abs(res))
    # activation-image input (the O projection's path): block sums are rebuilt from the digits
    act = D(lib.b2_act_bytes(cols))
    chk(lib, lib.b2_op_quant_act(dx.ptr, cols, act.ptr))
    dy2 = D(rows * 4)
    chk(lib, lib.b2_op_gemv(Q4K, dW.ptr, rows, cols, act.ptr, None, 0, 0, dy2.ptr))
    assert np.all(np.abs(dy2.to_numpy(np.float32, (rows,)) - want) <= 8e-6 * mag + 1e-30)


@pytest.mark.parametrize("V,H,plus_one", [(1000, 512, 0), (5000, 4096, 1)])
def test_q4k_norm_gemv_logits(lib, D, V, H, plus_one):
    rng = np.random.default_rng(V + H)
    blob = rand_blob(rng, V, H)
    x = (rng.standard_normal(H) * 2).astype(np.float32)
    w = (rng.standard_normal(H) * 0.1 + (0 if plus_one else 1)).astype(np.float32)
    eps = np.float32(1e-5)
    dW, dx, dw = pack_k(lib, D, blob, V, H), D.from_numpy(x), D.from_numpy(w)
    dy, dtok = D(V * 4), D(4)
    chk(lib, lib.b2_op_norm_gemv(Q4K, dW.ptr, V, H, dx.ptr, dw.ptr, eps, plus_one, None, 3, 0, dy.ptr, None, dtok.ptr))
    y = dy.to_numpy(np.float32, (V,))
    tx = torch.from_numpy(x)
    xn = (tx * (1.0 / torch.sqrt((tx * tx).mean() + eps))) * ((1.0 + torch.from_numpy(w)) if plus_one else torch.from_numpy(w))
    want, mag = ref(blob, xn.numpy())
    assert np.all(np.abs(y - want) <= 1e-5 * mag + 1e-30)
    assert int(dtok.to_numpy(np.int32, (1,))[0]) == int(np.argmax(y))


def test_q4k_gate_up(lib, D):
    rng = np.random.default_rng(5)
    I, H = 512, 1024
    g, u = rand_blob(rng, I, H, 0.05), rand_blob(rng, I, H, 0.05)
    x = rng.standard_normal(H).astype(np.float32)
    dW, dx, dw = pack_k(lib, D, g, I, H, u), D.from_numpy(x), D.from_numpy(np.ones(H, dtype=np.float32))
    dy = D(I * 4)
    chk(lib, lib.b2_op_norm_gemv(Q4K, dW.ptr, 2 * I, H, dx.ptr, dw.ptr, np.float32(1e-6), 0, None, 2, 0, dy.ptr, None, None))
    tx = torch.from_numpy(x)
    xn = (tx * (1.0 / torch.sqrt((tx * tx).mean() + np.float32(1e-6)))).numpy()
    gv, _ = ref(g, xn)
    uv, _ = ref(u, xn)
    gt = torch.from_numpy(gv)
    want = (gt / (1 + torch.exp(-gt))).numpy() * uv
    assert np.abs(dy.to_numpy(np.float32, (I,)) - want).max() <= 3e-5 * np.abs(want).max()


@pytest.mark.parametrize("key,head_qtype", [("qwen2:1.5b", "q4_k"), ("gemma:2b", "q4_0"), ("mistral:7b", "q4_k")])
def test_q4k_model_matches_oracle(key, head_qtype, gpu_engine):
    """The whole decode path on Q4_K layer matrices (and a Q4_K or Q4_0 head): teacher-forced logits and greedy tokens
    against the PyTorch oracle on the same GGML blocks (oracle/randinit.make_model quantises with the restated llama.cpp search)."""
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import archs, engine
    from oracle import archs as oarchs
    from oracle import randinit as orand
    from oracle.model import OracleModel
    a = archs.TINY[key]
    if a.H % 256 or a.I % 256 or a.q_dim % 256:
        pytest.skip("Q4_K needs 256-element super-blocks along K")
    tensors = orand.make_model(oarchs.TINY[key], seed=21, qtype="q4_k", head_qtype=head_qtype)
    m = engine.Model.from_tensors(gpu_engine, a, tensors, max_ctx=96)
    om = OracleModel(oarchs.TINY[key], tensors)
    prompt = [1, 5, 77, 300, 12, 9, 640, 2]
    got = m.session(96).forward_tokens(prompt)
    os_ = om.session(96)
    want = np.stack([os_.step(t).numpy() for t in prompt])
    assert np.abs(got - want).max() <= 2e-4 * np.abs(want).max()
    s = m.session(96)
    toks = s.generate_greedy(prompt, 40)
    assert list(map(int, toks)) == om.session(96).generate_greedy(prompt, 40)


@pytest.mark.parametrize("rows,cols,M", [(128, 256, 3), (1000, 512, 17), (4096, 4096, 33), (256, 14336, 64), (512, 4096, 128), (256, 14336, 90)])
def test_q4k_prefill_gemm(lib, D, rows, cols, M):
    """The tcgen05 prompt GEMM on a Q4_K matrix against the decode GEMV and the dequantised reference."""
    rng = np.random.default_rng(rows + cols + M)
    blob = rand_blob(rng, rows, cols)
    x = (rng.standard_normal((M, cols)) * np.exp(rng.standard_normal((M, cols)))).astype(np.float32)
    dW, dx = pack_k(lib, D, blob, rows, cols), D.from_numpy(x)
    dy = D(M * rows * 4)
    chk(lib, lib.b2_op_prefill_gemm(Q4K, dW.ptr, rows, cols, dx.ptr, M, None, 0.0, 0, None, 0, 0, dy.ptr))
    y = dy.to_numpy(np.float32, (M, rows))
    dy1, dx1 = D(rows * 4), D(cols * 4)
    for m in sorted({0, M // 2, M - 1}):
        want, mag = ref(blob, x[m])
        assert np.all(np.abs(y[m] - want) <= 8e-6 * mag + 1e-30), f"token {m} vs reference"
        chk(lib, lib.b2_dev_upload(dx1.ptr, x[m].ctypes.data, cols * 4))
        chk(lib, lib.b2_op_gemv_f32(Q4K, dW.ptr, rows, cols, dx1.ptr, None, 0, dy1.ptr))
        assert np.all(np.abs(y[m] - dy1.to_numpy(np.float32, (rows,))) <= 3e-6 * mag + 1e-30), f"token {m} vs decode kernel"
