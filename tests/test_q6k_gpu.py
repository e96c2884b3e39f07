"""Q6_K matrices (llama.cpp's format for output.weight and a tied token_embd.weight) through the streaming GEMV and the engine."""
import numpy as np
import pytest
import torch

from oracle import quants as Q
from oracle.model import fxp_round
from test_ops_gpu import chk, D  # noqa: F401

pytestmark = pytest.mark.gpu
Q6K = 14


def pack6(lib, D, blob, rows, cols):
    d = D(lib.b2_packed_bytes(Q6K, rows, cols))
    b1 = np.ascontiguousarray(blob)
    chk(lib, lib.b2_op_pack_weights(Q6K, rows, cols, b1.ctypes.data, None, d.ptr))
    return d


def ref(blob, x):
    W = Q.dequantize_q6_k(blob).astype(np.float64)
    xq = fxp_round(torch.from_numpy(x)).numpy().astype(np.float64)
    return W @ xq, np.abs(W) @ np.abs(xq)


@pytest.mark.parametrize("rows,cols", [(16, 256), (100, 512), (1000, 4096), (64, 14336), (777, 1536)])
def test_q6k_gemv(lib, D, rows, cols):
    rng = np.random.default_rng(rows + cols)
    blob = Q.quantize_q6_k((rng.standard_normal((rows, cols)) * 0.02).astype(np.float32))
    x = (rng.standard_normal(cols) * np.exp(rng.standard_normal(cols))).astype(np.float32)
    dW, dx, dy = pack6(lib, D, blob, rows, cols), D.from_numpy(x), D(rows * 4)
    want, mag = ref(blob, x)
    chk(lib, lib.b2_op_gemv_f32(Q6K, dW.ptr, rows, cols, dx.ptr, None, 0, dy.ptr))
    assert np.all(np.abs(dy.to_numpy(np.float32, (rows,)) - want) <= 8e-6 * mag + 1e-30)
    act = D(lib.b2_act_bytes(cols))
    chk(lib, lib.b2_op_quant_act(dx.ptr, cols, act.ptr))
    dy2 = D(rows * 4)
    chk(lib, lib.b2_op_gemv(Q6K, dW.ptr, rows, cols, act.ptr, None, 0, 0, dy2.ptr))
    assert np.all(np.abs(dy2.to_numpy(np.float32, (rows,)) - want) <= 8e-6 * mag + 1e-30)


@pytest.mark.parametrize("rows,cols,M", [(128, 256, 3), (1000, 512, 17), (4096, 4096, 33), (256, 14336, 64), (512, 4096, 128), (256, 14336, 90)])
def test_q6k_prefill_gemm(lib, D, rows, cols, M):
    """The tcgen05 prompt GEMM on a Q6_K matrix (llama.cpp's Q4_K_M keeps attn_v and ffn_down in this format) against the
    dequantised reference and the decode GEMV: the builder dequantises d*sc*(q-32) exactly and splits it into three bf16 parts."""
    rng = np.random.default_rng(rows + cols + M)
    blob = Q.quantize_q6_k((rng.standard_normal((rows, cols)) * 0.02).astype(np.float32))
    x = (rng.standard_normal((M, cols)) * np.exp(rng.standard_normal((M, cols)))).astype(np.float32)
    dW, dx = pack6(lib, D, blob, rows, cols), D.from_numpy(x)
    dy = D(M * rows * 4)
    chk(lib, lib.b2_op_prefill_gemm(Q6K, dW.ptr, rows, cols, dx.ptr, M, None, 0.0, 0, None, 0, 0, dy.ptr))
    y = dy.to_numpy(np.float32, (M, rows))
    dy1, dx1 = D(rows * 4), D(cols * 4)
    for m in sorted({0, M // 2, M - 1}):
        want, mag = ref(blob, x[m])
        assert np.all(np.abs(y[m] - want) <= 8e-6 * mag + 1e-30), f"token {m} vs reference"
        chk(lib, lib.b2_dev_upload(dx1.ptr, x[m].ctypes.data, cols * 4))
        chk(lib, lib.b2_op_gemv_f32(Q6K, dW.ptr, rows, cols, dx1.ptr, None, 0, dy1.ptr))
        assert np.all(np.abs(y[m] - dy1.to_numpy(np.float32, (rows,))) <= 8e-6 * mag + 1e-30), f"token {m} vs decode kernel"


def test_q6k_head_logits(lib, D):
    rng = np.random.default_rng(9)
    V, H = 5000, 4096
    blob = Q.quantize_q6_k((rng.standard_normal((V, H)) * 0.02).astype(np.float32))
    x = (rng.standard_normal(H) * 2).astype(np.float32)
    w = (rng.standard_normal(H) * 0.1 + 1).astype(np.float32)
    eps = np.float32(1e-5)
    dW, dx, dw = pack6(lib, D, blob, V, H), D.from_numpy(x), D.from_numpy(w)
    dy, dtok = D(V * 4), D(4)
    chk(lib, lib.b2_op_norm_gemv(Q6K, dW.ptr, V, H, dx.ptr, dw.ptr, eps, 0, None, 3, 0, dy.ptr, None, dtok.ptr))
    y = dy.to_numpy(np.float32, (V,))
    tx = torch.from_numpy(x)
    xn = (tx * (1.0 / torch.sqrt((tx * tx).mean() + eps))) * torch.from_numpy(w)
    want, mag = ref(blob, xn.numpy())
    assert np.all(np.abs(y - want) <= 1e-5 * mag + 1e-30)
    assert int(dtok.to_numpy(np.int32, (1,))[0]) == int(np.argmax(y))


@pytest.mark.parametrize("key,qtype", [("qwen2:1.5b", "q4_0"), ("gemma:2b", "q4_0"), ("mistral:7b", "q4_k")])
def test_model_with_q6k_head_matches_oracle(key, qtype, gpu_engine):
    """What a real llama.cpp "Q4_0" file looks like: Q4_0 (or Q4_K) layers, Q6_K output / tied embedding."""
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import archs, engine
    from oracle import archs as oarchs
    from oracle import randinit as orand
    from oracle.model import OracleModel
    a = archs.TINY[key]
    if a.H % 256:
        pytest.skip("K-quants need 256-element super-blocks along K")
    tensors = orand.make_model(oarchs.TINY[key], seed=33, qtype=qtype, head_qtype="q6_k")
    m = engine.Model.from_tensors(gpu_engine, a, tensors, max_ctx=96)
    om = OracleModel(oarchs.TINY[key], tensors)
    prompt = [1, 5, 77, 300, 12, 9, 640, 2]
    got = m.session(96).forward_tokens(prompt)
    os_ = om.session(96)
    want = np.stack([os_.step(t).numpy() for t in prompt])
    assert np.abs(got - want).max() <= 2e-4 * np.abs(want).max()
    s = m.session(96)
    assert list(map(int, s.generate_greedy(prompt, 40))) == om.session(96).generate_greedy(prompt, 40)
    # the prompt as ONE batch: q, k, v of different formats and the Q6_K ffn_down all go through the tcgen05 prompt GEMM
    long_prompt = [1 + (i * 37) % (a.V - 1) for i in range(70)]
    s2 = m.session(96)
    lg = s2.prefill(long_prompt, want_logits=True)
    want_lg = om.session(96).prefill(long_prompt).numpy()
    assert np.abs(lg - want_lg).max() <= 2e-4 * np.abs(want_lg).max()
    assert list(map(int, s2.decode(20))) == om.session(96).generate_greedy(long_prompt, 20)


def test_q4_k_m_style_mixed_formats_match_oracle(gpu_engine):
    """llama.cpp's Q4_K_M: Q4_K everywhere except Q6_K for attn_v, ffn_down and the head: q/k/v then run as three GEMVs."""
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import archs, engine
    from oracle import archs as oarchs
    from oracle import randinit as orand
    from oracle.model import OracleModel
    key = "mistral:7b"
    a, oa = archs.TINY[key], oarchs.TINY[key]
    tensors = orand.make_model(oa, seed=44, qtype="q4_k", head_qtype="q6_k")
    for name, kind, layer, rows, cols in oarchs.tensor_table(oa):
        if kind in ("attn_v", "ffn_down") and (kind == "attn_v" or layer % 2 == 0):
            tensors[name] = ("q6_k", orand.make_tensor(oa, 44, kind, layer, rows, cols, "q6_k"))
    m = engine.Model.from_tensors(gpu_engine, a, tensors, max_ctx=96)
    om = OracleModel(oa, tensors)
    prompt = [1, 5, 77, 300, 12, 9, 640, 2]
    got = m.session(96).forward_tokens(prompt)
    os_ = om.session(96)
    want = np.stack([os_.step(t).numpy() for t in prompt])
    assert np.abs(got - want).max() <= 2e-4 * np.abs(want).max()
    s = m.session(96)
    assert list(map(int, s.generate_greedy(prompt, 40))) == om.session(96).generate_greedy(prompt, 40)
    # the prompt as ONE batch: q, k, v of different formats and the Q6_K ffn_down all go through the tcgen05 prompt GEMM
    long_prompt = [1 + (i * 37) % (a.V - 1) for i in range(70)]
    s2 = m.session(96)
    lg = s2.prefill(long_prompt, want_logits=True)
    want_lg = om.session(96).prefill(long_prompt).numpy()
    assert np.abs(lg - want_lg).max() <= 2e-4 * np.abs(want_lg).max()
    assert list(map(int, s2.decode(20))) == om.session(96).generate_greedy(long_prompt, 20)
