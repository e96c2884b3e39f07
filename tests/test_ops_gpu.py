"""Parity of every decode-path kernel, called through the C ABI, against the oracle (bit-exact for
integer / byte / index work; float tolerances stated per test)."""
import ctypes as C
import math

import numpy as np
import pytest
import torch

from oracle import quants as Q
from oracle import sampler as osampler
from oracle.model import fxp_digits, fxp_round

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def D(lib):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.devmem import DevBuf
    return DevBuf


def chk(lib, rc):
    assert rc == 0, lib.b2_last_error().decode()


def a3_off(blk, c, e):
    """Byte offset of (block, digit plane, element) in the engine's activation image (csrc/layout.cuh)."""
    return (blk >> 2) * 384 + c * 128 + ((((e & 15) >> 2) * 4 + (blk & 3)) << 3) + ((e >> 4) << 2) + (e & 3)


def read_act(lib, buf, n):
    """-> (digits int8[3, n] in element order, scale float32[n/32]) from the engine's activation image."""
    nb = n // 32
    nbp = (nb + 3) // 4 * 4
    assert lib.b2_act_bytes(n) == nbp * 100
    raw = buf.to_numpy(np.uint8, (nbp * 100,))
    blk, c, e = np.meshgrid(np.arange(nbp), np.arange(3), np.arange(32), indexing="ij")
    planes = raw[: nbp * 96].view(np.int8)[a3_off(blk, c, e)]     # [nbp, 3, 32] in element order
    assert not planes[nb:].any()                                  # padding blocks are zero
    scale = raw[nbp * 96: nbp * 100].view(np.float32)[:nb]
    return planes[:nb].transpose(1, 0, 2).reshape(3, n), scale


def make_act(lib, D, x):
    n = len(x)
    dx = D.from_numpy(x.astype(np.float32))
    act = D(lib.b2_act_bytes(n))
    chk(lib, lib.b2_op_quant_act(dx.ptr, n, act.ptr))
    return act


@pytest.mark.parametrize("n", [32, 96, 4096, 14336])
def test_quant_act_bit_exact(lib, D, n):
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(n) * np.exp(rng.standard_normal(n) * 2)).astype(np.float32)
    x[:32] = 0.0                      # all-zero block
    if n > 64:
        x[32:64] = 1e-39              # denormal block
        x[64] = 2.0                   # exact power-of-two maximum
        x[65:96] = np.float32(2.0) * np.float32(1 - 2 ** -24)
    act = make_act(lib, D, x)
    dig, scale = read_act(lib, act, n)
    odig, oe = fxp_digits(x)
    assert np.array_equal(dig, odig)
    assert np.array_equal(scale, np.ldexp(np.float32(1), oe).astype(np.float32))


def test_quant_act_rejects_ragged(lib, D):
    d = D(1024)
    assert lib.b2_op_quant_act(d.ptr, 33, d.ptr) == -1
    assert b"multiple of 32" in lib.b2_last_error()
    assert lib.b2_op_quant_act(d.ptr, 0, d.ptr) == 0


@pytest.mark.parametrize("n,plus_one", [(256, 0), (288, 1), (4096, 0), (3584, 1)])
def test_rmsnorm_quant(lib, D, n, plus_one):
    rng = np.random.default_rng(n)
    x = rng.standard_normal(n).astype(np.float32) * 3
    w = (rng.standard_normal(n) * 0.1 + (0 if plus_one else 1)).astype(np.float32)
    eps = np.float32(1e-5)
    dx, dw, dy = D.from_numpy(x), D.from_numpy(w), D(n * 4)
    act = D(lib.b2_act_bytes(n))
    chk(lib, lib.b2_op_rmsnorm_quant(dx.ptr, dw.ptr, n, eps, plus_one, act.ptr, dy.ptr))
    y = dy.to_numpy(np.float32, (n,))
    tx = torch.from_numpy(x)
    ref = (tx * (1.0 / torch.sqrt((tx * tx).mean() + eps))) * ((1.0 + torch.from_numpy(w)) if plus_one else torch.from_numpy(w))
    # fp32, differs only by the summation order of mean(x^2): 1e-6 relative
    assert np.abs(y - ref.numpy()).max() <= 1e-6 * np.abs(ref.numpy()).max()
    # the digits are exactly the grid of the kernel's own y
    dig, scale = read_act(lib, act, n)
    odig, oe = fxp_digits(y)
    assert np.array_equal(dig, odig)
    assert np.array_equal(scale, np.ldexp(np.float32(1), oe).astype(np.float32))


def pack(lib, D, blob, rows, cols, blob2=None):
    total = rows * (2 if blob2 is not None else 1)
    d = D(lib.b2_packed_bytes(2, total, cols))
    b1 = np.ascontiguousarray(blob)
    b2 = np.ascontiguousarray(blob2) if blob2 is not None else None
    chk(lib, lib.b2_op_pack_weights(2, rows, cols, b1.ctypes.data, b2.ctypes.data if b2 is not None else None, d.ptr))
    return d


def ref_gemv(blob, x):
    """float64 reference on the activation grid: the CUDA kernel is exact per block, fp32 across blocks."""
    W = Q.dequantize_q4_0(blob).astype(np.float64)
    xq = fxp_round(torch.from_numpy(x)).numpy().astype(np.float64)
    return W @ xq, np.abs(W) @ np.abs(xq)


@pytest.mark.parametrize("rows,cols", [(1, 32), (7, 288), (64, 4096), (33, 1184), (16, 14336), (257, 1536), (4096, 4096), (6000, 2080), (9600, 8192)])
def test_gemv_modes(lib, D, rows, cols):
    rng = np.random.default_rng(rows * 131 + cols)
    Wf = (rng.standard_normal((rows, cols)) * 0.02).astype(np.float32)
    blob = Q.quantize_q4_0(Wf)
    x = (rng.standard_normal(cols) * np.exp(rng.standard_normal(cols))).astype(np.float32)
    act = make_act(lib, D, x)
    dW = pack(lib, D, blob, rows, cols)
    ref, mag = ref_gemv(blob, x)
    tol = 4e-6 * mag + 1e-30
    # mode 0 without and with bias
    dy = D(rows * 4)
    chk(lib, lib.b2_op_gemv(2, dW.ptr, rows, cols, act.ptr, None, 0, 0, dy.ptr))
    y = dy.to_numpy(np.float32, (rows,))
    assert np.all(np.abs(y - ref) <= tol)
    bias = rng.standard_normal(rows).astype(np.float32)
    db = D.from_numpy(bias)
    chk(lib, lib.b2_op_gemv(2, dW.ptr, rows, cols, act.ptr, db.ptr, 0, 0, dy.ptr))
    assert np.all(np.abs(dy.to_numpy(np.float32, (rows,)) - (ref + bias)) <= tol + 1e-6 * np.abs(bias))
    # mode 1: residual, in place
    dres = D.from_numpy(bias)
    chk(lib, lib.b2_op_gemv(2, dW.ptr, rows, cols, act.ptr, dres.ptr, 1, 0, dres.ptr))
    assert np.all(np.abs(dres.to_numpy(np.float32, (rows,)) - (ref + bias)) <= tol + 1e-6 * np.abs(bias))


@pytest.mark.parametrize("rows,cols", [(7, 288), (4096, 14336), (3072, 24576), (100, 32)])
def test_gemv_f32_input_is_gridded_in_kernel(lib, D, rows, cols):
    """The engine's down projection: fp32 vector in, activation grid applied inside the kernel."""
    rng = np.random.default_rng(rows + cols)
    blob = Q.quantize_q4_0((rng.standard_normal((rows, cols)) * 0.02).astype(np.float32))
    x = (rng.standard_normal(cols) * np.exp(rng.standard_normal(cols))).astype(np.float32)
    res = rng.standard_normal(rows).astype(np.float32)
    dW, dx, dres, dy = pack(lib, D, blob, rows, cols), D.from_numpy(x), D.from_numpy(res), D(rows * 4)
    ref, mag = ref_gemv(blob, x)
    chk(lib, lib.b2_op_gemv_f32(2, dW.ptr, rows, cols, dx.ptr, None, 0, dy.ptr))
    assert np.all(np.abs(dy.to_numpy(np.float32, (rows,)) - ref) <= 4e-6 * mag + 1e-30)
    chk(lib, lib.b2_op_gemv_f32(2, dW.ptr, rows, cols, dx.ptr, dres.ptr, 1, dres.ptr))
    assert np.all(np.abs(dres.to_numpy(np.float32, (rows,)) - (ref + res)) <= 4e-6 * mag + 1e-6 * np.abs(res))
    # it is the same function as gridding first and calling the activation-image entry point: bit-identical
    act = make_act(lib, D, x)
    dy2 = D(rows * 4)
    chk(lib, lib.b2_op_gemv(2, dW.ptr, rows, cols, act.ptr, None, 0, 0, dy2.ptr))
    chk(lib, lib.b2_op_gemv_f32(2, dW.ptr, rows, cols, dx.ptr, None, 0, dy.ptr))
    assert np.array_equal(dy.to_numpy(np.float32, (rows,)), dy2.to_numpy(np.float32, (rows,)))


@pytest.mark.parametrize("act_id,I,H", [(0, 416, 288), (1, 512, 256), (0, 1024, 4096)])
def test_gemv_gate_up_pairs(lib, D, act_id, I, H):
    rng = np.random.default_rng(I + H)
    g = Q.quantize_q4_0((rng.standard_normal((I, H)) * 0.05).astype(np.float32))
    u = Q.quantize_q4_0((rng.standard_normal((I, H)) * 0.05).astype(np.float32))
    x = rng.standard_normal(H).astype(np.float32)
    act = make_act(lib, D, x)
    dW = pack(lib, D, g, I, H, u)
    dy = D(I * 4)
    chk(lib, lib.b2_op_gemv(2, dW.ptr, 2 * I, H, act.ptr, None, 2, act_id, dy.ptr))
    y = dy.to_numpy(np.float32, (I,))
    gv, _ = ref_gemv(g, x)
    uv, _ = ref_gemv(u, x)
    gt = torch.from_numpy(gv)
    a = gt / (1 + torch.exp(-gt)) if act_id == 0 else 0.5 * gt * (1 + torch.tanh(0.7978845608028654 * (gt + 0.044715 * gt ** 3)))
    ref = a.numpy() * uv
    assert np.abs(y - ref).max() <= 2e-5 * np.abs(ref).max()
    assert lib.b2_op_gemv(2, dW.ptr, 2 * I + 1, H, act.ptr, None, 2, act_id, dy.ptr) == -1


@pytest.mark.parametrize("V,H,plus_one", [(1000, 288, 0), (1024, 512, 1), (32064, 3072, 0), (5000, 4096, 0)])
def test_norm_gemv_logits_and_argmax(lib, D, V, H, plus_one):
    """The engine's fused head: RMSNorm -> activation grid -> GEMV -> logits + greedy token."""
    rng = np.random.default_rng(V + H)
    blob = Q.quantize_q4_0((rng.standard_normal((V, H)) * 0.02).astype(np.float32))
    x = (rng.standard_normal(H) * 2).astype(np.float32)
    w = (rng.standard_normal(H) * 0.1 + (0 if plus_one else 1)).astype(np.float32)
    eps = np.float32(1e-5)
    dW, dx, dw = pack(lib, D, blob, V, H), D.from_numpy(x), D.from_numpy(w)
    dy, dtok = D(V * 4), D(4)
    chk(lib, lib.b2_op_norm_gemv(2, dW.ptr, V, H, dx.ptr, dw.ptr, eps, plus_one, None, 3, 0, dy.ptr, None, dtok.ptr))
    y = dy.to_numpy(np.float32, (V,))
    tx = torch.from_numpy(x)
    xn = (tx * (1.0 / torch.sqrt((tx * tx).mean() + eps))) * ((1.0 + torch.from_numpy(w)) if plus_one else torch.from_numpy(w))
    ref, mag = ref_gemv(blob, xn.numpy())
    assert np.all(np.abs(y - ref) <= 6e-6 * mag + 1e-30)
    assert int(dtok.to_numpy(np.int32, (1,))[0]) == int(np.argmax(y))      # arg-max of the kernel's own logits, lowest index


@pytest.mark.parametrize("I,H,act_id", [(416, 288, 0), (1024, 512, 1), (14336, 4096, 0)])
def test_norm_gemv_gate_up_emits_activation_grid(lib, D, I, H, act_id):
    rng = np.random.default_rng(I + H)
    g = Q.quantize_q4_0((rng.standard_normal((I, H)) * 0.05).astype(np.float32))
    u = Q.quantize_q4_0((rng.standard_normal((I, H)) * 0.05).astype(np.float32))
    x = rng.standard_normal(H).astype(np.float32)
    w = np.ones(H, dtype=np.float32)
    dW, dx, dw = pack(lib, D, g, I, H, u), D.from_numpy(x), D.from_numpy(w)
    dy, dact = D(I * 4), D(lib.b2_act_bytes(I))
    dact_zero = np.zeros(lib.b2_act_bytes(I), dtype=np.uint8)
    chk(lib, lib.b2_dev_upload(dact.ptr, dact_zero.ctypes.data, dact_zero.nbytes))
    chk(lib, lib.b2_op_norm_gemv(2, dW.ptr, 2 * I, H, dx.ptr, dw.ptr, np.float32(1e-6), 0, None, 2, act_id, dy.ptr, dact.ptr, None))
    h = dy.to_numpy(np.float32, (I,))
    dig, scale = read_act(lib, dact, I)
    odig, oe = fxp_digits(h)                      # the emitted image is exactly the grid of the kernel's own h
    assert np.array_equal(dig, odig)
    assert np.array_equal(scale, np.ldexp(np.float32(1), oe).astype(np.float32))
    tx = torch.from_numpy(x)
    xn = (tx * (1.0 / torch.sqrt((tx * tx).mean() + np.float32(1e-6)))).numpy()
    gv, _ = ref_gemv(g, xn)
    uv, _ = ref_gemv(u, xn)
    gt = torch.from_numpy(gv)
    a = gt / (1 + torch.exp(-gt)) if act_id == 0 else 0.5 * gt * (1 + torch.tanh(0.7978845608028654 * (gt + 0.044715 * gt ** 3)))
    assert np.abs(h - a.numpy() * uv).max() <= 2e-5 * np.abs(a.numpy() * uv).max()


def rope_tables(hd, theta, n):
    inv = 1.0 / (theta ** (np.arange(0, hd, 2, dtype=np.float64) / hd))
    ang = np.arange(n, dtype=np.float64)[:, None] * inv[None]
    return np.cos(ang).astype(np.float32), np.sin(ang).astype(np.float32)


@pytest.mark.parametrize("nh,nkv,hd", [(4, 1, 128), (3, 3, 96), (2, 2, 256), (7, 1, 128)])
def test_rope_kv_append_bit_exact(lib, D, nh, nkv, hd):
    rng = np.random.default_rng(nh * 7 + hd)
    n_pages, max_ctx = 6, 96
    cs, sn = rope_tables(hd, 1e4, max_ctx)
    page_table = np.array([5, 2, 0, 4, 1, 3], dtype=np.int32)
    kc = np.zeros((n_pages, nkv, 16, hd), dtype=np.float16)
    vc = np.zeros_like(kc)
    dk, dv = D.from_numpy(kc), D.from_numpy(vc)
    dcs, dsn, dpt = D.from_numpy(cs), D.from_numpy(sn), D.from_numpy(page_table)
    for pos in (0, 15, 16, 77):
        qkv = rng.standard_normal((nh + 2 * nkv) * hd).astype(np.float32)
        dq = D.from_numpy(qkv)
        chk(lib, lib.b2_op_rope_kv(dq.ptr, nh, nkv, hd, pos, dcs.ptr, dsn.ptr, dk.ptr, dv.ptr, dpt.ptr, 16))
        got = dq.to_numpy(np.float32, qkv.shape)
        t = torch.from_numpy(qkv[: (nh + nkv) * hd]).reshape(nh + nkv, hd)
        c, s = torch.from_numpy(cs[pos]), torch.from_numpy(sn[pos])
        t1, t2 = t[:, : hd // 2], t[:, hd // 2:]
        rot = torch.cat([t1 * c - t2 * s, t2 * c + t1 * s], dim=1).numpy()
        assert np.array_equal(got[: nh * hd], rot[:nh].reshape(-1))           # q rotated in place
        assert np.array_equal(got[nh * hd:], qkv[nh * hd:])                     # k, v inputs untouched
        kk = dk.to_numpy(np.float16, kc.shape)
        vv = dv.to_numpy(np.float16, vc.shape)
        page, slot = page_table[pos // 16], pos % 16
        assert np.array_equal(kk[page, :, slot], rot[nh:].astype(np.float16))
        assert np.array_equal(vv[page, :, slot], qkv[(nh + nkv) * hd:].reshape(nkv, hd).astype(np.float16))


@pytest.mark.parametrize("nh,nkv,hd,n_ctx", [(4, 1, 128, 1), (4, 1, 128, 17), (3, 3, 96, 130), (2, 2, 256, 300),
                                             (32, 8, 128, 2048), (7, 1, 128, 1000), (8, 1, 256, 129)])
def test_attn_decode(lib, D, nh, nkv, hd, n_ctx):
    rng = np.random.default_rng(nh + hd + n_ctx)
    n_pages = (n_ctx + 15) // 16
    perm = rng.permutation(n_pages + 3).astype(np.int32)[:n_pages]
    kc = (rng.standard_normal((n_pages + 3, nkv, 16, hd))).astype(np.float16)
    vc = (rng.standard_normal((n_pages + 3, nkv, 16, hd))).astype(np.float16)
    q = (rng.standard_normal((nh, hd)) * 1.5).astype(np.float32)
    dout = D(nh * hd * 4)
    dq, dk, dv, dpt = D.from_numpy(q), D.from_numpy(kc), D.from_numpy(vc), D.from_numpy(perm)   # keep the buffers alive
    chk(lib, lib.b2_op_attn_decode(dq.ptr, nh, nkv, hd, n_ctx, dk.ptr, dv.ptr, dpt.ptr, 16, dout.ptr))
    out = dout.to_numpy(np.float32, (nh, hd))
    K = torch.from_numpy(kc[perm].transpose(0, 2, 1, 3).reshape(n_pages * 16, nkv, hd)[:n_ctx]).float()
    V = torch.from_numpy(vc[perm].transpose(0, 2, 1, 3).reshape(n_pages * 16, nkv, hd)[:n_ctx]).float()
    qg = torch.from_numpy(q).reshape(nkv, nh // nkv, hd)
    sc = torch.einsum("ngd,tnd->ngt", qg, K) * np.float32(1.0 / math.sqrt(hd))
    ref = torch.einsum("ngt,tnd->ngd", torch.softmax(sc, -1), V).reshape(nh, hd).numpy()
    assert np.abs(out - ref).max() <= 2e-5 * max(1.0, np.abs(ref).max())   # fp32, order of summation only


@pytest.mark.parametrize("nh,nkv,hd,pos", [(4, 1, 128, 0), (4, 1, 128, 16), (3, 3, 96, 129), (2, 2, 256, 299),
                                           (32, 8, 128, 2047), (7, 1, 128, 999), (8, 1, 256, 128)])
def test_attn_fused_matches_unfused_semantics(lib, D, nh, nkv, hd, pos):
    """RoPE + append + attention + activation grid in one kernel == oracle ops on the same inputs."""
    rng = np.random.default_rng(nh + hd + pos)
    max_ctx = 2048
    n_pages = max_ctx // 16
    perm = rng.permutation(n_pages).astype(np.int32)
    kc = rng.standard_normal((n_pages, nkv, 16, hd)).astype(np.float16)
    vc = rng.standard_normal((n_pages, nkv, 16, hd)).astype(np.float16)
    qkv = (rng.standard_normal((nh + 2 * nkv) * hd) * 1.5).astype(np.float32)
    cs, sn = rope_tables(hd, 1e4, max_ctx)
    dq, dk, dv, dpt, dcs, dsn = (D.from_numpy(x) for x in (qkv, kc, vc, perm, cs, sn))
    dout, dact = D(nh * hd * 4), D(lib.b2_act_bytes(nh * hd))
    z = np.zeros(lib.b2_act_bytes(nh * hd), dtype=np.uint8)
    chk(lib, lib.b2_dev_upload(dact.ptr, z.ctypes.data, z.nbytes))
    chk(lib, lib.b2_op_attn_fused(dq.ptr, nh, nkv, hd, pos, dcs.ptr, dsn.ptr, dk.ptr, dv.ptr, dpt.ptr, 16, max_ctx, dout.ptr, dact.ptr))
    out = dout.to_numpy(np.float32, (nh, hd))
    # oracle: rotate, append (f16), attend
    t = torch.from_numpy(qkv[: (nh + nkv) * hd]).reshape(nh + nkv, hd)
    c, s = torch.from_numpy(cs[pos]), torch.from_numpy(sn[pos])
    t1, t2 = t[:, : hd // 2], t[:, hd // 2:]
    rot = torch.cat([t1 * c - t2 * s, t2 * c + t1 * s], dim=1)
    kk, vv = kc.copy(), vc.copy()
    kk[perm[pos // 16], :, pos % 16] = rot[nh:].numpy().astype(np.float16)
    vv[perm[pos // 16], :, pos % 16] = qkv[(nh + nkv) * hd:].reshape(nkv, hd).astype(np.float16)
    assert np.array_equal(dk.to_numpy(np.float16, kc.shape), kk)            # append is bit-exact, nothing else touched
    assert np.array_equal(dv.to_numpy(np.float16, vc.shape), vv)
    n_ctx = pos + 1
    K = torch.from_numpy(kk[perm].transpose(0, 2, 1, 3).reshape(n_pages * 16, nkv, hd)[:n_ctx]).float()
    V = torch.from_numpy(vv[perm].transpose(0, 2, 1, 3).reshape(n_pages * 16, nkv, hd)[:n_ctx]).float()
    qg = rot[:nh].reshape(nkv, nh // nkv, hd)
    sc = torch.einsum("ngd,tnd->ngt", qg, K) * np.float32(1.0 / math.sqrt(hd))
    ref = torch.einsum("ngt,tnd->ngd", torch.softmax(sc, -1), V).reshape(nh, hd).numpy()
    assert np.abs(out - ref).max() <= 2e-5 * max(1.0, np.abs(ref).max())
    dig, scale = read_act(lib, dact, nh * hd)                                # the image is the grid of the kernel's own output
    odig, oe = fxp_digits(out.reshape(-1))
    assert np.array_equal(dig, odig)
    assert np.array_equal(scale, np.ldexp(np.float32(1), oe).astype(np.float32))


def test_attn_decode_rejects_bad_geometry(lib, D):
    d = D(4096)
    assert lib.b2_op_attn_decode(d.ptr, 4, 3, 128, 5, d.ptr, d.ptr, d.ptr, 16, d.ptr) == -1
    assert lib.b2_op_attn_decode(d.ptr, 4, 1, 100, 5, d.ptr, d.ptr, d.ptr, 16, d.ptr) == -1
    assert lib.b2_op_attn_decode(d.ptr, 4, 1, 128, 0, d.ptr, d.ptr, d.ptr, 16, d.ptr) == -1


@pytest.mark.parametrize("n", [1, 31, 1000, 32064, 128256, 256000])
def test_argmax_with_ties(lib, D, n):
    rng = np.random.default_rng(n)
    x = rng.standard_normal(n).astype(np.float32)
    dout = D(4)
    for variant in range(3):
        if variant == 1 and n > 2:      # exact tie: lowest index must win
            i, j = sorted(rng.choice(n, 2, replace=False))
            x[i] = x[j] = x.max() + 1
        if variant == 2:
            x[:] = -np.inf
            x[n - 1] = -1e30
        dx = D.from_numpy(x)
        chk(lib, lib.b2_op_argmax(dx.ptr, n, dout.ptr))
        assert int(dout.to_numpy(np.int32, (1,))[0]) == osampler.argmax_lowest(x)


@pytest.mark.parametrize("n,k", [(1000, 40), (128256, 40), (32064, 1), (300, 64)])
def test_sample_topk(lib, D, n, k):
    rng = np.random.default_rng(n + k)
    dout = D(4)
    checked = 0
    for trial in range(12):
        x = (rng.standard_normal(n) * 2).astype(np.float32)
        if trial == 0 and n > 10:
            x[3] = x[9] = x.max() + 0.5          # tie inside the top-k
        u = float(rng.random())
        T, top_p = 0.8, 0.9 if trial % 2 == 0 else 1.0
        tok, kept, margin = osampler.sample_topk(x, k, T, top_p, u)
        dx = D.from_numpy(x)
        chk(lib, lib.b2_op_sample_topk(dx.ptr, n, k, T, top_p, u, dout.ptr))
        got = int(dout.to_numpy(np.int32, (1,))[0])
        assert got in set(map(int, kept)) or margin < 1e-5
        if margin > 1e-5:                        # away from a CDF step the draw is identical
            assert got == tok
            checked += 1
    assert checked >= 8
    assert lib.b2_op_sample_topk(dout.ptr, n, 65, 0.8, 0.9, 0.5, dout.ptr) == -1


@pytest.mark.parametrize("n,k,last_n", [(1000, 40, 64), (32064, 20, 8), (300, 64, 500)])
def test_sample_topk_with_repeat_penalty(lib, D, n, k, last_n):
    """options.repeat_penalty (oracle/sampler.py:apply_repeat_penalty): tokens of the window lose their place among the candidates;
    a token repeated inside the window is penalised once; tokens before the window are not."""
    rng = np.random.default_rng(7 * n + k)
    dout = D(4)
    checked = 0
    for trial in range(10):
        x = (rng.standard_normal(n) * 2).astype(np.float32)
        top = np.argsort(-x)[:8]
        ctx = np.concatenate([rng.integers(0, n, size=100), top[:4], top[:2]]).astype(np.int32)   # the favourites sit in the window, two of them twice
        u = float(rng.random())
        pen = osampler.apply_repeat_penalty(x, ctx, 1.5, last_n)
        tok, kept, margin = osampler.sample_topk(pen, k, 0.8, 0.9, u)
        dx, dc = D.from_numpy(x), D.from_numpy(ctx)
        chk(lib, lib.b2_op_sample_topk_penalty(dx.ptr, n, k, 0.8, 0.9, u, dc.ptr, len(ctx), 1.5, last_n, dout.ptr))
        got = int(dout.to_numpy(np.int32, (1,))[0])
        if margin > 1e-5:
            assert got == tok
            checked += 1
        # penalty 1.0 and window 0 are both "off": the plain sampler's answer
        plain, _, pm = osampler.sample_topk(x, k, 0.8, 0.9, u)
        for pv, ln in ((1.0, last_n), (1.5, 0)):
            chk(lib, lib.b2_op_sample_topk_penalty(dx.ptr, n, k, 0.8, 0.9, u, dc.ptr, len(ctx), pv, ln, dout.ptr))
            assert pm < 1e-5 or int(dout.to_numpy(np.int32, (1,))[0]) == plain
    assert checked >= 6
