"""ORACLE (test infrastructure, never on the product path): a plain PyTorch CPU forward of the seven
architectures, the checker BASELINE.json's north_star names as the fallback when Ollama's CPU
backend cannot be installed offline (it cannot: SURVEY.md §8c).

What it restates: the stages of Ollama's `/api/generate` that the reference spends its time in
(call site `experiment/RunnerConfig.py:128-131`; stage list SURVEY.md §3.3 / §8 a3.1-a3.8):
embed -> per layer [RMSNorm -> QKV (+bias) -> RoPE -> KV append -> attention -> O -> +res ->
RMSNorm -> gate/up -> act*mul -> down -> +res] -> final norm -> LM head -> greedy argmax.
Architecture semantics (gemma (1+w) norm and sqrt(H) embed scale, qwen2 QKV bias, llama3 rope
scaling, tied heads, GQA/MQA/MHA) are cross-checked against the Hugging Face model classes in
`tests/golden/make_hf_golden.py`; those fixtures are the pin for this file.

Numerics contract (DESIGN.md "Numerics"), identical on the CUDA path:
  * weights: exact dequantisation of the GGML blocks (oracle/quants.py), fp32;
  * activations are fp32 between ops; the vector entering every quantised matmul is first rounded
    to a block-fixed-point grid: per 32 elements, step 2^(e-19) with e = max(floor(log2 max|x|), -100) (-100 for an all-zero block), i.e. 21
    significant bits relative to the block maximum (`fxp_round`).  The CUDA kernels carry that
    integer as three int8 digits and do exact integer dot products; here it is a rounding;
  * KV cache holds f16 (round-to-nearest-even) keys (after RoPE) and values;
  * RoPE cos/sin tables are computed in float64 and rounded to fp32;
  * everything else (norms, softmax, activations, accumulation) is fp32; only summation order and
    libm's expf/tanhf differ from the GPU (about 1e-6 relative).
Parity status: architecture pinned against HF fixtures; formats pinned against gguf-py; the
activation grid is this project's own contract (no counterpart in the unpinned Ollama dependency).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

from . import quants
from .archs import Arch

ACT_BITS = 20  # |X| <= 2^20


def fxp_round(x: torch.Tensor, bits: int = ACT_BITS) -> torch.Tensor:
    """Round each block of 32 to the grid 2^(e-bits+1), e = max(floor(log2(max|x|)), -100)."""
    if bits <= 0:
        return x
    xb = x.reshape(-1, 32)
    m = xb.abs().amax(dim=1, keepdim=True)
    _, ex = torch.frexp(m)                       # m = f * 2^ex, f in [0.5, 1)
    e = torch.clamp(torch.where(m == 0, torch.full_like(ex, -100), ex - 1), min=-100)
    s = torch.ldexp(torch.ones_like(m), e - (bits - 1))
    return (torch.round(xb / s) * s).reshape(x.shape)


def fxp_digits(x: np.ndarray, bits: int = ACT_BITS):
    """The integer view of `fxp_round`: (digits int8[3, n], exponent int32[n/32]) with
    X = d2*16384 + d1*128 + d0 (balanced base 128) and x_hat = X * 2^exponent."""
    t = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).reshape(-1, 32)
    m = t.abs().amax(dim=1, keepdim=True)
    _, ex = torch.frexp(m)
    e = torch.clamp(torch.where(m == 0, torch.full_like(ex, -100), ex - 1), min=-100) - (bits - 1)
    s = torch.ldexp(torch.ones_like(m), e)
    X = torch.round(t / s).to(torch.int64).numpy().reshape(-1)
    d0 = ((X + 64) & 127) - 64
    X1 = (X - d0) >> 7
    d1 = ((X1 + 64) & 127) - 64
    d2 = (X1 - d1) >> 7
    return np.stack([d0, d1, d2]).astype(np.int8), e.reshape(-1).to(torch.int32).numpy()


def rope_inv_freq(a: Arch) -> np.ndarray:
    """float64[hd/2]; llama3 frequency scaling as in the Llama-3.1 model card."""
    j = np.arange(0, a.hd, 2, dtype=np.float64)
    inv = 1.0 / (a.rope_theta ** (j / a.hd))
    if a.rope_llama3 is not None:
        factor, lo, hi, orig = a.rope_llama3
        wavelen = 2.0 * math.pi / inv
        lo_wl, hi_wl = orig / lo, orig / hi
        scaled = np.where(wavelen > lo_wl, inv / factor, inv)
        smooth = (orig / wavelen - lo) / (hi - lo)
        mid = (1.0 - smooth) * inv / factor + smooth * inv
        is_mid = ~(wavelen < hi_wl) & ~(wavelen > lo_wl)
        inv = np.where(is_mid, mid, scaled)
    return inv


def rope_tables(a: Arch, max_ctx: int):
    ang = np.arange(max_ctx, dtype=np.float64)[:, None] * rope_inv_freq(a)[None, :]
    return (torch.from_numpy(np.cos(ang).astype(np.float32)), torch.from_numpy(np.sin(ang).astype(np.float32)))


class QLinear:
    """y = W x for a GGML-quantised W [rows, cols]; the input is put on the activation grid first."""

    def __init__(self, qtype: str, blob: np.ndarray, rows: int, cols: int, cache: bool, act_bits: int):
        self.qtype, self.rows, self.cols, self.act_bits = qtype, rows, cols, act_bits
        self.blob = np.ascontiguousarray(blob).reshape(rows, -1)
        self._w: Optional[torch.Tensor] = None
        self.cache = cache

    def dense(self, r0: int = 0, r1: Optional[int] = None) -> torch.Tensor:
        if self._w is not None:
            return self._w[r0:r1]
        b = self.blob[r0:r1]
        w = b if self.qtype == "f32" else quants.dequantize(b, self.qtype)
        w = torch.from_numpy(np.ascontiguousarray(w, dtype=np.float32))
        if self.cache and r0 == 0 and r1 is None:
            self._w = w
        return w

    def row(self, r: int) -> torch.Tensor:
        return self.dense(r, r + 1)[0].clone()

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        xq = fxp_round(x, self.act_bits)
        if self._w is not None or self.cache:
            return self.dense() @ xq
        out = torch.empty(self.rows, dtype=torch.float32)
        step = 4096
        for r0 in range(0, self.rows, step):
            out[r0:r0 + step] = self.dense(r0, min(r0 + step, self.rows)) @ xq
        return out


def _cat_rows(parts):
    qts = {p[0] for p in parts}
    assert len(qts) == 1
    return qts.pop(), np.concatenate([np.ascontiguousarray(p[1]).reshape(p[1].shape[0], -1) for p in parts], axis=0)


class QCat:
    """Row-wise concatenation of matrices in DIFFERENT block formats (llama.cpp's Q4_K_M keeps attn_v in Q6_K next to Q4_K
    attn_q / attn_k): y = [W0 x; W1 x; ...]."""

    def __init__(self, parts):
        self.parts = parts

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        return torch.cat([p(x) for p in self.parts])


class OracleModel:
    def __init__(self, arch: Arch, tensors: dict, cache_dense: bool = True, act_bits: int = ACT_BITS):
        a = self.a = arch
        self.act_bits = act_bits
        g = lambda n: tensors[n]
        vec = lambda n: torch.from_numpy(np.asarray(tensors[n][1], dtype=np.float32).copy())
        lin = lambda qt, blob, r, c: QLinear(qt, blob, r, c, cache_dense, act_bits)
        self.embd = lin(*g("token_embd.weight"), a.V, a.H)
        self.head = self.embd if a.tied else lin(*g("output.weight"), a.V, a.H)
        self.out_norm = vec("output_norm.weight")
        self.layers = []
        for i in range(a.L):
            p = f"blk.{i}."
            qkv_lin = None
            if p + "attn_qkv.weight" in tensors:
                qkv = g(p + "attn_qkv.weight")
            else:
                parts = [g(p + "attn_q.weight"), g(p + "attn_k.weight"), g(p + "attn_v.weight")]
                if len({t[0] for t in parts}) == 1:
                    qkv = _cat_rows(parts)
                else:
                    qkv = None
                    qkv_lin = QCat([lin(*t, r, a.H) for t, r in zip(parts, (a.q_dim, a.kv_dim, a.kv_dim))])
            bias = None
            if a.qkv_bias:
                bias = torch.cat([vec(p + "attn_q.bias"), vec(p + "attn_k.bias"), vec(p + "attn_v.bias")])
            self.layers.append(dict(
                attn_norm=vec(p + "attn_norm.weight"), qkv=qkv_lin if qkv_lin is not None else lin(*qkv, a.q_dim + 2 * a.kv_dim, a.H), qkv_bias=bias,
                o=lin(*g(p + "attn_output.weight"), a.H, a.q_dim), ffn_norm=vec(p + "ffn_norm.weight"),
                gate=lin(*g(p + "ffn_gate.weight"), a.I, a.H), up=lin(*g(p + "ffn_up.weight"), a.I, a.H),
                down=lin(*g(p + "ffn_down.weight"), a.H, a.I)))

    def session(self, max_ctx: Optional[int] = None, kv_dtype=torch.float16) -> "OracleSession":
        return OracleSession(self, max_ctx or self.a.max_ctx, kv_dtype)


class OracleSession:
    def __init__(self, model: OracleModel, max_ctx: int, kv_dtype=torch.float16):
        self.m, a = model, model.a
        self.max_ctx, self.pos, self.kv_dtype = max_ctx, 0, kv_dtype
        self.k = [torch.zeros(max_ctx, a.nkv, a.hd, dtype=kv_dtype) for _ in range(a.L)]
        self.v = [torch.zeros(max_ctx, a.nkv, a.hd, dtype=kv_dtype) for _ in range(a.L)]
        self.cos, self.sin = rope_tables(a, max_ctx)

    def reset(self):
        self.pos = 0

    def _norm(self, x, w):
        a = self.m.a
        var = (x * x).mean()
        xh = x * (1.0 / torch.sqrt(var + np.float32(a.rms_eps)))
        return xh * (1.0 + w) if a.norm_plus_one else xh * w

    def _rope(self, t, pos):  # t [heads, hd], rotate-half convention
        h = self.m.a.hd // 2
        c, s = self.cos[pos], self.sin[pos]
        t1, t2 = t[:, :h], t[:, h:]
        return torch.cat([t1 * c - t2 * s, t2 * c + t1 * s], dim=1)

    def _act(self, g):
        if self.m.a.act == "silu":
            return g / (1.0 + torch.exp(-g))
        return 0.5 * g * (1.0 + torch.tanh(np.float32(0.7978845608028654) * (g + np.float32(0.044715) * g * g * g)))

    def hidden(self, token: int) -> torch.Tensor:
        """Advance one position; returns the final-norm hidden state (before the LM head)."""
        m, a, p = self.m, self.m.a, self.pos
        assert p < self.max_ctx, "context full"
        x = m.embd.row(int(token))
        if a.embed_scale:
            x = x * np.float32(math.sqrt(a.H))
        grp = a.nh // a.nkv
        for li, ly in enumerate(m.layers):
            qkv = ly["qkv"](self._norm(x, ly["attn_norm"]))
            if ly["qkv_bias"] is not None:
                qkv = qkv + ly["qkv_bias"]
            q = self._rope(qkv[:a.q_dim].reshape(a.nh, a.hd), p)
            k = self._rope(qkv[a.q_dim:a.q_dim + a.kv_dim].reshape(a.nkv, a.hd), p)
            v = qkv[a.q_dim + a.kv_dim:].reshape(a.nkv, a.hd)
            self.k[li][p] = k.to(self.kv_dtype)
            self.v[li][p] = v.to(self.kv_dtype)
            K = self.k[li][:p + 1].float()                      # [T, nkv, hd]
            Vv = self.v[li][:p + 1].float()
            qg = q.reshape(a.nkv, grp, a.hd)
            sc = torch.einsum("ngd,tnd->ngt", qg, K) * np.float32(1.0 / math.sqrt(a.hd))
            pr = torch.softmax(sc, dim=-1)
            att = torch.einsum("ngt,tnd->ngd", pr, Vv).reshape(a.q_dim)
            x = x + ly["o"](att)
            hn = self._norm(x, ly["ffn_norm"])
            x = x + ly["down"](self._act(ly["gate"](hn)) * ly["up"](hn))
        self.pos += 1
        return self._norm(x, m.out_norm)

    def step(self, token: int) -> torch.Tensor:
        return self.m.head(self.hidden(token))

    def prefill(self, tokens) -> torch.Tensor:
        """Token-by-token prefill (the same math as a causal batched forward); logits of the last."""
        for t in tokens[:-1]:
            self.hidden(int(t))
        return self.step(int(tokens[-1]))

    def generate_greedy(self, prompt, n_new: int, keep_logits: bool = False):
        logits = self.prefill(prompt)
        out, keep = [], []
        for i in range(n_new):
            tok = int(torch.argmax(logits))
            out.append(tok)
            if keep_logits:
                keep.append(logits.clone())
            if i + 1 < n_new:
                logits = self.step(tok)
        return (out, keep) if keep_logits else out
