"""Architecture records of the seven model tags the reference sweeps
(`experiment/RunnerConfig.py:80`: llama3.1:8b, gemma:2b, gemma:7b, phi3:3.8b, qwen2:1.5b, qwen2:7b,
mistral:7b).  The reference only names Ollama tags; hyper-parameters are the upstream model cards'
(SURVEY.md §8d).  `to_c()` produces the `b2_arch_cfg` the C ABI takes."""
from __future__ import annotations

import ctypes
from dataclasses import dataclass, replace
from typing import Optional


@dataclass(frozen=True)
class Arch:
    name: str
    family: str
    H: int
    L: int
    nh: int
    nkv: int
    hd: int
    I: int
    V: int
    tied: bool
    rope_theta: float
    rms_eps: float
    qkv_bias: bool = False
    act: str = "silu"
    norm_plus_one: bool = False
    embed_scale: bool = False
    rope_llama3: Optional[tuple] = None
    max_ctx: int = 2048

    @property
    def q_dim(self) -> int:
        return self.nh * self.hd

    @property
    def kv_dim(self) -> int:
        return self.nkv * self.hd

    def matmul_weights(self) -> int:
        per_layer = (self.q_dim + 2 * self.kv_dim) * self.H + self.H * self.q_dim + 3 * self.I * self.H
        return self.L * per_layer + self.V * self.H

    def weight_bytes_q4(self) -> int:
        """Algorithmic bytes per decoded token at 4.5 bit/weight (Q4_0 18 B/32, Q4_K 144 B/256)."""
        return self.matmul_weights() * 18 // 32

    def kv_bytes_per_ctx_token(self) -> int:
        return self.L * 2 * self.kv_dim * 2


class CArchCfg(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in (
        "hidden", "n_layers", "n_heads", "n_kv_heads", "head_dim", "ffn", "vocab", "tied", "qkv_bias", "act",
        "norm_plus_one", "embed_scale", "rope_llama3", "max_ctx")] + [(n, ctypes.c_float) for n in (
            "rope_theta", "rms_eps", "rope_factor", "rope_low_freq", "rope_high_freq", "rope_orig_ctx")]


def to_c(a: Arch, max_ctx: Optional[int] = None) -> CArchCfg:
    f = a.rope_llama3 or (1.0, 1.0, 1.0, 1.0)
    return CArchCfg(a.H, a.L, a.nh, a.nkv, a.hd, a.I, a.V, int(a.tied), int(a.qkv_bias), 0 if a.act == "silu" else 1,
                    int(a.norm_plus_one), int(a.embed_scale), int(a.rope_llama3 is not None), max_ctx or a.max_ctx,
                    a.rope_theta, a.rms_eps, *f)


ARCHS = {
    "qwen2:1.5b": Arch("qwen2:1.5b", "qwen2", H=1536, L=28, nh=12, nkv=2, hd=128, I=8960, V=151936, tied=True,
                       rope_theta=1e6, rms_eps=1e-6, qkv_bias=True),
    "gemma:2b": Arch("gemma:2b", "gemma", H=2048, L=18, nh=8, nkv=1, hd=256, I=16384, V=256000, tied=True,
                     rope_theta=1e4, rms_eps=1e-6, act="gelu_tanh", norm_plus_one=True, embed_scale=True),
    "phi3:3.8b": Arch("phi3:3.8b", "phi3", H=3072, L=32, nh=32, nkv=32, hd=96, I=8192, V=32064, tied=False,
                      rope_theta=1e4, rms_eps=1e-5),
    "gemma:7b": Arch("gemma:7b", "gemma", H=3072, L=28, nh=16, nkv=16, hd=256, I=24576, V=256000, tied=True,
                     rope_theta=1e4, rms_eps=1e-6, act="gelu_tanh", norm_plus_one=True, embed_scale=True),
    "qwen2:7b": Arch("qwen2:7b", "qwen2", H=3584, L=28, nh=28, nkv=4, hd=128, I=18944, V=152064, tied=False,
                     rope_theta=1e6, rms_eps=1e-6, qkv_bias=True),
    "mistral:7b": Arch("mistral:7b", "mistral", H=4096, L=32, nh=32, nkv=8, hd=128, I=14336, V=32768, tied=False,
                       rope_theta=1e6, rms_eps=1e-5),
    "llama3.1:8b": Arch("llama3.1:8b", "llama", H=4096, L=32, nh=32, nkv=8, hd=128, I=14336, V=128256, tied=False,
                        rope_theta=5e5, rms_eps=1e-5, rope_llama3=(8.0, 1.0, 4.0, 8192.0)),
}

# structurally identical tiny variants (2 layers, real head dims) for fast parity tests
TINY = {
    "qwen2:1.5b": replace(ARCHS["qwen2:1.5b"], name="tiny-qwen2:1.5b", H=768, L=2, nh=6, nkv=1, hd=128, I=512, V=1000),
    "gemma:2b": replace(ARCHS["gemma:2b"], name="tiny-gemma:2b", H=256, L=2, nh=2, nkv=1, hd=256, I=512, V=1000),
    "phi3:3.8b": replace(ARCHS["phi3:3.8b"], name="tiny-phi3:3.8b", H=288, L=2, nh=3, nkv=3, hd=96, I=416, V=1000),
    "gemma:7b": replace(ARCHS["gemma:7b"], name="tiny-gemma:7b", H=256, L=2, nh=2, nkv=2, hd=256, I=768, V=1000),
    "qwen2:7b": replace(ARCHS["qwen2:7b"], name="tiny-qwen2:7b", H=896, L=2, nh=7, nkv=1, hd=128, I=1184, V=1000),
    "mistral:7b": replace(ARCHS["mistral:7b"], name="tiny-mistral:7b", H=512, L=2, nh=4, nkv=1, hd=128, I=1024, V=1000),
    "llama3.1:8b": replace(ARCHS["llama3.1:8b"], name="tiny-llama3.1:8b", H=512, L=2, nh=4, nkv=1, hd=128, I=1024, V=1024),
}

KIND = {"token_embd": 0, "output_norm": 1, "output": 2, "attn_norm": 3, "attn_q": 4, "attn_k": 5, "attn_v": 6,
        "attn_q_bias": 7, "attn_k_bias": 8, "attn_v_bias": 9, "attn_output": 10, "ffn_norm": 11, "ffn_gate": 12,
        "ffn_up": 13, "ffn_down": 14}
GLOBAL_LAYER = 0xFFFF


def tensor_id(kind: str, layer: int = GLOBAL_LAYER) -> int:
    return layer * 16 + KIND[kind]


def tensor_table(a: Arch):
    """[(gguf_name, kind, layer, rows, cols)]; cols == 0 marks an f32 vector."""
    t = [("token_embd.weight", "token_embd", GLOBAL_LAYER, a.V, a.H)]
    for i in range(a.L):
        p = f"blk.{i}."
        t += [(p + "attn_norm.weight", "attn_norm", i, a.H, 0),
              (p + "attn_q.weight", "attn_q", i, a.q_dim, a.H),
              (p + "attn_k.weight", "attn_k", i, a.kv_dim, a.H),
              (p + "attn_v.weight", "attn_v", i, a.kv_dim, a.H)]
        if a.qkv_bias:
            t += [(p + "attn_q.bias", "attn_q_bias", i, a.q_dim, 0),
                  (p + "attn_k.bias", "attn_k_bias", i, a.kv_dim, 0),
                  (p + "attn_v.bias", "attn_v_bias", i, a.kv_dim, 0)]
        t += [(p + "attn_output.weight", "attn_output", i, a.H, a.q_dim),
              (p + "ffn_norm.weight", "ffn_norm", i, a.H, 0),
              (p + "ffn_gate.weight", "ffn_gate", i, a.I, a.H),
              (p + "ffn_up.weight", "ffn_up", i, a.I, a.H),
              (p + "ffn_down.weight", "ffn_down", i, a.H, a.I)]
    t.append(("output_norm.weight", "output_norm", GLOBAL_LAYER, a.H, 0))
    if not a.tied:
        t.append(("output.weight", "output", GLOBAL_LAYER, a.V, a.H))
    return t


def init_recipe(a: Arch, kind: str):
    """(sigma, offset) of the deterministic random initialiser for a tensor kind."""
    if kind in ("attn_norm", "ffn_norm", "output_norm"):
        return 0.1, (0.0 if a.norm_plus_one else 1.0)
    if kind.endswith("_bias"):
        return 0.1, 0.0
    return 0.02, 0.0
