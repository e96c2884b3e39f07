"""ORACLE (test infrastructure): the seven model architectures the reference sweeps
(`experiment/RunnerConfig.py:80`, `README.md:30`) as hyper-parameter records, plus structurally
identical tiny variants for fast CPU tests.

The reference names only Ollama tags; the hyper-parameters come from the upstream model cards
(SURVEY.md §8d table).  This table is an independent restatement of the product's
`archs.py`; `tests/test_archs.py` checks the two agree and checks the weight-byte counts against
SURVEY.md §8d.
"""
from __future__ import annotations

from dataclasses import dataclass, field, replace
from typing import Optional


@dataclass(frozen=True)
class Arch:
    name: str
    family: str            # llama | mistral | qwen2 | gemma | phi3
    H: int                 # hidden size
    L: int                 # layers
    nh: int                # query heads
    nkv: int               # key/value heads
    hd: int                # head dim
    I: int                 # MLP intermediate size
    V: int                 # vocabulary
    tied: bool             # LM head shares the embedding matrix
    rope_theta: float
    rms_eps: float
    qkv_bias: bool = False
    act: str = "silu"      # silu | gelu_tanh
    norm_plus_one: bool = False   # gemma: y = x_hat * (1 + w)
    embed_scale: bool = False     # gemma: h0 = embed * sqrt(H)
    rope_llama3: Optional[tuple] = None  # (factor, low_freq_factor, high_freq_factor, original_ctx)
    max_ctx: int = 2048    # Ollama default num_ctx

    @property
    def q_dim(self) -> int:
        return self.nh * self.hd

    @property
    def kv_dim(self) -> int:
        return self.nkv * self.hd

    def matmul_weights(self) -> int:
        """Number of weights read once per decoded token (all layers + LM head)."""
        per_layer = (self.q_dim + 2 * self.kv_dim) * self.H + self.H * self.q_dim + 3 * self.I * self.H
        return self.L * per_layer + self.V * self.H

    def kv_bytes_per_ctx_token(self) -> int:
        return self.L * 2 * self.kv_dim * 2  # f16 K and V


ARCHS = {
    "qwen2:1.5b": Arch("qwen2:1.5b", "qwen2", 1536, 28, 12, 2, 128, 8960, 151936, True, 1e6, 1e-6, qkv_bias=True),
    "gemma:2b": Arch("gemma:2b", "gemma", 2048, 18, 8, 1, 256, 16384, 256000, True, 1e4, 1e-6,
                     act="gelu_tanh", norm_plus_one=True, embed_scale=True),
    "phi3:3.8b": Arch("phi3:3.8b", "phi3", 3072, 32, 32, 32, 96, 8192, 32064, False, 1e4, 1e-5),
    "gemma:7b": Arch("gemma:7b", "gemma", 3072, 28, 16, 16, 256, 24576, 256000, True, 1e4, 1e-6,
                     act="gelu_tanh", norm_plus_one=True, embed_scale=True),
    "qwen2:7b": Arch("qwen2:7b", "qwen2", 3584, 28, 28, 4, 128, 18944, 152064, False, 1e6, 1e-6, qkv_bias=True),
    "mistral:7b": Arch("mistral:7b", "mistral", 4096, 32, 32, 8, 128, 14336, 32768, False, 1e6, 1e-5),
    "llama3.1:8b": Arch("llama3.1:8b", "llama", 4096, 32, 32, 8, 128, 14336, 128256, False, 5e5, 1e-5,
                        rope_llama3=(8.0, 1.0, 4.0, 8192)),
}

# Tiny variants: same family switches, real head dims, 2 layers. K dims are multiples of 256 where the
# family is also served as Q4_K (llama) so every quant type can be exercised.
TINY = {
    "qwen2:1.5b": replace(ARCHS["qwen2:1.5b"], name="tiny-qwen2:1.5b", H=768, L=2, nh=6, nkv=1, hd=128, I=512, V=1000),
    "gemma:2b": replace(ARCHS["gemma:2b"], name="tiny-gemma:2b", H=256, L=2, nh=2, nkv=1, hd=256, I=512, V=1000),
    "phi3:3.8b": replace(ARCHS["phi3:3.8b"], name="tiny-phi3:3.8b", H=288, L=2, nh=3, nkv=3, hd=96, I=416, V=1000),
    "gemma:7b": replace(ARCHS["gemma:7b"], name="tiny-gemma:7b", H=256, L=2, nh=2, nkv=2, hd=256, I=768, V=1000),
    "qwen2:7b": replace(ARCHS["qwen2:7b"], name="tiny-qwen2:7b", H=896, L=2, nh=7, nkv=1, hd=128, I=1184, V=1000),
    "mistral:7b": replace(ARCHS["mistral:7b"], name="tiny-mistral:7b", H=512, L=2, nh=4, nkv=1, hd=128, I=1024, V=1000),
    "llama3.1:8b": replace(ARCHS["llama3.1:8b"], name="tiny-llama3.1:8b", H=512, L=2, nh=4, nkv=1, hd=128, I=1024, V=1024),
}

# Tensor kinds: the id feeding the deterministic random initialiser (oracle/randinit.py) and the
# GGUF-style name each maps to.
KIND = {
    "token_embd": 0, "output_norm": 1, "output": 2,
    "attn_norm": 3, "attn_q": 4, "attn_k": 5, "attn_v": 6,
    "attn_q_bias": 7, "attn_k_bias": 8, "attn_v_bias": 9,
    "attn_output": 10, "ffn_norm": 11, "ffn_gate": 12, "ffn_up": 13, "ffn_down": 14,
}
GLOBAL_LAYER = 0xFFFF


def tensor_id(kind: str, layer: int = GLOBAL_LAYER) -> int:
    return layer * 16 + KIND[kind]


def tensor_table(a: Arch):
    """[(name, kind, layer, rows, cols)] for every tensor of the model; cols == 0 marks a vector."""
    t = [("token_embd.weight", "token_embd", GLOBAL_LAYER, a.V, a.H)]
    for i in range(a.L):
        p = f"blk.{i}."
        t.append((p + "attn_norm.weight", "attn_norm", i, a.H, 0))
        t.append((p + "attn_q.weight", "attn_q", i, a.q_dim, a.H))
        t.append((p + "attn_k.weight", "attn_k", i, a.kv_dim, a.H))
        t.append((p + "attn_v.weight", "attn_v", i, a.kv_dim, a.H))
        if a.qkv_bias:
            t.append((p + "attn_q.bias", "attn_q_bias", i, a.q_dim, 0))
            t.append((p + "attn_k.bias", "attn_k_bias", i, a.kv_dim, 0))
            t.append((p + "attn_v.bias", "attn_v_bias", i, a.kv_dim, 0))
        t.append((p + "attn_output.weight", "attn_output", i, a.H, a.q_dim))
        t.append((p + "ffn_norm.weight", "ffn_norm", i, a.H, 0))
        t.append((p + "ffn_gate.weight", "ffn_gate", i, a.I, a.H))
        t.append((p + "ffn_up.weight", "ffn_up", i, a.I, a.H))
        t.append((p + "ffn_down.weight", "ffn_down", i, a.H, a.I))
    t.append(("output_norm.weight", "output_norm", GLOBAL_LAYER, a.H, 0))
    if not a.tied:
        t.append(("output.weight", "output", GLOBAL_LAYER, a.V, a.H))
    return t
