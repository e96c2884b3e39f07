"""GGUF checkpoints for the engine: what Ollama does when it opens a model blob
(`README.md:29-31` "install Ollama and pull the seven models"; the reference never touches the files itself).

`load(path)` reads a llama.cpp / Ollama GGUF file with the `gguf` package (third-party reader, present in the image) and
returns the `Arch` record, the tensor dictionary `Model.from_tensors` takes, and the tokenizer metadata.  Conventions of
llama.cpp's converters that differ from the engine's (which follows the Hugging Face layout the oracle is pinned to):
  * llama / mistral: the rows of attn_q and attn_k are permuted per head so that llama.cpp's adjacent-pair RoPE equals
    rotate-half; they are permuted back here (rows are quantised independently, so this is exact);
  * gemma: the converter already stored (1 + w) in the norm weights -> norm_plus_one is off for GGUF gemma;
  * llama 3.1: the RoPE frequency scaling travels as the tensor rope_freqs.weight -> the Llama-3.1 constants are assumed.
The engine's kernels take Q4_0, Q4_K and Q6_K matrices -- what llama.cpp's Q4_0 and Q4_K_S/M files contain (output.weight and a
tied token_embd.weight in Q6_K, Q4_K_M with Q6_K attn_v / ffn_down; q, k, v of different formats run as three GEMVs; gate and up
must share one format) -- and F32 vectors.  Models with Q6_K layer matrices evaluate the prompt token by token.  Other formats
(Q5_K, Q8_0, F16 matrices, ...) are refused unless `requantize=True`, which dequantises every non-Q4_0 matrix and re-quantises it to Q4_0
with the gguf package's own reference code (lossy; reported in `notes`)."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np

from .archs import Arch

Q4_0_BLOCK, Q4_0_BYTES = 32, 18
LLAMA3_ROPE = (8.0, 1.0, 4.0, 8192.0)      # factor, low_freq_factor, high_freq_factor, original context (Llama 3.1 model card)


@dataclass
class TokenizerData:
    model: str = ""                         # "gpt2" (byte-level BPE) or "llama" (SentencePiece-style BPE with scores)
    pre: str = ""
    tokens: List[str] = field(default_factory=list)
    scores: List[float] = field(default_factory=list)
    token_types: List[int] = field(default_factory=list)
    merges: List[str] = field(default_factory=list)
    bos_id: Optional[int] = None
    eos_id: Optional[int] = None
    add_bos: bool = True
    chat_template: str = ""


@dataclass
class GGUFModel:
    arch: Arch
    tensors: Dict[str, Tuple[str, np.ndarray]]
    tokenizer: TokenizerData
    notes: List[str]


def _field(reader, key, default=None):
    f = reader.fields.get(key)
    if f is None:
        return default
    return f.contents()


def _unpermute_rows(raw: np.ndarray, n_head: int) -> np.ndarray:
    """Inverse of convert_hf_to_gguf's permute(): rows [n_head][hd/2][2] -> [n_head][2][hd/2]."""
    rows = raw.shape[0]
    hd = rows // n_head
    return np.ascontiguousarray(raw.reshape(n_head, hd // 2, 2, *raw.shape[1:]).swapaxes(1, 2).reshape(raw.shape))


def permute_rows(raw: np.ndarray, n_head: int) -> np.ndarray:
    """convert_hf_to_gguf's permute() (used by the tests to write llama-style files)."""
    rows = raw.shape[0]
    hd = rows // n_head
    return np.ascontiguousarray(raw.reshape(n_head, 2, hd // 2, *raw.shape[1:]).swapaxes(1, 2).reshape(raw.shape))


_FAMILY = {"llama": "llama", "qwen2": "qwen2", "gemma": "gemma", "phi3": "phi3"}


def load(path: str, requantize: bool = False, name: Optional[str] = None) -> GGUFModel:
    import gguf
    from gguf import GGMLQuantizationType as QT
    reader = gguf.GGUFReader(path)
    notes: List[str] = []
    garch = _field(reader, "general.architecture")
    if garch not in _FAMILY:
        raise ValueError(f"GGUF architecture '{garch}' is not one of {sorted(_FAMILY)}")
    gname = name or _field(reader, "general.name", garch)
    key = lambda k, d=None: _field(reader, f"{garch}.{k}", d)
    H, L = int(key("embedding_length")), int(key("block_count"))
    nh = int(key("attention.head_count"))
    nkv = int(key("attention.head_count_kv", nh))
    hd = int(key("attention.key_length", H // nh))
    I = int(key("feed_forward_length"))
    eps = float(key("attention.layer_norm_rms_epsilon", 1e-5))
    theta = float(key("rope.freq_base", 10000.0))
    ctx = int(key("context_length", 2048))
    tensors_by_name = {t.name: t for t in reader.tensors}
    family = _FAMILY[garch]
    if family == "llama" and "mistral" in str(gname).lower():
        family = "mistral"
    tied = "output.weight" not in tensors_by_name
    emb = tensors_by_name.get("token_embd.weight")
    if emb is None:
        raise ValueError("GGUF file has no token_embd.weight")
    V = int(emb.shape[1])                     # GGUF shapes are fastest-dimension first: (cols, rows)
    rope_llama3 = LLAMA3_ROPE if "rope_freqs.weight" in tensors_by_name else None
    if rope_llama3:
        notes.append("rope_freqs.weight present: Llama-3.1 frequency scaling (8, 1, 4, 8192) assumed")
    arch = Arch(str(gname), family, H=H, L=L, nh=nh, nkv=nkv, hd=hd, I=I, V=V, tied=tied, rope_theta=theta, rms_eps=eps,
                qkv_bias=("blk.0.attn_q.bias" in tensors_by_name), act="gelu_tanh" if family == "gemma" else "silu",
                norm_plus_one=False, embed_scale=(family == "gemma"), rope_llama3=rope_llama3, max_ctx=min(ctx, 2048))
    if family == "gemma":
        notes.append("gemma: the converter stored 1+w in the norm weights; norm_plus_one is off")
    out: Dict[str, Tuple[str, np.ndarray]] = {}
    for t in reader.tensors:
        if t.name == "rope_freqs.weight":
            continue
        data = np.asarray(t.data)
        if len(t.shape) == 1:
            if t.tensor_type == QT.F32:
                vec = data.astype(np.float32)
            elif t.tensor_type == QT.F16:
                vec = data.astype(np.float32)
            else:
                raise ValueError(f"{t.name}: vector of type {t.tensor_type.name} is not supported")
            out[t.name] = ("f32", np.ascontiguousarray(vec.reshape(-1)))
            continue
        cols, rows = int(t.shape[0]), int(t.shape[1])
        qt_name = "q4_0"
        if t.tensor_type == QT.Q4_0:
            raw = np.ascontiguousarray(data).reshape(rows, cols // Q4_0_BLOCK * Q4_0_BYTES).view(np.uint8)
        elif t.tensor_type == QT.Q4_K and not requantize:      # native in the decode GEMV (csrc/layout.cuh)
            raw = np.ascontiguousarray(data).reshape(rows, cols // 256 * 144).view(np.uint8)
            qt_name = "q4_k"
        elif t.tensor_type == QT.Q6_K and not requantize:       # llama.cpp's format for the LM head, and Q4_K_M's for attn_v / ffn_down
            raw = np.ascontiguousarray(data).reshape(rows, cols // 256 * 210).view(np.uint8)
            qt_name = "q6_k"
        else:
            if not requantize:
                raise ValueError(f"{t.name} is {t.tensor_type.name}; the engine's matrices are Q4_0, Q4_K or Q6_K "
                                 f"(pass requantize=True to convert everything else to Q4_0, lossy)")
            if t.tensor_type in (QT.F32, QT.F16):
                dense = data.astype(np.float32).reshape(rows, cols)
            else:
                dense = gguf.quants.dequantize(data, t.tensor_type).astype(np.float32).reshape(rows, cols)
            raw = np.ascontiguousarray(gguf.quants.quantize(dense, QT.Q4_0)).reshape(rows, -1).view(np.uint8)
            notes.append(f"{t.name}: {t.tensor_type.name} -> Q4_0 (requantised, lossy)")
        if family in ("llama", "mistral"):
            if t.name.endswith("attn_q.weight"):
                raw = _unpermute_rows(raw, nh)
            elif t.name.endswith("attn_k.weight"):
                raw = _unpermute_rows(raw, nkv)
        out[t.name] = (qt_name, raw)
    if family in ("llama", "mistral") and arch.qkv_bias:
        for l in range(L):
            for nm, n in (("attn_q.bias", nh), ("attn_k.bias", nkv)):
                k = f"blk.{l}.{nm}"
                if k in out:
                    out[k] = ("f32", _unpermute_rows(out[k][1], n))
    tok = TokenizerData(
        model=str(_field(reader, "tokenizer.ggml.model", "")), pre=str(_field(reader, "tokenizer.ggml.pre", "")),
        tokens=list(_field(reader, "tokenizer.ggml.tokens", []) or []), scores=list(_field(reader, "tokenizer.ggml.scores", []) or []),
        token_types=list(_field(reader, "tokenizer.ggml.token_type", []) or []), merges=list(_field(reader, "tokenizer.ggml.merges", []) or []),
        bos_id=_field(reader, "tokenizer.ggml.bos_token_id"), eos_id=_field(reader, "tokenizer.ggml.eos_token_id"),
        add_bos=bool(_field(reader, "tokenizer.ggml.add_bos_token", True)), chat_template=str(_field(reader, "tokenizer.chat_template", "") or ""))
    return GGUFModel(arch, out, tok, notes)
