"""GGUF checkpoints and tokenizers.  No real checkpoint exists offline: the files are written here with the
gguf package in llama.cpp's conventions (permuted q/k rows for llama, rope_freqs tensor, fastest-dimension-first shapes) and
must come back as exactly the tensors they were made from."""
import numpy as np
import pytest

from cain_2025_device_remote_llm_energy_rep_pkg_b200 import archs, gguf_io, tokenizer
from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import host_random_tensors

TOKENS = ["<unk>", "<s>", "</s>"] + ["<0x%02X>" % b for b in range(256)]


def write_gguf(path, a, tensors, garch, tok=None, dense_output=False):
    import gguf
    from gguf import GGMLQuantizationType as QT
    w = gguf.GGUFWriter(str(path), garch)
    w.add_name(a.name)
    w.add_block_count(a.L)
    w.add_embedding_length(a.H)
    w.add_feed_forward_length(a.I)
    w.add_head_count(a.nh)
    w.add_head_count_kv(a.nkv)
    w.add_key_length(a.hd)
    w.add_layer_norm_rms_eps(a.rms_eps)
    w.add_rope_freq_base(a.rope_theta)
    w.add_context_length(a.max_ctx)
    if tok:
        w.add_tokenizer_model(tok["model"])
        w.add_token_list(tok["tokens"])
        if "scores" in tok:
            w.add_token_scores(tok["scores"])
        if "types" in tok:
            w.add_token_types(tok["types"])
        if "merges" in tok:
            w.add_token_merges(tok["merges"])
        w.add_bos_token_id(tok["bos"])
        w.add_eos_token_id(tok["eos"])
    for name, (qt, arr) in tensors.items():
        arr = np.ascontiguousarray(arr)
        if qt == "f32":
            if garch == "llama" and name.endswith(("attn_q.bias", "attn_k.bias")):
                arr = gguf_io.permute_rows(arr, a.nh if "attn_q" in name else a.nkv)
            w.add_tensor(name, arr.astype(np.float32))
            continue
        rows = arr.shape[0]
        cols = arr.shape[1] // 18 * 32
        if garch == "llama" and name.endswith("attn_q.weight"):
            arr = gguf_io.permute_rows(arr, a.nh)
        if garch == "llama" and name.endswith("attn_k.weight"):
            arr = gguf_io.permute_rows(arr, a.nkv)
        if dense_output and name == "output.weight":      # a matrix that is not Q4_0 (llama.cpp keeps output.weight in Q6_K even in "Q4_0" files)
            dense = gguf.quants.dequantize(arr, QT.Q4_0)
            w.add_tensor(name, dense.astype(np.float32), raw_dtype=QT.F32)
            continue
        w.add_tensor(name, arr, raw_shape=(rows, arr.shape[1]), raw_dtype=QT.Q4_0)
        assert cols % 32 == 0
    if a.rope_llama3:
        w.add_tensor("rope_freqs.weight", np.ones(a.hd // 2, dtype=np.float32))
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()


@pytest.mark.parametrize("key,garch", [("llama3.1:8b", "llama"), ("qwen2:1.5b", "qwen2"), ("phi3:3.8b", "phi3")])
def test_gguf_round_trip(tmp_path, lib, key, garch):
    a = archs.TINY[key]
    tensors = host_random_tensors(a, seed=3)
    p = tmp_path / "m.gguf"
    write_gguf(p, a, tensors, garch)
    g = gguf_io.load(str(p))
    b = g.arch
    assert (b.H, b.L, b.nh, b.nkv, b.hd, b.I, b.V, b.tied, b.qkv_bias) == (a.H, a.L, a.nh, a.nkv, a.hd, a.I, a.V, a.tied, a.qkv_bias)
    assert b.family == a.family and abs(b.rope_theta - a.rope_theta) < 1e-3 * a.rope_theta and (b.rope_llama3 is None) == (a.rope_llama3 is None)
    assert set(g.tensors) == set(tensors)
    for name, (qt, arr) in tensors.items():
        qt2, arr2 = g.tensors[name]
        assert qt2 == qt and np.array_equal(np.ascontiguousarray(arr).reshape(arr2.shape), arr2), name


def test_gguf_refuses_or_requantizes_other_formats(tmp_path, lib):
    a = archs.TINY["mistral:7b"]
    tensors = host_random_tensors(a, seed=4)
    p = tmp_path / "m.gguf"
    write_gguf(p, a, tensors, "llama", dense_output=True)
    with pytest.raises(ValueError) as e:
        gguf_io.load(str(p))
    assert "output.weight" in str(e.value) and "requantize" in str(e.value)
    g = gguf_io.load(str(p), requantize=True)
    assert any("output.weight" in n for n in g.notes)
    # Q4_0 -> f32 -> Q4_0 with the same reference quantiser reproduces the blocks
    assert np.array_equal(g.tensors["output.weight"][1], np.ascontiguousarray(tensors["output.weight"][1]).reshape(a.V, -1))
    assert g.arch.family == "mistral"


def test_permute_is_the_converters_and_inverts():
    raw = np.arange(4 * 8 * 3, dtype=np.uint8).reshape(32, 3)      # 4 heads, head dim 8
    p = gguf_io.permute_rows(raw, 4)
    # convert_hf_to_gguf.permute: reshape(n_head, 2, hd/2, cols).swapaxes(1, 2): HF row (h, half, j) -> GGUF row (h, j, half)
    assert np.array_equal(p[1], raw[4]) and np.array_equal(p[2], raw[1]) and np.array_equal(p[8 + 3], raw[8 + 5])
    assert np.array_equal(gguf_io._unpermute_rows(p, 4), raw)


def test_spm_tokenizer_from_gguf_metadata():
    pieces = ["▁", "h", "e", "l", "o", "▁h", "ll", "he", "▁he", "llo", "▁hello", "w", "r", "d", "▁w", "or", "▁wor", "ld", "▁world"]
    toks = TOKENS + pieces
    scores = [0.0] * len(TOKENS) + [-float(i) for i in range(len(pieces), 0, -1)]    # later pieces score higher
    data = gguf_io.TokenizerData(model="llama", tokens=toks, scores=scores, token_types=[2, 3, 3] + [6] * 256 + [1] * len(pieces),
                                 bos_id=1, eos_id=2, add_bos=True)
    tk = tokenizer.GGUFTokenizer(data, family="")
    ids = tk.encode("hello world", template=False)
    assert ids == [1, toks.index("▁hello"), toks.index("▁world")]
    assert tk.decode(ids[1:]) == " hello world"
    ids = tk.encode("hé", template=False)                       # é is not in the vocabulary: byte fallback
    assert ids[:2] == [1, toks.index("▁h")] and ids[2:] == [toks.index("<0xC3>"), toks.index("<0xA9>")]
    assert tk.decode(ids[1:]) == " hé"
    assert tk.is_eos(2) and not tk.is_eos(5)


def test_bpe_tokenizer_from_gguf_metadata():
    b2u = tokenizer._bytes_to_unicode()
    base = [b2u[b] for b in range(256)]
    sp = b2u[ord(" ")]
    merges = ["h e", "l l", "he ll", "hell o", f"{sp} w", f"{sp}w o", "r l", f"{sp}wo rl", f"{sp}worl d"]
    toks = base + ["".join(m.split(" ")) for m in merges] + ["<|begin|>", "<|eot|>"]
    types = [1] * (len(toks) - 2) + [3, 3]
    data = gguf_io.TokenizerData(model="gpt2", tokens=toks, merges=merges, token_types=types, bos_id=len(toks) - 2, eos_id=len(toks) - 1)
    tk = tokenizer.GGUFTokenizer(data, family="")
    ids = tk.encode("hello world", template=False)
    assert ids == [len(toks) - 2, toks.index("hello"), toks.index(sp + "world")]
    assert tk.decode(ids[1:]) == "hello world"
    ids = tk.encode("hello<|eot|>hé", template=False)              # special tokens are matched verbatim, other bytes singly
    assert ids[1:3] == [toks.index("hello"), len(toks) - 1] and tk.decode(ids[3:]) == "hé"


@pytest.mark.gpu
def test_gguf_model_runs_like_the_tensors_it_came_from(tmp_path, gpu_engine):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import engine
    a = archs.TINY["llama3.1:8b"]
    tensors = host_random_tensors(a, seed=6)
    p = tmp_path / "m.gguf"
    write_gguf(p, a, tensors, "llama")
    g = gguf_io.load(str(p))
    prompt = [1, 17, 400, 23, 999, 5]
    l1 = engine.Model.from_tensors(gpu_engine, a, tensors, max_ctx=64).session(64).forward_tokens(prompt)
    l2 = engine.Model.from_tensors(gpu_engine, g.arch, g.tensors, max_ctx=64).session(64).forward_tokens(prompt)
    assert np.array_equal(l1, l2)
