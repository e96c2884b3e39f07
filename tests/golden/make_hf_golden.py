"""Generates tests/golden/hf_logits.npz: logits of the Hugging Face model classes (transformers, CPU,
fp32, eager attention) for the tiny variant of each of the seven architectures, with OUR deterministic
random-init weights (oracle/randinit.py, unquantised fp32) loaded into them.  tests/test_oracle_model.py
rebuilds the same weights from the seed and checks oracle/model.py (act grid off, fp32 weights) against
these logits: this pins the oracle's architecture semantics to an independent implementation.
Run once in the build container:  python tests/golden/make_hf_golden.py"""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import numpy as np, torch
import transformers
from transformers import (LlamaConfig, LlamaForCausalLM, MistralConfig, MistralForCausalLM, Qwen2Config,
                          Qwen2ForCausalLM, GemmaConfig, GemmaForCausalLM, Phi3Config, Phi3ForCausalLM)
from oracle.archs import TINY
from oracle import randinit

SEED = 7
TOKENS = [3, 17, 999, 0, 512, 77, 78, 79, 3, 3, 640, 5]

def hf_model(a):
    common = dict(hidden_size=a.H, intermediate_size=a.I, num_hidden_layers=a.L, num_attention_heads=a.nh,
                  num_key_value_heads=a.nkv, vocab_size=a.V, rms_norm_eps=a.rms_eps,
                  max_position_embeddings=(131072 if a.family == "llama" else 8192), tie_word_embeddings=a.tied, attn_implementation="eager")
    rope = dict(rope_theta=a.rope_theta)
    if a.family == "llama":
        f, lo, hi, orig = a.rope_llama3
        cfg = LlamaConfig(**common, head_dim=a.hd, rope_parameters=dict(rope_type="llama3", rope_theta=a.rope_theta, factor=f,
                          low_freq_factor=lo, high_freq_factor=hi, original_max_position_embeddings=orig))
        return LlamaForCausalLM(cfg)
    if a.family == "mistral":
        return MistralForCausalLM(MistralConfig(**common, head_dim=a.hd, sliding_window=None,
                                                rope_parameters=dict(rope_type="default", rope_theta=a.rope_theta)))
    if a.family == "qwen2":
        return Qwen2ForCausalLM(Qwen2Config(**common, rope_parameters=dict(rope_type="default", rope_theta=a.rope_theta)))
    if a.family == "gemma":
        return GemmaForCausalLM(GemmaConfig(**common, head_dim=a.hd, hidden_activation="gelu_pytorch_tanh",
                                            rope_parameters=dict(rope_type="default", rope_theta=a.rope_theta)))
    if a.family == "phi3":
        return Phi3ForCausalLM(Phi3Config(**common, pad_token_id=0,
                                          rope_parameters=dict(rope_type="default", rope_theta=a.rope_theta)))
    raise ValueError(a.family)

def load(a, model, t):
    sd = {}
    T = lambda n: torch.from_numpy(np.asarray(t[n][1], dtype=np.float32))
    sd["model.embed_tokens.weight"] = T("token_embd.weight")
    sd["model.norm.weight"] = T("output_norm.weight")
    if not a.tied:
        sd["lm_head.weight"] = T("output.weight")
    for i in range(a.L):
        p, h = f"blk.{i}.", f"model.layers.{i}."
        sd[h + "input_layernorm.weight"] = T(p + "attn_norm.weight")
        sd[h + "post_attention_layernorm.weight"] = T(p + "ffn_norm.weight")
        sd[h + "self_attn.o_proj.weight"] = T(p + "attn_output.weight")
        sd[h + "mlp.down_proj.weight"] = T(p + "ffn_down.weight")
        if a.family == "phi3":
            sd[h + "self_attn.qkv_proj.weight"] = torch.cat([T(p + "attn_q.weight"), T(p + "attn_k.weight"), T(p + "attn_v.weight")])
            sd[h + "mlp.gate_up_proj.weight"] = torch.cat([T(p + "ffn_gate.weight"), T(p + "ffn_up.weight")])
        else:
            for s, d in (("q", "q_proj"), ("k", "k_proj"), ("v", "v_proj")):
                sd[h + f"self_attn.{d}.weight"] = T(p + f"attn_{s}.weight")
                if a.qkv_bias:
                    sd[h + f"self_attn.{d}.bias"] = T(p + f"attn_{s}.bias")
            sd[h + "mlp.gate_proj.weight"] = T(p + "ffn_gate.weight")
            sd[h + "mlp.up_proj.weight"] = T(p + "ffn_up.weight")
    missing, unexpected = model.load_state_dict(sd, strict=False)
    missing = [m for m in missing if "rotary" not in m and not (a.tied and m == "lm_head.weight")]
    assert not missing and not unexpected, (missing, unexpected)

out = {"tokens": np.array(TOKENS, dtype=np.int32), "seed": np.array(SEED), "transformers": np.array(transformers.__version__)}
torch.manual_seed(0)
for key, a in TINY.items():
    t = randinit.make_model(a, SEED, "f32")
    m = hf_model(a).eval().float()
    load(a, m, t)
    with torch.no_grad():
        lg = m(torch.tensor([TOKENS])).logits[0].float().numpy()
    out[key] = lg
    print(key, lg.shape, float(np.abs(lg).max()))
np.savez_compressed(os.path.join(os.path.dirname(__file__), "hf_logits.npz"), **out)
