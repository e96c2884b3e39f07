"""The K-quant configuration at llama3.1:8b's real matrix shapes: Q4_K layer matrices and a Q6_K LM head /
embedding from host super-blocks (oracle/quants.py) through b2_model_create, against the committed oracle fixture
(tests/golden/make_kquant_golden.py): 64 greedy tokens identical, logits within 1e-3 of the logit scale.  Two layers, all other
sizes real (tests/kquant_full.py); the host quantisation takes about a minute of the box's cores."""
import os

import numpy as np
import pytest

import kquant_full as K

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "kquant_full_width.npz")


@pytest.mark.timeout(1200)
def test_q4k_layers_q6k_head_at_full_width(gpu_engine):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import engine
    G = np.load(GOLD)
    a, _ = K.arch_pair()
    prompt = K.prompt_tokens()
    assert np.array_equal(np.array(prompt, dtype=np.int32), G["prompt"])
    tensors = K.build_tensors()
    assert tensors["blk.0.ffn_down.weight"][0] == "q4_k" and tensors["output.weight"][0] == "q6_k" and tensors["token_embd.weight"][0] == "q6_k"
    model = engine.Model.from_tensors(gpu_engine, a, tensors, max_ctx=256)
    sess = model.session(256)
    sess.prefill(prompt)            # the prompt runs through the tcgen05 GEMM (Q4_K layer matrices), the head through the Q6_K GEMV
    toks, lg = sess.decode(K.N_NEW, want_logits=True)
    same = toks == G["tokens"]
    first_bad = int(np.argmin(same)) if not same.all() else -1
    assert same.all(), f"greedy tokens diverge at step {first_bad} (oracle margin there {G['margin'][first_bad]:.3e}, logit scale {G['absmax'][first_bad]:.2f})"
    got = lg[[0, 1, 31, 63]][:, ::K.STRIDE]
    scale = G["absmax"][[0, 1, 31, 63]].max()
    assert np.abs(got - G["logits_sub"]).max() <= 1e-3 * scale, float(np.abs(got - G["logits_sub"]).max() / scale)
    model.close()
