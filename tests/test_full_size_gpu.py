"""North-star parity at BASELINE.json's full sizes: for each of the seven architectures, greedy token sequences of the
CUDA engine are identical to the PyTorch CPU oracle over 256 tokens, and logits agree far inside the 1e-2 the north
star allows.  The oracle's outputs are the committed fixtures of tests/golden/make_full_golden.py (the oracle needs up
to 40 GB and minutes per model, so it ran once in the build container); the weights are rebuilt here from the same
seed on the device (b2_model_create_random, byte-identical to the host generator: test_model_gpu.py)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "full_size.npz"))
MODELS = sorted({k.split("/")[0] for k in G.files})
STRIDE = 97


@pytest.mark.parametrize("key", MODELS)
def test_full_size_greedy_256_and_logits(gpu_engine, key):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import archs, engine, tokenizer
    a = archs.ARCHS[key]
    prompt = tokenizer.encode("In 100 words, please give me information about United States", a.V, a.family)
    assert np.array_equal(np.array(prompt, dtype=np.int32), G[key + "/prompt"])
    model = engine.Model.random(gpu_engine, a, seed=1, max_ctx=512)
    sess = model.session(512)
    sess.prefill(prompt)
    toks, lg = sess.decode(256, want_logits=True)
    want = G[key + "/tokens"]
    same = toks == want
    first_bad = int(np.argmin(same)) if not same.all() else -1
    assert same.all(), (f"{key}: greedy tokens diverge at step {first_bad} (oracle top-1/top-2 margin there "
                        f"{G[key + '/margin'][first_bad]:.3e}, logit scale {G[key + '/absmax'][first_bad]:.2f})")
    ref = G[key + "/logits_sub"]
    got = lg[[0, 1, 127, 255]][:, ::STRIDE]
    scale = G[key + "/absmax"][[0, 1, 127, 255]][:, None]
    # fp32 on both sides; differences are summation order and libm through up to 32 layers: measured ~2e-4 of the logit
    # scale at full size; bar 1e-3 (north star: 1e-2)
    assert np.abs(got - ref).max() <= 1e-3 * scale.max(), float(np.abs(got - ref).max() / scale.max())
    model.close()
