"""Several GPUs driven by ONE process (server.py --devices: one engine per GPU): every device must give the oracle's
tokens.  Kernel function attributes (the large dynamic shared memory of the GEMV and of the prompt GEMM) belong to a device, not to
the process; the first multi-GPU server run of round 2 failed on every GPU but the first for exactly that reason.  Needs two GPUs."""
import os
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "full_size.npz")


def _n_gpus() -> int:
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_n_gpus() < 2, reason="needs two GPUs")
def test_two_engines_in_one_process_both_match_the_oracle(lib):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import archs, engine, tokenizer
    G = np.load(GOLD)
    key = "qwen2:1.5b"
    a = archs.ARCHS[key]
    prompt = tokenizer.encode("In 100 words, please give me information about United States", a.V, a.family)
    out = {}

    def work(dev):
        eng = engine.Engine(dev)
        m = engine.Model.random(eng, a, seed=1, max_ctx=512)
        s = m.session(512)
        s.prefill(prompt)                      # batched prompt evaluation (tcgen05 GEMM) ...
        out[dev] = s.decode(64)                # ... and the decode GEMVs, both with > 48 KB of dynamic shared memory
        s.close()
        m.close()

    n = min(_n_gpus(), 4)
    threads = [threading.Thread(target=work, args=(d,)) for d in range(n)]   # concurrently, as the server's request threads do
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for d in range(n):
        assert d in out and np.array_equal(out[d], G[key + "/tokens"][:64]), f"device {d}"
