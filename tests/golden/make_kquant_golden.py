"""Generates tests/golden/kquant_full_width.npz: 64 greedy tokens and sampled logits of the PyTorch CPU oracle for llama3.1:8b's
real matrix shapes with Q4_K layer matrices and a Q6_K head (two layers; tests/kquant_full.py says why), the K-quant case of
SURVEY.md §8d config 4.  Weights: deterministic initialiser -> oracle/quants.py quantisers (the Q4_K / Q6_K dequantisers are pinned
to gguf-py; the quantisers restate llama.cpp's search and are unpinned — the GPU side receives the very same super-blocks, so
the comparison does not depend on them).  Takes a few minutes:
    python tests/golden/make_kquant_golden.py
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import kquant_full as K  # noqa: E402
from oracle.model import OracleModel  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "kquant_full_width.npz")


def main():
    torch.set_num_threads(os.cpu_count())
    t0 = time.time()
    tensors = K.build_tensors()
    print(f"tensors: {time.time() - t0:.0f}s", flush=True)
    _, oa = K.arch_pair()
    om = OracleModel(oa, tensors)
    prompt = K.prompt_tokens()
    toks, logits = om.session(len(prompt) + K.N_NEW + 1).generate_greedy(prompt, K.N_NEW, keep_logits=True)
    lg = np.stack([l.numpy() for l in logits])
    top2 = np.sort(lg, axis=1)[:, -2:]
    np.savez_compressed(OUT, prompt=np.array(prompt, dtype=np.int32), tokens=np.array(toks, dtype=np.int32),
                        logits_sub=lg[[0, 1, 31, 63]][:, ::K.STRIDE].astype(np.float32), margin=(top2[:, 1] - top2[:, 0]).astype(np.float32),
                        absmax=np.abs(lg).max(axis=1).astype(np.float32))
    print(f"done in {time.time() - t0:.0f}s, min margin {float((top2[:, 1] - top2[:, 0]).min()):.3e}, first tokens {toks[:8]}", flush=True)


if __name__ == "__main__":
    main()
