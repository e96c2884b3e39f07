"""Generates tests/golden/full_size.npz: greedy token sequences (256 steps) and sampled logits of the PyTorch CPU
oracle for the FULL-SIZE random-init model of each of the seven architectures (seed 1, Q4_0), the north star's
parity bar ("greedy token sequences identical over 256 tokens").  Weights come from the deterministic
initialiser (b2_randinit_tensor == oracle/randinit.py, checked in tests/test_abi.py); the GPU test rebuilds the
same bytes on the device with b2_model_create_random.  Takes tens of minutes and up to ~40 GB of RAM:
    python tests/golden/make_full_golden.py [model ...]
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
sys.path.insert(0, ROOT)
from cain_2025_device_remote_llm_energy_rep_pkg_b200 import tokenizer  # noqa: E402
from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS  # noqa: E402
from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import host_random_tensors  # noqa: E402
from oracle.archs import ARCHS as OARCHS  # noqa: E402
from oracle.model import OracleModel  # noqa: E402

SEED, N_NEW, STRIDE = 1, 256, 97
PROMPT = "In 100 words, please give me information about United States"   # RunnerConfig.py:120 + topics.csv:2
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "full_size.npz")


def main():
    keys = sys.argv[1:] or sorted(ARCHS, key=lambda k: ARCHS[k].matmul_weights())
    data = dict(np.load(OUT)) if os.path.exists(OUT) else {}
    torch.set_num_threads(os.cpu_count())
    for key in keys:
        a = ARCHS[key]
        t0 = time.time()
        tensors = host_random_tensors(a, SEED)
        om = OracleModel(OARCHS[key], tensors)
        prompt = tokenizer.encode(PROMPT, a.V, a.family)
        toks, logits = om.session(len(prompt) + N_NEW + 1).generate_greedy(prompt, N_NEW, keep_logits=True)
        lg = np.stack([l.numpy() for l in logits])
        top2 = np.sort(lg, axis=1)[:, -2:]
        data[key + "/prompt"] = np.array(prompt, dtype=np.int32)
        data[key + "/tokens"] = np.array(toks, dtype=np.int32)
        data[key + "/logits_sub"] = lg[[0, 1, 127, 255]][:, ::STRIDE].astype(np.float32)
        data[key + "/margin"] = (top2[:, 1] - top2[:, 0]).astype(np.float32)      # top-1 / top-2 gap at every step
        data[key + "/absmax"] = np.abs(lg).max(axis=1).astype(np.float32)
        np.savez_compressed(OUT, **data)
        print(f"{key}: {time.time() - t0:.0f}s, min margin {data[key + '/margin'].min():.3e}, first tokens {toks[:8]}", flush=True)
        del om, tensors, logits, lg


if __name__ == "__main__":
    main()
