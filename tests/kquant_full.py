"""Helper (not a test): the K-quant configuration of SURVEY.md §8d config 4 at the real matrix shapes — llama3.1:8b's hidden /
ffn / head sizes and vocabulary with Q4_K layer matrices and a Q6_K LM head and embedding (what Ollama ships for that tag is a
K-quant mix), quantised on the host by oracle/quants.py from the deterministic initialiser.  Two layers instead of 32: the
kernels a layer runs do not depend on how many layers follow, and the numpy Q4_K search quantiser manages ~1 M weights/s per core
(all 8 G weights would take hours).  Used by tests/golden/make_kquant_golden.py (oracle side) and tests/test_kquant_full_gpu.py."""
import dataclasses
import os
from concurrent.futures import ProcessPoolExecutor

import numpy as np

SEED, N_LAYERS, N_NEW, STRIDE = 3, 2, 64, 97
PROMPT_TEXT = "In 100 words, please give me information about United States"   # RunnerConfig.py:120 + topics.csv:2


def arch_pair():
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
    from oracle.archs import ARCHS as OARCHS
    return dataclasses.replace(ARCHS["llama3.1:8b"], L=N_LAYERS), dataclasses.replace(OARCHS["llama3.1:8b"], L=N_LAYERS)


def _chunk(job):
    from oracle import randinit
    from oracle.archs import ARCHS as OARCHS
    kind, layer, rows, cols, qt, r0, n = job
    oa = dataclasses.replace(OARCHS["llama3.1:8b"], L=N_LAYERS)
    return randinit.make_tensor(oa, SEED, kind, layer, rows, cols, qt, r0, n)


def build_tensors(workers=None):
    """name -> (qtype, ndarray) as oracle.randinit.make_model(qtype="q4_k", head_qtype="q6_k") would return it, built by a pool of
    processes over row chunks (rows quantise independently, so the bytes are the same)."""
    from oracle import randinit
    from oracle.archs import tensor_table
    _, oa = arch_pair()
    jobs, spans, out = [], [], {}
    for name, kind, layer, rows, cols in tensor_table(oa):
        if cols == 0:
            out[name] = ("f32", randinit.make_tensor(oa, SEED, kind, layer, rows, 0, "f32"))
            continue
        qt = randinit.qtype_for(kind, "q4_k", "q6_k")
        step = max(16, (1 << 22) // cols)
        first = len(jobs)
        jobs += [(kind, layer, rows, cols, qt, r0, min(step, rows - r0)) for r0 in range(0, rows, step)]
        spans.append((name, qt, first, len(jobs)))
    with ProcessPoolExecutor(max_workers=workers or os.cpu_count()) as pool:
        parts = list(pool.map(_chunk, jobs, chunksize=1))
    for name, qt, a, b in spans:
        out[name] = (qt, np.concatenate(parts[a:b], axis=0))
    return out


def prompt_tokens():
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import tokenizer
    a, _ = arch_pair()
    return tokenizer.encode(PROMPT_TEXT, a.V, a.family)
