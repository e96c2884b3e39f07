"""ORACLE (test infrastructure): restatement of the sampler contract of csrc/sampler.cu.

Ollama's default request (the one `experiment/RunnerConfig.py:128` sends, no `options`) samples with
top-k 40 / top-p 0.9 / temperature 0.8 (external and unpinned).  The contract used here:
candidates = the k largest logits sorted by (logit desc, index asc); p = softmax(logit / T) over them;
keep the shortest prefix whose cumulative probability reaches top_p; draw the first kept candidate whose
cumulative (unnormalised) mass reaches u * kept_mass.  Greedy = argmax with the lowest index on ties.
Parity unpinned against Ollama (no golden vectors exist); pinned only against this file."""
import numpy as np


def argmax_lowest(logits: np.ndarray) -> int:
    return int(np.argmax(logits))  # numpy returns the first maximum


def topk_sorted(logits: np.ndarray, k: int):
    idx = np.lexsort((np.arange(len(logits)), -logits.astype(np.float64)))[:k]
    return idx, logits[idx]


def sample_topk(logits: np.ndarray, k: int, temperature: float, top_p: float, u: float):
    """Returns (token, kept_indices, margin) — margin is the distance of u*mass from the nearest CDF step
    (tests skip draws that sit on a step within float rounding)."""
    idx, v = topk_sorted(logits, k)
    w = np.exp((v.astype(np.float32) - v[0]) / np.float32(temperature)).astype(np.float32)
    p = w / w.sum(dtype=np.float32)
    cum = np.cumsum(p, dtype=np.float32)
    keep = int(np.argmax(cum >= np.float32(top_p))) + 1 if (cum >= np.float32(top_p)).any() else len(idx)
    mass = w[:keep].sum(dtype=np.float32)
    run = np.cumsum(w[:keep], dtype=np.float32)
    target = np.float32(u) * mass
    hit = run >= target
    j = int(np.argmax(hit)) if hit.any() else keep - 1
    margin = float(np.min(np.abs(run - target)) / mass)
    cut_margin = float(np.min(np.abs(cum - np.float32(top_p)))) if top_p < 1.0 else float("inf")
    return int(idx[j]), idx[:keep], min(margin, cut_margin)


def apply_repeat_penalty(logits: np.ndarray, context, penalty: float, last_n: int) -> np.ndarray:
    """Ollama's options.repeat_penalty / repeat_last_n (llama.cpp's rule, external and unpinned): every distinct token among the
    last `last_n` tokens of the context has its logit divided by the penalty when positive, multiplied by it otherwise."""
    out = logits.astype(np.float32).copy()
    if not penalty or penalty == 1.0 or last_n <= 0:
        return out
    for t in set(int(x) for x in list(context)[-last_n:]):
        out[t] = out[t] / np.float32(penalty) if out[t] > 0 else out[t] * np.float32(penalty)
    return out


def _splitmix64(z: int) -> int:
    z = (z + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
    return z ^ (z >> 31)


def uniform(seed: int, i: int) -> float:
    """The uniform of draw i of a session (csrc/engine.cu:k_make_u): top 24 bits of splitmix64(splitmix64(seed) + i)."""
    h = _splitmix64((_splitmix64(seed & 0xFFFFFFFFFFFFFFFF) + i) & 0xFFFFFFFFFFFFFFFF)
    return float(np.float32(h >> 40) * np.float32(2.0 ** -24))
