"""Generates tests/golden/quant_golden.npz from gguf-py (gguf.quants), the only independent
implementation of the GGML block formats available offline.  Run once in the build
container:  python tests/golden/make_quant_golden.py
The .npz holds inputs and gguf-py's outputs; tests compare oracle/quants.py against them bit for bit."""
import numpy as np
from gguf import quants as gq, GGMLQuantizationType as T

rng = np.random.default_rng(20251106)
out = {}
# float inputs: gaussian, heavy-tailed, tiny, zero block, exact ties (x/d + 8.5 on .0 boundaries)
x = rng.standard_normal((24, 256)).astype(np.float32) * 0.02
x[1] *= 100.0
x[2] = rng.standard_cauchy(256).astype(np.float32)
x[3] = 0.0
x[4, :32] = np.linspace(-8, 7, 32, dtype=np.float32)  # d = 1 exactly: every t lands on n+.5
x[5] *= 1e-7
x[6, 7] = -3.0
x[7, 9] = 3.0
out["x"] = x
out["q4_0"] = gq.quantize(x, T.Q4_0)
out["q8_0"] = gq.quantize(x, T.Q8_0)
out["dq4_0"] = gq.dequantize(out["q4_0"], T.Q4_0)
out["dq8_0"] = gq.dequantize(out["q8_0"], T.Q8_0)
# random *bytes* for the K-quant dequantisers (all bit patterns of scales/nibbles; finite f16 d)
def rand_blob(nblk, nbytes, f16_offsets):
    b = rng.integers(0, 256, size=(nblk, nbytes), dtype=np.uint8)
    for off in f16_offsets:
        d = (rng.standard_normal(nblk) * 0.01).astype(np.float16)
        b[:, off:off + 2] = d.view(np.uint8).reshape(nblk, 2)
    return b
b4k = rand_blob(48, 144, (0, 2)).reshape(6, 8 * 144)
b6k = rand_blob(48, 210, (208,)).reshape(6, 8 * 210)
b40 = rand_blob(64, 18, (0,)).reshape(4, 16 * 18)
out["b4k"], out["dq4_k"] = b4k, gq.dequantize(b4k, T.Q4_K)
out["b6k"], out["dq6_k"] = b6k, gq.dequantize(b6k, T.Q6_K)
out["b40"], out["dqb40"] = b40, gq.dequantize(b40, T.Q4_0)
np.savez_compressed(__file__.replace("make_quant_golden.py", "quant_golden.npz"), **out)
print({k: (v.shape, v.dtype) for k, v in out.items()})
