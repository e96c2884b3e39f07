"""Timeline of the GEMV kernels inside a real decode step (llama3.1:8b): %globaltimer stamps of thread 0 of every CTA,
collected through the development hook b2dbg_gemv_profile_ring.  Graphs are disabled so that every launch gets its own slot."""
import ctypes as C, os, sys
import numpy as np
os.environ["B2_NO_GRAPH"] = "1"
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from cain_2025_device_remote_llm_energy_rep_pkg_b200 import _lib
from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
from cain_2025_device_remote_llm_energy_rep_pkg_b200.devmem import DevBuf
from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import Engine, Model
lib = _lib.load()
eng = Engine(0)
a = ARCHS[sys.argv[1] if len(sys.argv) > 1 else "llama3.1:8b"]
m = Model.random(eng, a, seed=1)
s = m.session(256)
s.prefill(list(range(1, 33)))
s.decode(8)
per_tok = 4 * a.L + 1
slots = 2 * per_tok
dbg = DevBuf(slots * 148 * 8 * 8)
lib.b2dbg_gemv_profile_ring.argtypes = [C.c_void_p, C.c_int]
lib.b2dbg_gemv_profile_ring(dbg.ptr, slots)
s.decode(2)
lib.b2dbg_gemv_profile_ring(None, 1)
d = dbg.to_numpy(np.int64, (slots, 148, 8)).astype(np.float64)
names = ["qkv", "o", "gate/up", "down"]
base = d[per_tok + 4 * 10, :, 0].min()      # second token, layer 10
print("kernel: [ns since layer start]  first CTA start | median: loads issued, wait done, input staged, stream done, flushed | last CTA flushed")
for k in range(9):
    x = d[per_tok + 4 * 10 + k] - base
    print(f"{names[k % 4]:8s} start {x[:,0].min():7.0f} | " + " ".join(f"{np.median(x[:, i]):7.0f}" for i in range(1, 6)) + f" | end {x[:,5].max():7.0f}")
    if len(sys.argv) > 2:   # spread over the CTAs: min / max of every stamp
        print("          min    " + " ".join(f"{x[:, i].min():7.0f}" for i in range(0, 6)))
        print("          max    " + " ".join(f"{x[:, i].max():7.0f}" for i in range(0, 6)))
tok = d[per_tok:, :, :]
print("token time (first start -> last flush of the LM head):", (tok[-1, :, 5].max() - tok[0, :, 0].min()) / 1e3, "us")
