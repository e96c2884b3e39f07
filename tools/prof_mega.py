"""Timeline of the stages inside the persistent token kernel (mega.cu): %globaltimer stamps of thread 0 of every CTA.
Needs the profiling build:  EXTRA=-DMEGA_PROF csrc/build.sh   (rebuild without it afterwards).
Stamps per stage: 0 stage entered, 1 first weight loads issued (before the grid barrier), 2 barrier passed + input copy issued,
3 input staged, 4 stream done, 5 flushed.  Attention: 0 entered, 1 pages prefetched, 2 barrier passed, 5 done."""
import ctypes as C, os, sys
import numpy as np
os.environ["B2_MEGA"] = "1"
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from cain_2025_device_remote_llm_energy_rep_pkg_b200 import _lib
from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
from cain_2025_device_remote_llm_energy_rep_pkg_b200.devmem import DevBuf
from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import Engine, Model
lib = _lib.load()
eng = Engine(0)
a = ARCHS[sys.argv[1] if len(sys.argv) > 1 else "llama3.1:8b"]
ctx0 = int(sys.argv[2]) if len(sys.argv) > 2 else 32
m = Model.random(eng, a, seed=1)
s = m.session(max(256, ctx0 + 64))
s.prefill([1 + (i % 1000) for i in range(ctx0)])
s.decode(8)
per_tok = 1 + 5 * a.L + 1          # token begin, 5 stages per layer, LM head
ntok = 3
cap = ntok * per_tok
G = 148
dbg = DevBuf(cap * G * 8 * 8)
lib.b2dbg_mega_profile.argtypes = [C.c_void_p, C.c_int]
rc = lib.b2dbg_mega_profile(dbg.ptr, cap)
assert rc == 0, "not a MEGA_PROF build"
s.decode(ntok)
lib.b2dbg_mega_profile(None, 0)
d = dbg.to_numpy(np.int64, (cap, G, 8)).astype(np.float64)
names = ["qkv", "attn", "o", "gate/up", "down"]
t = 1                                # second profiled token
lay = min(10, a.L - 1)
first = t * per_tok + 1 + 5 * lay
base = d[first, :, 0].min()
print("stage: [ns since layer start] first CTA in | median: loads issued, barrier passed, input staged, stream done, flushed | last CTA out")
for k in range(11):
    x = d[first + k] - base
    nm = names[k % 5]
    if nm == "attn":
        print(f"{nm:8s} in {x[:,0].min():7.0f} | {np.median(x[:,1]):7.0f} {np.median(x[:,2]):7.0f} {'':7s} {'':7s} {np.median(x[:,5]):7.0f} | out {x[:,5].max():7.0f}")
    else:
        print(f"{nm:8s} in {x[:,0].min():7.0f} | " + " ".join(f"{np.median(x[:, i]):7.0f}" for i in range(1, 6)) + f" | out {x[:,5].max():7.0f}")
# averages over all layers of the token
def stage_stats(k):
    idx = [t * per_tok + 1 + 5 * l + k for l in range(a.L)]
    x = d[idx]
    return x
print("\nper-stage averages over the layers of one token (us): wait at barrier (1->2, median CTA), staging (2->3), stream (3->4), flush (4->5), whole stage (first in -> last out)")
tot = 0.0
for k, nm in enumerate(names):
    x = stage_stats(k)
    whole = (x[:, :, 5].max(axis=1) - x[:, :, 0].min(axis=1)).mean() / 1e3
    if nm == "attn":
        print(f"{nm:8s} barrier {np.median(x[:,:,2]-x[:,:,1],axis=1).mean()/1e3:6.2f} (after last arrival {(np.median(x[:,:,2],axis=1)-x[:,:,1].max(axis=1)).mean()/1e3:5.2f})  work {np.median(x[:,:,5]-x[:,:,2],axis=1).mean()/1e3:6.2f} (max CTA {(x[:,:,5]-x[:,:,2]).max(axis=1).mean()/1e3:6.2f})  whole {whole:6.2f}")
    else:
        print(f"{nm:8s} barrier {np.median(x[:,:,2]-x[:,:,1],axis=1).mean()/1e3:6.2f} (after last arrival {(np.median(x[:,:,2],axis=1)-x[:,:,1].max(axis=1)).mean()/1e3:5.2f})  copy {np.median(x[:,:,6]-x[:,:,2],axis=1).mean()/1e3:5.2f}  staging {np.median(x[:,:,3]-x[:,:,2],axis=1).mean()/1e3:6.2f}  "
              f"stream {np.median(x[:,:,4]-x[:,:,3],axis=1).mean()/1e3:6.2f} (max {(x[:,:,4]-x[:,:,3]).max(axis=1).mean()/1e3:6.2f})  flush {np.median(x[:,:,5]-x[:,:,4],axis=1).mean()/1e3:6.2f}  whole {whole:6.2f}")
    tot += whole
print(f"sum of stage spans per layer: {tot:.2f} us")
tok = d[t * per_tok:(t + 1) * per_tok]
print("token time (token begin -> LM head flushed):", (tok[-1, :, 5].max() - tok[0, :, 0].min()) / 1e3, "us;  LM head:",
      (tok[-1, :, 5].max() - tok[-1, :, 0].min()) / 1e3, "us;  token begin:", (tok[0, :, 5].max() - tok[0, :, 0].min()) / 1e3, "us")
