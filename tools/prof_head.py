"""The LM-head GEMV alone (llama3.1:8b: 128256 x 4096, 295 MB): in-kernel %globaltimer stamps per CTA -> achieved stream rate."""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from cain_2025_device_remote_llm_energy_rep_pkg_b200 import _lib
from cain_2025_device_remote_llm_energy_rep_pkg_b200.devmem import DevBuf
lib = _lib.load()
chk = _lib.check
V, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (128256, 4096)
rng = np.random.default_rng(0)
d = DevBuf(lib.b2_packed_bytes(2, V, H))
raw = rng.integers(0, 256, size=(V, H // 32 * 18), dtype=np.uint8)
raw[:, 1::18] = 0x20
chk(lib.b2_op_pack_weights(2, V, H, raw.ctypes.data, None, d.ptr))
x = DevBuf.from_numpy(rng.standard_normal(H).astype(np.float32))
nw = DevBuf.from_numpy(np.ones(H, dtype=np.float32))
y, tok = DevBuf(V * 4), DevBuf(4)
dbg = DevBuf(148 * 8 * 8)
lib.b2dbg_gemv_profile.argtypes = [C.c_void_p]
lib.b2dbg_gemv_profile(dbg.ptr)
nbytes = V * (H // 32) * 18
for rep in range(4):
    chk(lib.b2_op_norm_gemv(2, d.ptr, V, H, x.ptr, nw.ptr, np.float32(1e-5), 0, None, 3, 0, y.ptr, None, tok.ptr))
    chk(lib.b2_dev_sync())
    t = dbg.to_numpy(np.int64, (148, 8)).astype(np.float64)
    t0 = t[:, 0].min()
    whole = t[:, 5].max() - t0
    stream = t[:, 4] - t[:, 3]
    print(f"rep {rep}: kernel {whole/1e3:7.1f} us = {nbytes/whole:6.0f} GB/s | stream per CTA median {np.median(stream)/1e3:6.1f} max {stream.max()/1e3:6.1f} us "
          f"({nbytes/148/np.median(stream):5.1f} GB/s per SM) | first start->staged median {np.median(t[:,3]-t0)/1e3:5.1f} us | stream done spread {(t[:,4].max()-t[:,4].min())/1e3:5.1f} us")
