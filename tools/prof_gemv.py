"""Standalone driver for profiling the streaming GEMV (ncu --set full): llama3.1:8b shapes, op-level C ABI."""
import ctypes as C
import os
import sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from cain_2025_device_remote_llm_energy_rep_pkg_b200 import _lib
from cain_2025_device_remote_llm_energy_rep_pkg_b200.devmem import DevBuf

lib = _lib.load()
chk = _lib.check
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
H, I = 4096, 14336
rng = np.random.default_rng(0)


def rand_packed(rows, cols):
    d = DevBuf(lib.b2_packed_bytes(2, rows, cols))
    raw = rng.integers(0, 256, size=(rows, cols // 32 * 18), dtype=np.uint8)
    raw[:, 1::18] = 0x20    # keep the f16 scales small and finite
    chk(lib.b2_op_pack_weights(2, rows, cols, raw.ctypes.data, None, d.ptr))
    return d


w_down, w_gu = rand_packed(H, I), rand_packed(2 * I, H)
x = DevBuf.from_numpy(rng.standard_normal(H).astype(np.float32))
nw = DevBuf.from_numpy(np.ones(H, dtype=np.float32))
h = DevBuf.from_numpy(rng.standard_normal(I).astype(np.float32))
act_h = DevBuf(lib.b2_act_bytes(I))
chk(lib.b2_op_quant_act(h.ptr, I, act_h.ptr))
y, hh, act_o = DevBuf(H * 4), DevBuf(I * 4), DevBuf(lib.b2_act_bytes(I))
dbg = DevBuf(148 * 8 * 8)
prof = os.environ.get("B2_PROF") == "1"
if prof:
    lib.b2dbg_gemv_profile.argtypes = [C.c_void_p]
    lib.b2dbg_gemv_profile(dbg.ptr)


def report(name):
    if not prof:
        return
    chk(lib.b2_dev_sync())
    d = dbg.to_numpy(np.int64, (148, 8)).astype(np.float64)
    t0 = d[:, 0].min()
    rel = d[:, :6] - t0
    names = ["start", "loads issued", "wait done", "input staged", "stream done", "flushed"]
    print(name + ": ns since first CTA start (median / max over CTAs): " +
          ", ".join(f"{n} {np.median(rel[:, i]):.0f}/{rel[:, i].max():.0f}" for i, n in enumerate(names)))


for _ in range(reps):
    chk(lib.b2_op_gemv(2, w_down.ptr, H, I, act_h.ptr, x.ptr, 1, 0, y.ptr))
    report("down 4096x14336 (IN_ACT, RES)")
    chk(lib.b2_op_norm_gemv(2, w_gu.ptr, 2 * I, H, x.ptr, nw.ptr, np.float32(1e-5), 0, None, 2, 0, hh.ptr, act_o.ptr, None))
    report("gate/up 28672x4096 (IN_NORM, ACTMUL)")
chk(lib.b2_dev_sync())
print("ok", float(np.abs(y.to_numpy(np.float32, (H,))).max()))
