import ctypes as C, time, sys
sys.path.insert(0, '.')
from cain_2025_device_remote_llm_energy_rep_pkg_b200 import _lib
lib = _lib.load()
v = C.c_uint64(); g = C.c_uint32(); m = C.c_uint32()
t=time.time(); lib.b2_nvml_energy_mj(0, C.byref(v)); print("first energy call (NVML start-up)", round(time.time()-t,3), "s")
for name, fn in (("energy", lambda: lib.b2_nvml_energy_mj(0, C.byref(v))), ("util", lambda: lib.b2_nvml_util(0, C.byref(g), C.byref(m)))):
    t=time.time()
    for _ in range(200): fn()
    print(name, "per call", round((time.time()-t)/200*1e3,3), "ms")
import pynvml
t=time.time(); pynvml.nvmlInit(); print("pynvml.nvmlInit", round(time.time()-t,3), "s")
