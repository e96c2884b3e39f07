"""ctypes binding of libb2llm.so (C ABI: include/b2llm.h).  Fails loudly when the library is missing:
there is no CPU fallback for any compute call."""
from __future__ import annotations

import ctypes as C
import os

from .archs import CArchCfg

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B2_LIB") or os.path.join(_HERE, "libb2llm.so")   # B2_LIB: an experimental build (csrc/build.sh with OUT=, EXTRA=)

T_F32, T_Q4_0, T_Q8_0, T_Q4_K, T_Q6_K = 0, 2, 8, 12, 14
QTYPE = {"f32": T_F32, "q4_0": T_Q4_0, "q8_0": T_Q8_0, "q4_k": T_Q4_K, "q6_k": T_Q6_K}


class B2Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"b2llm error {code}: {msg}")
        self.code = code


class TensorDesc(C.Structure):
    _fields_ = [("name", C.c_char_p), ("type", C.c_int32), ("rows", C.c_int64), ("cols", C.c_int64),
                ("data", C.c_void_p), ("nbytes", C.c_uint64)]


class SamplerCfg(C.Structure):
    _fields_ = [("top_k", C.c_int32), ("temperature", C.c_float), ("top_p", C.c_float), ("seed", C.c_uint64),
                ("repeat_penalty", C.c_float), ("repeat_last_n", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("prompt_eval_count", C.c_uint64), ("eval_count", C.c_uint64), ("prompt_eval_ns", C.c_uint64),
                ("eval_ns", C.c_uint64), ("kernel_launches", C.c_uint64)]


P = C.POINTER
vp, i32, i64, u64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_float

# name -> (restype, argtypes); every symbol include/b2llm.h declares
SIGNATURES = {
    "b2_last_error": (C.c_char_p, []),
    "b2_version": (C.c_char_p, []),
    "b2_engine_create": (i32, [i32, i32, i32, P(vp)]),
    "b2_engine_destroy": (i32, [vp]),
    "b2_engine_device": (i32, [vp, P(i32)]),
    "b2_engine_tp_handle": (i32, [vp, vp]),
    "b2_engine_tp_connect": (i32, [vp, vp]),
    "b2_model_create": (i32, [vp, P(CArchCfg), P(TensorDesc), i32, P(vp)]),
    "b2_model_create_random": (i32, [vp, P(CArchCfg), u64, i32, i32, P(vp)]),
    "b2_model_destroy": (i32, [vp]),
    "b2_model_weight_bytes": (i32, [vp, P(u64)]),
    "b2_session_create": (i32, [vp, i32, P(vp)]),
    "b2_session_reset": (i32, [vp]),
    "b2_session_destroy": (i32, [vp]),
    "b2_session_pos": (i32, [vp, P(i32)]),
    "b2_session_stats": (i32, [vp, P(Stats)]),
    "b2_prefill": (i32, [vp, vp, i32, vp]),
    "b2_decode": (i32, [vp, i32, P(SamplerCfg), vp, vp]),
    "b2_forward_tokens": (i32, [vp, vp, i32, vp]),
    "b2_nvml_energy_mj": (i32, [i32, P(u64)]),
    "b2_nvml_power_mw": (i32, [i32, P(C.c_uint32)]),
    "b2_nvml_util": (i32, [i32, P(C.c_uint32), P(C.c_uint32)]),
    "b2_nvml_clocks": (i32, [i32, P(C.c_uint32), P(C.c_uint32), P(u64)]),
    "b2_randinit_tensor": (i32, [u64, C.c_uint32, i64, i64, i32, f32, f32, i64, i64, vp, u64]),
    "b2_dev_alloc": (i32, [u64, P(vp)]),
    "b2_dev_free": (i32, [vp]),
    "b2_dev_upload": (i32, [vp, vp, u64]),
    "b2_dev_download": (i32, [vp, vp, u64]),
    "b2_dev_sync": (i32, []),
    "b2_packed_bytes": (u64, [i32, i64, i64]),
    "b2_op_pack_weights": (i32, [i32, i64, i64, vp, vp, vp]),
    "b2_act_bytes": (u64, [i64]),
    "b2_op_quant_act": (i32, [vp, i64, vp]),
    "b2_op_rmsnorm_quant": (i32, [vp, vp, i64, f32, i32, vp, vp]),
    "b2_op_gemv": (i32, [i32, vp, i64, i64, vp, vp, i32, i32, vp]),
    "b2_op_gemv_f32": (i32, [i32, vp, i64, i64, vp, vp, i32, vp]),
    "b2_op_norm_gemv": (i32, [i32, vp, i64, i64, vp, vp, f32, i32, vp, i32, i32, vp, vp, vp]),
    "b2_op_prefill_gemm": (i32, [i32, vp, i64, i64, vp, i32, vp, f32, i32, vp, i32, i32, vp]),
    "b2_op_rope_kv": (i32, [vp, i32, i32, i32, i32, vp, vp, vp, vp, vp, i32]),
    "b2_op_attn_decode": (i32, [vp, i32, i32, i32, i32, vp, vp, vp, i32, vp]),
    "b2_op_attn_fused": (i32, [vp, i32, i32, i32, i32, vp, vp, vp, vp, vp, i32, i32, vp, vp]),
    "b2_op_argmax": (i32, [vp, i64, vp]),
    "b2_op_argmax_noalloc": (i32, [vp, i64, vp]),
    "b2_op_sample_topk": (i32, [vp, i64, i32, f32, f32, f32, vp]),
    "b2_op_sample_topk_penalty": (i32, [vp, i64, i32, f32, f32, f32, vp, i32, f32, i32, vp]),
}

_lib = None


def load() -> C.CDLL:
    """Loads libb2llm.so and sets every prototype.  Raises if the library has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                    "(or csrc/build.sh); there is no CPU fallback")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise B2Error(rc, load().b2_last_error().decode("utf-8", "replace"))
