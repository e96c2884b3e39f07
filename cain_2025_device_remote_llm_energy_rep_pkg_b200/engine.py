"""Python face of the C ABI: Engine / Model / Session, mirroring the stages of one `/api/generate`
request (model load -> prompt evaluation -> decode loop; reference call site
`experiment/RunnerConfig.py:128-131`).  All compute happens in libb2llm.so on the GPU."""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from .archs import Arch, init_recipe, tensor_id, tensor_table, to_c
from ._lib import QTYPE, SamplerCfg, Stats, TensorDesc, check


def host_random_tensors(a: Arch, seed: int, qtype: str = "q4_0") -> Dict[str, Tuple[str, np.ndarray]]:
    """The deterministic random-init model as host GGML blobs (via b2_randinit_tensor; the same bytes as
    oracle/randinit.py and as b2_model_create_random generates on the device)."""
    lib = _lib.load()
    out = {}
    for name, kind, layer, rows, cols in tensor_table(a):
        sigma, off = init_recipe(a, kind)
        if cols == 0:
            buf = np.empty(rows, dtype=np.float32)
            check(lib.b2_randinit_tensor(seed, tensor_id(kind, layer), rows, 0, QTYPE["f32"], sigma, off, 0, rows,
                                         buf.ctypes.data, buf.nbytes))
            out[name] = ("f32", buf)
        else:
            assert qtype == "q4_0"
            buf = np.empty((rows, cols // 32 * 18), dtype=np.uint8)
            check(lib.b2_randinit_tensor(seed, tensor_id(kind, layer), rows, cols, QTYPE[qtype], sigma, off, 0, rows,
                                         buf.ctypes.data, buf.nbytes))
            out[name] = (qtype, buf)
    return out


TP_HANDLE_BYTES = 64


def gather_tp_handles(mine: bytes, rank: int, size: int, all_gather) -> bytes:
    """Rank-ordered concatenation of every rank's exchange handle.  `all_gather(obj) -> list` is the host transport
    (torch.distributed.all_gather_object wrapped by the caller, a pipe, ...); entries are checked before they reach the
    C ABI, because a short or misplaced handle would map the wrong memory."""
    if len(mine) != TP_HANDLE_BYTES:
        raise ValueError(f"handle of rank {rank} has {len(mine)} bytes, expected {TP_HANDLE_BYTES}")
    got = all_gather((rank, mine))
    if len(got) != size:
        raise ValueError(f"gathered {len(got)} handles for a group of {size}")
    by_rank = dict(got)
    if sorted(by_rank) != list(range(size)):
        raise ValueError(f"handles of ranks {sorted(by_rank)} for a group of {size}")
    if any(len(h) != TP_HANDLE_BYTES for h in by_rank.values()):
        raise ValueError("a peer sent a handle of the wrong size")
    return b"".join(by_rank[r] for r in range(size))


class Engine:
    """tp_size > 1: rank `tp_rank` of a tensor-parallel group (one process and one GPU per rank).  Call `tp_connect`
    (or `tp_connect_torch`) on every rank before creating a model; afterwards every rank makes the same calls."""

    def __init__(self, device: int = 0, tp_rank: int = 0, tp_size: int = 1):
        self.lib = _lib.load()
        self.h = C.c_void_p()
        check(self.lib.b2_engine_create(device, tp_rank, tp_size, C.byref(self.h)))
        self.device, self.tp_rank, self.tp_size = device, tp_rank, tp_size

    def tp_handle(self) -> bytes:
        buf = C.create_string_buffer(TP_HANDLE_BYTES)
        check(self.lib.b2_engine_tp_handle(self.h, buf))
        return buf.raw

    def tp_connect(self, all_gather) -> None:
        handles = gather_tp_handles(self.tp_handle(), self.tp_rank, self.tp_size, all_gather)
        check(self.lib.b2_engine_tp_connect(self.h, handles))

    def tp_connect_torch(self, group=None) -> None:
        """Exchange the handles over an initialised torch.distributed process group (nccl or gloo)."""
        import torch.distributed as dist

        def all_gather(obj):
            out = [None] * dist.get_world_size(group)
            dist.all_gather_object(out, obj, group=group)
            return out

        self.tp_connect(all_gather)
        dist.barrier(group)

    def close(self):
        if self.h:
            self.lib.b2_engine_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Model:
    def __init__(self, engine: Engine, arch: Arch, handle: C.c_void_p):
        self.engine, self.arch, self.h, self.lib = engine, arch, handle, engine.lib

    @classmethod
    def from_tensors(cls, engine: Engine, arch: Arch, tensors: Dict[str, Tuple[str, np.ndarray]],
                     max_ctx: Optional[int] = None) -> "Model":
        descs = (TensorDesc * len(tensors))()
        keep = []
        for i, (name, (qt, arr)) in enumerate(tensors.items()):
            arr = np.ascontiguousarray(arr)
            keep.append(arr)
            if qt == "f32" and arr.ndim == 1:
                rows, cols = arr.shape[0], 0
            elif qt == "f32":
                rows, cols = arr.shape
            else:
                blk, nbytes = {"q4_0": (32, 18), "q8_0": (32, 34), "q4_k": (256, 144), "q6_k": (256, 210)}[qt]
                rows, cols = arr.shape[0], arr.shape[1] // nbytes * blk
            descs[i] = TensorDesc(name.encode(), QTYPE[qt], rows, cols, arr.ctypes.data, arr.nbytes)
        cfg = to_c(arch, max_ctx)
        h = C.c_void_p()
        check(engine.lib.b2_model_create(engine.h, C.byref(cfg), descs, len(tensors), C.byref(h)))
        return cls(engine, arch, h)

    @classmethod
    def random(cls, engine: Engine, arch: Arch, seed: int, qtype: str = "q4_0", head_qtype: Optional[str] = None,
               max_ctx: Optional[int] = None) -> "Model":
        cfg = to_c(arch, max_ctx)
        h = C.c_void_p()
        check(engine.lib.b2_model_create_random(engine.h, C.byref(cfg), seed, QTYPE[qtype], QTYPE[head_qtype or qtype], C.byref(h)))
        return cls(engine, arch, h)

    def weight_bytes(self) -> int:
        v = C.c_uint64()
        check(self.lib.b2_model_weight_bytes(self.h, C.byref(v)))
        return v.value

    def session(self, max_ctx: Optional[int] = None) -> "Session":
        return Session(self, max_ctx or self.arch.max_ctx)

    def close(self):
        if self.h:
            self.lib.b2_model_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Session:
    def __init__(self, model: Model, max_ctx: int):
        self.model, self.lib, self.max_ctx = model, model.lib, max_ctx
        self.h = C.c_void_p()
        check(self.lib.b2_session_create(model.h, max_ctx, C.byref(self.h)))

    def reset(self):
        check(self.lib.b2_session_reset(self.h))

    @property
    def pos(self) -> int:
        v = C.c_int32()
        check(self.lib.b2_session_pos(self.h, C.byref(v)))
        return v.value

    def stats(self) -> dict:
        st = Stats()
        check(self.lib.b2_session_stats(self.h, C.byref(st)))
        return {k: getattr(st, k) for k, _ in Stats._fields_}

    def prefill(self, tokens: Sequence[int], want_logits: bool = False) -> Optional[np.ndarray]:
        t = np.ascontiguousarray(tokens, dtype=np.int32)
        lg = np.empty(self.model.arch.V, dtype=np.float32) if want_logits else None
        check(self.lib.b2_prefill(self.h, t.ctypes.data, len(t), lg.ctypes.data if want_logits else None))
        return lg

    def decode(self, n_steps: int, want_logits: bool = False, sampler: Optional[SamplerCfg] = None):
        out = np.empty(n_steps, dtype=np.int32)
        lg = np.empty((n_steps, self.model.arch.V), dtype=np.float32) if want_logits else None
        check(self.lib.b2_decode(self.h, n_steps, C.byref(sampler) if sampler else None, out.ctypes.data,
                                 lg.ctypes.data if want_logits else None))
        return (out, lg) if want_logits else out

    def forward_tokens(self, tokens: Sequence[int]) -> np.ndarray:
        t = np.ascontiguousarray(tokens, dtype=np.int32)
        lg = np.empty((len(t), self.model.arch.V), dtype=np.float32)
        check(self.lib.b2_forward_tokens(self.h, t.ctypes.data, len(t), lg.ctypes.data))
        return lg

    def generate_greedy(self, prompt: Sequence[int], n_new: int) -> np.ndarray:
        self.prefill(prompt)
        return self.decode(n_new)

    def close(self):
        if self.h:
            self.lib.b2_session_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def nvml_energy_mj(device: int = 0) -> int:
    v = C.c_uint64()
    check(_lib.load().b2_nvml_energy_mj(device, C.byref(v)))
    return v.value
