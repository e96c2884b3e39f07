"""The C-ABI library loads on a CPU-only machine, exports every symbol include/b2llm.h declares, and
its compute entry points fail loudly (no CPU fallback) when there is no CUDA device."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import HAS_GPU, ROOT


def header_functions():
    src = open(os.path.join(ROOT, "include", "b2llm.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b2_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import _lib
    names = header_functions()
    assert len(names) >= 35
    for n in names:
        assert hasattr(lib, n), f"{n} declared in b2llm.h but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes prototype"
    assert set(_lib.SIGNATURES) == set(names)


def test_struct_layout_matches_header(lib):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import CArchCfg
    from cain_2025_device_remote_llm_energy_rep_pkg_b200._lib import SamplerCfg, Stats, TensorDesc
    assert ctypes.sizeof(CArchCfg) == 14 * 4 + 6 * 4
    assert ctypes.sizeof(TensorDesc) == 48
    assert ctypes.sizeof(SamplerCfg) == 32
    assert ctypes.sizeof(Stats) == 40


def test_size_helpers(lib):
    assert lib.b2_packed_bytes(2, 10, 64) == 1 * 4 * 288          # 1 tile of 16 rows, 2 blocks padded to 4, 288 B per block column
    assert lib.b2_packed_bytes(2, 4096, 4096) == 4096 * 4096 // 32 * 18   # no padding at real sizes: 4.5 bit/weight
    assert lib.b2_packed_bytes(2, 10, 33) == 0
    assert lib.b2_act_bytes(64) == 4 * 100


@pytest.mark.skipif(HAS_GPU, reason="checks the no-device behaviour")
def test_compute_fails_loudly_without_gpu(lib):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200._lib import B2Error
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import Engine
    with pytest.raises(B2Error) as ei:
        Engine(0)
    assert ei.value.code == -2 and "cuda" in str(ei.value).lower()


def test_missing_library_is_an_error(monkeypatch, lib):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libb2llm.so")
    with pytest.raises(FileNotFoundError):
        _lib.load()


def test_host_randinit_matches_numpy_oracle(lib):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import archs
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import host_random_tensors
    from oracle import archs as oarchs
    from oracle import randinit
    for k in ("qwen2:1.5b", "gemma:2b", "phi3:3.8b"):
        t = host_random_tensors(archs.TINY[k], 11)
        o = randinit.make_model(oarchs.TINY[k], 11, "q4_0")
        assert t.keys() == o.keys()
        for n in t:
            assert t[n][0] == o[n][0] and np.array_equal(t[n][1], o[n][1]), (k, n)
    g = randinit.gauss(1, 2, 200000)
    assert abs(g.mean()) < 0.01 and abs(g.std() - 1.0) < 0.01


def test_product_and_oracle_arch_tables_agree():
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import archs
    from oracle import archs as oarchs
    import dataclasses
    for table, otable in ((archs.ARCHS, oarchs.ARCHS), (archs.TINY, oarchs.TINY)):
        assert table.keys() == otable.keys()
        for k in table:
            assert dataclasses.asdict(table[k]) == dataclasses.asdict(otable[k]), k
            assert archs.tensor_table(table[k]) == oarchs.tensor_table(otable[k])
    assert set(archs.ARCHS) == {"llama3.1:8b", "gemma:2b", "gemma:7b", "phi3:3.8b", "qwen2:1.5b", "qwen2:7b", "mistral:7b"}


def test_prompt_gemm_k_split_against_wave_quantisation(lib):
    """Host logic of csrc/prefill.cu (launch_prefill_gemm_qt): work items are dealt round robin to one CTA per SM, so a launch takes
    ceil(groups * pieces / SMs) / pieces item times; the chooser takes the smallest number of K pieces within 5 % of the best."""
    import math
    f = lib.b2dbg_pf_ksplit
    f.argtypes, f.restype = [ctypes.c_int] * 3, ctypes.c_int
    assert f(224, 148, 4) == 3      # llama3.1:8b gate/up: two rounds for 1.5 rounds of work -> five rounds of a third
    assert f(296, 148, 4) == 1      # qwen2:7b gate/up: exactly two rounds
    assert f(32, 148, 4) == 4       # O / down projection: 128 items on 148 SMs
    assert f(32, 148, 8) == 4       # more pieces would not shorten the launch
    assert f(48, 148, 4) == 3       # QKV: 144 items
    assert f(384, 148, 3) == 3      # gemma:7b gate/up
    assert f(5, 148, 1) == 1 and f(0, 148, 8) == 1 and f(10, 0, 8) == 1
    for groups in (1, 7, 16, 96, 148, 149, 224, 500):
        for kmax in (1, 2, 4, 8):
            ks = f(groups, 148, kmax)
            assert 1 <= ks <= kmax
            t = lambda k: math.ceil(groups * k / 148) / k
            assert t(ks) <= t(1) + 1e-12
            assert t(ks) <= min(t(k) for k in range(1, kmax + 1)) / 0.95 / 0.95 + 1e-12   # within the tolerance of the greedy scan
