#!/usr/bin/env python
"""Development aid: error of the prompt GEMM against a float64 product on the activation grid, as a multiple of sum |w||x|
(the unit of the tolerances in tests/test_prefill_gpu.py: 6e-6 against the oracle, 2e-6 against the decode GEMV).
Run it with B2_LIB pointing at an experimental build to compare accumulation schemes:
    python tests/pf_accumulation_error.py ; B2_LIB=$PWD/tools/bin/lib_x.so python tests/pf_accumulation_error.py
Not collected by pytest (a measurement, no assertions); it lives under tests/ because it uses oracle/ as the checker, like the tests it mirrors."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))   # test_ops_gpu helpers

from oracle import quants as Q   # noqa: E402
from test_ops_gpu import pack, ref_gemv   # noqa: E402
from cain_2025_device_remote_llm_energy_rep_pkg_b200 import _lib   # noqa: E402
from cain_2025_device_remote_llm_energy_rep_pkg_b200.devmem import DevBuf as D   # noqa: E402


def main():
    lib = _lib.load()
    print("library:", _lib.LIB_PATH)
    for rows, cols, M, bias_mean in [(256, 4096, 64, 0.0), (128, 14336, 64, 0.0), (128, 14336, 64, 1.0), (256, 4096, 128, 0.0), (128, 14336, 128, 1.0)]:
        rng = np.random.default_rng(rows + cols + M)
        w = (rng.standard_normal((rows, cols)) * 0.02 + 0.02 * bias_mean).astype(np.float32)   # bias_mean 1: sums that do not cancel
        blob = Q.quantize_q4_0(w)
        x = (rng.standard_normal((M, cols)) * np.exp(rng.standard_normal((M, cols)))).astype(np.float32)
        if bias_mean:
            x = np.abs(x)
        dW, dx = pack(lib, D, blob, rows, cols), D.from_numpy(x)
        dy, dy1, dx1 = D(M * rows * 4), D(rows * 4), D(cols * 4)
        rc = lib.b2_op_prefill_gemm(2, dW.ptr, rows, cols, dx.ptr, M, None, 0.0, 0, None, 0, 0, dy.ptr)
        if rc:
            print(f"rows {rows} cols {cols} M {M}: error {lib.b2_last_error().decode()}")
            continue
        y = dy.to_numpy(np.float32, (M, rows))
        e_or = e_dec = e_rel = 0.0
        for m in sorted({0, M // 2, M - 1}):
            ref, mag = ref_gemv(blob, x[m])
            e_or = max(e_or, float(np.max(np.abs(y[m] - ref) / mag)))
            e_rel = max(e_rel, float(np.max(np.abs(y[m] - ref) / np.maximum(np.abs(ref), 1e-30))))
            lib.b2_dev_upload(dx1.ptr, x[m].ctypes.data, cols * 4)
            lib.b2_op_gemv_f32(2, dW.ptr, rows, cols, dx1.ptr, None, 0, dy1.ptr)
            g = dy1.to_numpy(np.float32, (rows,))
            e_dec = max(e_dec, float(np.max(np.abs(y[m] - g) / mag)))
        print(f"rows {rows:5d} cols {cols:6d} M {M:3d} {'same-sign' if bias_mean else 'random   '}: vs float64 {e_or:.2e} of sum|w||x| "
              f"(bar 6e-6; {e_rel:.2e} of |y|)   vs decode GEMV {e_dec:.2e} (bar 2e-6)", flush=True)


if __name__ == "__main__":
    main()
