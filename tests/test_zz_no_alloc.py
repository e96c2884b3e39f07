"""b2_op_argmax_noalloc: the greedy sampler for callers that may not allocate on the device (a server's sampling thread
during a measured request — cudaMalloc/cudaFree stall the device and would show up in the recorded time and energy —
or a process that has already filled the card).  Its scratch is a fixed array inside the library, so every call on a
device shares it and starts from whatever the previous call left behind: the test makes two calls in a row and checks
each against b2_op_argmax (index work: bit-exact).  Collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_noalloc_argmax_matches_argmax_twice_in_a_row(lib):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.devmem import DevBuf
    rng = np.random.default_rng(20251106)
    out_ref, out_new = DevBuf(4), DevBuf(4)
    for n in (128256, 1000):                      # second call: different size, stale partials from the first in the scratch
        x = rng.standard_normal(n).astype(np.float32)
        i, j = sorted(rng.choice(n, 2, replace=False))
        x[i] = x[j] = x.max() + 1.0               # a tie: the lowest index wins in both entry points
        dx = DevBuf.from_numpy(x)
        assert lib.b2_op_argmax(dx.ptr, n, out_ref.ptr) == 0, lib.b2_last_error().decode()
        assert lib.b2_op_argmax_noalloc(dx.ptr, n, out_new.ptr) == 0, lib.b2_last_error().decode()
        assert lib.b2_dev_sync() == 0
        want = int(out_ref.to_numpy(np.int32, (1,))[0])
        got = int(out_new.to_numpy(np.int32, (1,))[0])
        assert got == want == int(i)


def test_noalloc_argmax_rejects_oversize_and_bad_arguments(lib):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.devmem import DevBuf
    d = DevBuf(64)
    assert lib.b2_op_argmax_noalloc(d.ptr, 1048576 + 1, d.ptr) == -1
    assert b"maximum" in lib.b2_last_error()
    assert lib.b2_op_argmax_noalloc(d.ptr, 0, d.ptr) == -1
    assert lib.b2_op_argmax_noalloc(None, 8, d.ptr) == -1
