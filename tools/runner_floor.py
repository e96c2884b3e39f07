#!/usr/bin/env python
"""How long does the UNCHANGED reference runner take for the run table when the server answers instantly?  The floor under any
multi-GPU sweep time: K runner processes (launch.py --shards K) against a server whose backends are test doubles without
arithmetic (tests/stub_backend.py, `delay` seconds per request).  Runs on the CPU; needs /root/reference or baseline/_ref.
    python tools/runner_floor.py [--reps 1] [--delay 0.0] [--shards 1,8]"""
import argparse
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from stub_backend import StubBackend   # noqa: E402
from cain_2025_device_remote_llm_energy_rep_pkg_b200 import server   # noqa: E402
from cain_2025_device_remote_llm_energy_rep_pkg_b200.runner import launch   # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=1)
    ap.add_argument("--delay", type=float, default=0.0)
    ap.add_argument("--shards", default="1,8")
    ap.add_argument("--reference-dir", default="/root/reference" if os.path.isdir("/root/reference") else os.path.join(ROOT, "baseline", "_ref"))
    args = ap.parse_args()
    for k in [int(v) for v in args.shards.split(",")]:
        backs = []
        for d in range(max(k, 1)):
            b = StubBackend(delay_s=args.delay)
            b.device = d
            backs.append(b)
        srv = server.serve(server.MultiBackend(backs), host="127.0.0.1", port=11434, verbose=False)
        work = tempfile.mkdtemp(prefix=f"floor_{k}_")
        t0 = time.time()
        try:
            argv = ["--reference-dir", args.reference_dir, "--work-dir", work, "--name", "floor", "--cooldown-ms", "0", "--repetitions", str(args.reps),
                    "--seed", "0"]
            if k > 1:
                argv += ["--shards", str(k), "--nvml-devices", ",".join(str(d) for d in range(k))]
            rc = launch.main(argv)
        finally:
            srv.shutdown()
            srv.server_close()
        n = sum(len(b.requests) for b in backs)
        print(f"shards {k}: rc {rc}, {n} requests, {time.time() - t0:.1f} s wall ({args.delay} s per request in the stub backend)", flush=True)


if __name__ == "__main__":
    main()
