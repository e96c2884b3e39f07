"""Prompt-evaluation throughput (tokens/s) of b2_prefill for a few prompt lengths; B2_NO_PREFILL_GEMM=1 gives the token-by-token path."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import Engine, Model
eng = Engine(0)
name = sys.argv[1] if len(sys.argv) > 1 else "llama3.1:8b"
a = ARCHS[name]
m = Model.random(eng, a, seed=1)
lens = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else [32, 64, 512, 2000]
s = m.session(min(max(lens) + 16, a.max_ctx if hasattr(a, "max_ctx") else 2048))
for n in lens:
    prompt = [1 + (i * 7919) % (a.V - 1) for i in range(n)]
    best = None
    for rep in range(3):
        s.reset()
        s.prefill(prompt)
        ns = s.stats()["prompt_eval_ns"]
        best = ns if best is None else min(best, ns)
    wbytes = a.weight_bytes_q4()
    per_pass = int(os.environ.get("B2_PF_TOKENS", "128"))   # tokens per pass over the weights (engine.cu: prefill_batched)
    passes = -(-n // per_pass)
    print(f"{name} prompt {n:5d} tokens: {best/1e6:9.3f} ms  {n/(best/1e9):10.0f} tokens/s   weight passes {passes}  "
          f"({passes*wbytes/(best/1e9)/1e9:7.0f} GB/s of weights)")
