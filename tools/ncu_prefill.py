"""One batched prompt evaluation for an ncu launch list / capture."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import Engine, Model
eng = Engine(0)
a = ARCHS[sys.argv[1] if len(sys.argv) > 1 else "llama3.1:8b"]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
m = Model.random(eng, a, seed=1)
s = m.session(256)
s.prefill([1 + (i * 7919) % (a.V - 1) for i in range(n)])
print("done", s.stats())
