"""One short launch of the persistent token kernel for an ncu capture (ncu --kernel-name k_token_loop -c 1 ...)."""
import os, sys
os.environ["B2_MEGA"] = "1"
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import Engine, Model
eng = Engine(0)
a = ARCHS[sys.argv[1] if len(sys.argv) > 1 else "llama3.1:8b"]
m = Model.random(eng, a, seed=1)
s = m.session(256)
s.prefill(list(range(1, 1 + int(sys.argv[2]) if len(sys.argv) > 2 else 5)))
print("done", s.stats())
