"""Test double for the server's backend: same interface as server.B2Backend, no arithmetic (it returns a
fixed token pattern).  Lives under tests/ — the package has no CPU backend."""
import time

from cain_2025_device_remote_llm_energy_rep_pkg_b200 import tokenizer
from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
from cain_2025_device_remote_llm_energy_rep_pkg_b200.server import default_num_predict


class StubBackend:
    def __init__(self, delay_s: float = 0.0):
        self.delay_s = delay_s
        self.requests = []
        self._loaded = []

    def list_models(self):
        return list(ARCHS)

    def gpu_status(self, energy=True):
        """Side channel of the real backend (GET /b2/gpus): a fake energy counter that advances 1 J per call, 42 % utilisation."""
        self._mj = getattr(self, "_mj", 5_000_000) + 1000
        return {str(getattr(self, "device", 0)): {"energy_mj": self._mj, "util": 42}}

    def loaded(self):
        return self._loaded

    def generate(self, tag, prompt, options, raw=False):
        a = ARCHS[tag]
        if tag not in self._loaded:
            self._loaded.append(tag)
        ptoks = tokenizer.encode(prompt, a.V, a.family, template=not raw)
        budget = int(options.get("num_predict", default_num_predict(prompt)))
        self.requests.append((tag, prompt, dict(options)))
        time.sleep(self.delay_s)
        toks = [(7 * i + len(prompt)) % a.V for i in range(budget)]
        if options.get("_stream"):
            for i in range(0, budget, 16):
                yield {"tokens": toks[i:i + 16]}
        else:
            yield {"tokens": toks}
        yield {"done": True, "done_reason": "length", "prompt_tokens": ptoks, "tokens": toks, "total_duration": 1000, "load_duration": 0,
               "prompt_eval_count": len(ptoks), "prompt_eval_duration": 10, "eval_count": budget, "eval_duration": max(1, budget) * 1000000,
               "energy_mj": 1234, "device": getattr(self, "device", 0), "prompt": options.get("_prompt", prompt)}
