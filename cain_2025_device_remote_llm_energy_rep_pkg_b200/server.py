"""Ollama-compatible HTTP front end (drop-in boundary b1).

The unchanged reference issues, per run, `curl http://<host>:11434/api/generate -d '{"model": ..., "prompt": ...,
"stream": false}'` (`experiment/RunnerConfig.py:128-131`): an HTTP/1.1 POST whose JSON body arrives with
`Content-Type: application/x-www-form-urlencoded` (curl -d default), so the body is parsed as JSON whatever the
header says.  The response is Ollama's single JSON object (fields in SURVEY.md §3.3); durations are ns.

The server owns no arithmetic: every token comes from a backend object; `B2Backend` drives libb2llm.so (the
CUDA engine) through ctypes.  There is no CPU backend in this package.
"""
from __future__ import annotations

import argparse
import datetime
import json
import math
import os
import re
import threading
import time
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Dict, Iterator, List, Optional

from . import tokenizer
from .archs import ARCHS, Arch

DEFAULT_PORT = 11434
WORDS_RE = re.compile(r"\bIn (\d+) words\b")


def default_num_predict(prompt: str) -> int:
    """Token budget when the request has no options.num_predict.  Random-init weights never emit a meaningful
    EOS, so the budget follows the reference's prompt grammar "In N words, ..." (RunnerConfig.py:120):
    ceil(1.33 * N) tokens; 128 for any other prompt."""
    m = WORDS_RE.search(prompt)
    return int(math.ceil(1.33 * int(m.group(1)))) if m else 128


# Ollama's defaults when a request asks for sampling without giving every knob (external, unpinned)
OLLAMA_TOP_K, OLLAMA_TOP_P, OLLAMA_TEMPERATURE = 40, 0.9, 0.8


OLLAMA_REPEAT_PENALTY, OLLAMA_REPEAT_LAST_N = 1.1, 64


def repeat_from_options(options: dict, default_sampling: bool = False):
    """options.{repeat_penalty, repeat_last_n} -> (penalty, window).  Off (1.0) unless the request names a penalty or the
    server applies Ollama's defaults (--ollama-sampling: 1.1 over the last 64 tokens)."""
    try:
        pen = float(options.get("repeat_penalty", OLLAMA_REPEAT_PENALTY if default_sampling else 1.0))
        last = int(options.get("repeat_last_n", OLLAMA_REPEAT_LAST_N))
    except (TypeError, ValueError) as ex:
        raise ValueError(f"bad repeat option: {ex}")
    if pen <= 0.0 or last == 0:
        return 1.0, 0
    return pen, (1 << 30) if last < 0 else last      # Ollama: -1 = the whole context


def sampler_from_options(options: dict, default_sampling: bool = False):
    """options.{temperature, top_k, top_p, seed} -> (top_k, temperature, top_p, seed) or None for greedy.
    A request without any of them is greedy (deterministic, what the parity tests pin) unless the server was started
    with --ollama-sampling, which applies Ollama's stochastic defaults as the reference's requests got them."""
    keys = ("temperature", "top_k", "top_p", "seed")
    if not default_sampling and not any(k in options for k in keys):
        return None
    try:
        temperature = float(options.get("temperature", OLLAMA_TEMPERATURE))
        top_k = int(options.get("top_k", OLLAMA_TOP_K))
        top_p = float(options.get("top_p", OLLAMA_TOP_P))
        seed = int(options.get("seed", 0))
    except (TypeError, ValueError) as ex:
        raise ValueError(f"bad sampling option: {ex}")
    if temperature <= 0.0 or top_k == 1:
        return None                      # Ollama: temperature 0 = greedy
    if top_k <= 0 or top_k > 64:
        top_k = 64                       # the device sampler keeps at most 64 candidates (0 = "disabled" in Ollama)
    if not (0.0 < top_p <= 1.0):
        top_p = 1.0
    return top_k, temperature, top_p, seed & 0xFFFFFFFFFFFFFFFF


# Chat forms of the five model families (what their GGUF files carry in tokenizer.chat_template, restated as
# (per-message format by role, generation prompt)); `{c}` is the message content.
CHAT_FORMS = {
    "llama": ({"*": "<|start_header_id|>{role}<|end_header_id|>\n\n{c}<|eot_id|>"}, "<|start_header_id|>assistant<|end_header_id|>\n\n"),
    "qwen2": ({"*": "<|im_start|>{role}\n{c}<|im_end|>\n"}, "<|im_start|>assistant\n"),
    "gemma": ({"user": "<start_of_turn>user\n{c}<end_of_turn>\n", "system": "<start_of_turn>user\n{c}<end_of_turn>\n",
               "assistant": "<start_of_turn>model\n{c}<end_of_turn>\n"}, "<start_of_turn>model\n"),
    "phi3": ({"*": "<|{role}|>\n{c}<|end|>\n"}, "<|assistant|>\n"),
    "mistral": ({"user": "[INST] {c} [/INST]", "system": "[INST] {c} [/INST]", "assistant": "{c}</s>"}, ""),
}


def chat_family(family: str, template: str = "") -> str:
    """The chat form to apply: recognised from the GGUF file's tokenizer.chat_template when it carries one (its marker
    tokens decide), else the architecture family's."""
    for fam, marker in (("llama", "<|start_header_id|>"), ("qwen2", "<|im_start|>"), ("gemma", "<start_of_turn>"),
                        ("phi3", "<|assistant|>"), ("mistral", "[INST]")):
        if marker in (template or ""):
            return fam
    return family


def chat_prompt(messages, family: str = "", template: str = "") -> str:
    """/api/chat: the conversation rendered with the model's chat form (already templated: encode with template=False)."""
    if not isinstance(messages, list) or not messages:
        raise ValueError("messages must be a non-empty list")
    for m in messages:
        if not isinstance(m, dict) or not isinstance(m.get("content", ""), str):
            raise ValueError("each message needs a string 'content'")
    fam = chat_family(family, template)
    if fam not in CHAT_FORMS:
        return "\n".join(f"{m.get('role', 'user')}: {m.get('content', '')}" for m in messages) + "\nassistant:"
    forms, gen = CHAT_FORMS[fam]
    out = ""
    for m in messages:
        role = str(m.get("role", "user"))
        out += forms.get(role, forms.get("*", "{c}")).format(role=role, c=m.get("content", ""))
    return out + gen


def parse_keep_alive(v, default: float) -> float:
    """Ollama's keep_alive: seconds as a number, or "10m" / "30s" / "1h"; negative = keep for good; 0 = unload at once."""
    if v is None:
        return default
    if isinstance(v, (int, float)):
        return float(v)
    m = re.fullmatch(r"\s*(-?\d+(?:\.\d+)?)\s*(ms|s|m|h)?\s*", str(v))
    if not m:
        raise ValueError(f"bad keep_alive: {v!r}")
    return float(m.group(1)) * {"ms": 1e-3, "s": 1.0, "m": 60.0, "h": 3600.0, None: 1.0}[m.group(2)]


class GenerationResult:
    def __init__(self, tokens: List[int], prompt_tokens: List[int], stats: dict, done_reason: str):
        self.tokens, self.prompt_tokens, self.stats, self.done_reason = tokens, prompt_tokens, stats, done_reason


class B2Backend:
    """Model residency + one serial session per model on one GPU (requests are serialised per backend)."""

    def __init__(self, device: int = 0, seed: int = 1, qtype: str = "q4_0", stream_chunk: int = 16, default_sampling: bool = False,
                 gguf: Optional[Dict[str, str]] = None, requantize: bool = False, keep_alive: float = 300.0):
        from .engine import Engine
        self.engine = Engine(device)
        self.device, self.seed, self.qtype, self.stream_chunk = device, seed, qtype, stream_chunk
        self.default_sampling = default_sampling
        self.gguf = dict(gguf or {})          # tag -> GGUF file: real weights and tokenizer instead of random init + byte stand-in
        self.requantize = requantize
        self.keep_alive = keep_alive          # seconds a model stays resident after its last request (Ollama: 5 minutes); < 0: for good
        self.tokenizers: Dict[str, object] = {}
        self.templates: Dict[str, str] = {}   # tag -> tokenizer.chat_template of its GGUF file
        self.models: Dict[str, tuple] = {}
        self.expiry: Dict[str, float] = {}    # tag -> time.monotonic() after which the model may be unloaded
        self.lock = threading.Lock()

    def chat_template(self, tag: str) -> str:
        return self.templates.get(tag, "")

    def family(self, tag: str) -> str:
        try:
            return self.arch(tag).family
        except KeyError:
            return "unknown"

    def _unload(self, tag: str):
        model, sess, _ = self.models.pop(tag)
        self.expiry.pop(tag, None)
        sess.close()
        model.close()

    def _evict_expired(self, keep: str):
        now = time.monotonic()
        for tag in [t for t, e in self.expiry.items() if t != keep and 0 <= e < now]:
            self._unload(tag)

    def list_models(self) -> List[str]:
        return list(ARCHS) + [t for t in self.gguf if t not in ARCHS]

    def loaded(self) -> List[str]:
        return list(self.models)

    def arch(self, tag: str) -> Arch:
        if tag in self.models:
            return self.models[tag][0].arch
        if tag not in ARCHS:
            raise KeyError(tag)
        return ARCHS[tag]

    def _load(self, tag: str):
        from .engine import Model
        if tag not in self.models:
            t0 = time.perf_counter_ns()
            if tag in self.gguf:
                from . import gguf_io
                g = gguf_io.load(self.gguf[tag], requantize=self.requantize, name=tag)
                a = g.arch
                model = Model.from_tensors(self.engine, a, g.tensors)
                if g.tokenizer.tokens:
                    self.tokenizers[tag] = tokenizer.GGUFTokenizer(g.tokenizer, a.family)
                self.templates[tag] = g.tokenizer.chat_template or ""
            else:
                a = self.arch(tag)
                model = Model.random(self.engine, a, seed=self.seed, qtype=self.qtype)
            sess = model.session(a.max_ctx)
            # one throw-away generation: the CUDA graphs of both attention variants (contexts below / above 32 positions) and of the
            # multi-token decode are captured here, not inside the first measured request (0.8 s instead of 0.2 s on the first hit)
            try:
                sess.prefill([1, 2, 3])
                sess.decode(min(48, a.max_ctx - 4))
                sess.reset()
            except Exception:
                pass
            self.models[tag] = (model, sess, time.perf_counter_ns() - t0)
            return self.models[tag], self.models[tag][2]
        return self.models[tag], 0

    def energy_mj(self) -> Optional[int]:
        from .engine import nvml_energy_mj
        try:
            return nvml_energy_mj(self.device)
        except Exception:
            return None

    def gpu_status(self, energy: bool = True) -> dict:
        """NVML counters of this backend's GPU, read in the server process (which holds NVML open) for the runner's measurement
        processes, which are forked per run.  The energy counter costs 4.5 ms per read (the utilisation 25 us): pollers ask without it."""
        import ctypes as C
        lib = self.engine.lib
        gpu, mem = C.c_uint32(), C.c_uint32()
        ok = lib.b2_nvml_util(self.device, C.byref(gpu), C.byref(mem)) == 0
        return {str(self.device): {"energy_mj": self.energy_mj() if energy else None, "util": gpu.value if ok else None}}

    def generate(self, tag: str, prompt: str, options: dict, raw: bool = False) -> Iterator[dict]:
        """Yields {"tokens": [...]} chunks, then a final {"done": True, ...stats}."""
        with self.lock:
            t_start = time.perf_counter_ns()
            e0 = self.energy_mj()
            keep_s = parse_keep_alive(options.get("_keep_alive"), self.keep_alive)
            self._evict_expired(tag)
            (model, sess, _), load_ns = self._load(tag)
            a = model.arch
            tk = self.tokenizers.get(tag)
            ptoks = tk.encode(prompt, template=not raw) if tk else tokenizer.encode(prompt, a.V, a.family, template=not raw)
            detok = tk.decode if tk else tokenizer.decode
            budget = int(options.get("num_predict", default_num_predict(prompt)))
            try:   # options.num_ctx: the context window of this request (Ollama default 2048 = the session's size; never larger)
                ctx = min(a.max_ctx, max(1, int(options.get("num_ctx", a.max_ctx))))
            except (TypeError, ValueError) as ex:
                raise ValueError(f"bad num_ctx: {ex}")
            room = ctx - len(ptoks)
            if room <= 0:
                raise ValueError(f"prompt of {len(ptoks)} tokens does not fit the {ctx}-token context")
            reason = "length"
            if budget < 0 or budget > room:
                budget = room
            samp = sampler_from_options(options, self.default_sampling)
            cfg = None
            if samp is not None:
                from ._lib import SamplerCfg
                pen, last = repeat_from_options(options, self.default_sampling)
                cfg = SamplerCfg(top_k=samp[0], temperature=samp[1], top_p=samp[2], seed=samp[3], repeat_penalty=pen, repeat_last_n=last)
            sess.reset()
            sess.prefill(ptoks)
            out: List[int] = []
            while len(out) < budget:
                n = min(self.stream_chunk if options.get("_stream") else budget, budget - len(out))
                toks = [int(t) for t in sess.decode(n, sampler=cfg)]
                if tk and any(tk.is_eos(t) for t in toks):       # a real vocabulary has a real end-of-sequence token
                    toks = toks[:next(i for i, t in enumerate(toks) if tk.is_eos(t))]
                    reason = "stop"
                out += toks
                yield {"tokens": toks, "text": detok(toks)}
                if reason == "stop":
                    break
            st = sess.stats()
            e1 = self.energy_mj()
            self.expiry[tag] = -1.0 if keep_s < 0 else time.monotonic() + keep_s
            if keep_s == 0:
                self._unload(tag)
            yield {"done": True, "done_reason": reason, "prompt_tokens": ptoks, "tokens": out, "text": detok(out),
                   "total_duration": time.perf_counter_ns() - t_start, "load_duration": load_ns,
                   "prompt_eval_count": st["prompt_eval_count"], "prompt_eval_duration": st["prompt_eval_ns"],
                   "eval_count": st["eval_count"], "eval_duration": st["eval_ns"],
                   "energy_mj": (e1 - e0) if (e0 is not None and e1 is not None) else None, "device": self.device,
                   "prompt": options.get("_prompt", prompt)}


class MultiBackend:
    """One backend per GPU behind ONE endpoint: the unchanged reference always calls port 11434
    (experiment/RunnerConfig.py:130), so K runner processes that share a box all arrive here.  Every GPU may hold every
    model (the seven tags are 21.4 GB of Q4 weights together against 180 GB per GPU); a request goes to an idle GPU that
    already has the model resident, else to the least-loaded GPU (which then loads it).  Backends serialise their own
    requests; different GPUs run concurrently (the C ABI calls release the GIL).  No collective anywhere."""

    def __init__(self, backends):
        self.backends = list(backends)
        if not self.backends:
            raise ValueError("MultiBackend needs at least one backend")
        self.active = [0] * len(self.backends)
        self._lock = threading.Lock()

    def list_models(self) -> List[str]:
        return self.backends[0].list_models()

    def loaded(self) -> List[str]:
        out = []
        for b in self.backends:
            out += [t for t in b.loaded() if t not in out]
        return out

    def family(self, tag: str) -> str:
        return self.backends[0].family(tag) if hasattr(self.backends[0], "family") else "unknown"

    def chat_template(self, tag: str) -> str:
        b = self.backends[0]
        return b.chat_template(tag) if hasattr(b, "chat_template") else ""

    def busy_devices(self) -> List[int]:
        with self._lock:
            return [getattr(b, "device", i) for i, b in enumerate(self.backends) if self.active[i] > 0]

    def gpu_status(self, energy: bool = True) -> dict:
        out = {}
        for b in self.backends:
            if hasattr(b, "gpu_status"):
                out.update(b.gpu_status(energy))
        return out

    def pick(self, tag: str) -> int:
        with self._lock:
            i = min(range(len(self.backends)),
                    key=lambda j: (self.active[j], 0 if tag in self.backends[j].loaded() else 1, len(self.backends[j].loaded()), j))
            self.active[i] += 1
            return i

    def generate(self, tag: str, prompt: str, options: dict, raw: bool = False) -> Iterator[dict]:
        i = self.pick(tag)
        try:
            yield from self.backends[i].generate(tag, prompt, options, raw=raw)
        finally:
            with self._lock:
                self.active[i] -= 1


def parse_devices(spec: str) -> List[int]:
    """"0-7" / "0,2,5" / "0-3,6" -> device ordinals."""
    out: List[int] = []
    for part in filter(None, (p.strip() for p in spec.split(","))):
        if "-" in part:
            a, b = part.split("-", 1)
            out += list(range(int(a), int(b) + 1))
        else:
            out.append(int(part))
    if not out or len(set(out)) != len(out) or min(out) < 0:
        raise ValueError(f"bad device list: {spec!r}")
    return out


def _now() -> str:
    return datetime.datetime.now(datetime.timezone.utc).isoformat().replace("+00:00", "Z")


class Handler(BaseHTTPRequestHandler):
    protocol_version = "HTTP/1.1"
    server_version = "b2llm"

    def log_message(self, fmt, *args):  # requests are logged by the server object; the side channel's polls are not
        if not str(self.path).startswith("/b2/"):
            self.server.log(fmt % args)

    # ------------------------------------------------------------------ helpers
    def _send_json(self, code: int, obj, extra_headers=None):
        body = (json.dumps(obj) if not isinstance(obj, (bytes, str)) else obj)
        body = body.encode() if isinstance(body, str) else body
        self.send_response(code)
        self.send_header("Content-Type", "application/json; charset=utf-8")
        self.send_header("Content-Length", str(len(body)))
        for k, v in (extra_headers or {}).items():
            self.send_header(k, v)
        self.end_headers()
        self.wfile.write(body)

    def _error(self, code: int, msg: str):
        self.server.log(f"error {code}: {msg}")
        self._send_json(code, {"error": msg})

    def _read_body(self) -> bytes:
        n = int(self.headers.get("Content-Length") or 0)
        return self.rfile.read(n) if n else b""

    # ------------------------------------------------------------------ routes
    def do_GET(self):
        b = self.server.backend
        if self.path in ("/", ""):
            body = b"Ollama is running"
            self.send_response(200)
            self.send_header("Content-Type", "text/plain; charset=utf-8")
            self.send_header("Content-Length", str(len(body)))
            self.end_headers()
            self.wfile.write(body)
        elif self.path == "/api/version":
            self._send_json(200, {"version": "0.0.0-b2llm"})
        elif self.path == "/api/tags":
            fam = lambda m: b.family(m) if hasattr(b, "family") else (ARCHS[m].family if m in ARCHS else "unknown")
            self._send_json(200, {"models": [{"name": m, "model": m, "details": {"quantization_level": "Q4_0", "family": fam(m)}}
                                             for m in b.list_models()]})
        elif self.path == "/api/ps":
            self._send_json(200, {"models": [{"name": m, "model": m} for m in b.loaded()]})
        elif self.path.startswith("/b2/active"):   # devices serving a request right now (the powermetrics stand-in samples those)
            self._send_json(200, {"devices": b.busy_devices() if hasattr(b, "busy_devices") else []})
        elif self.path.startswith("/b2/gpus"):     # NVML energy counters and utilisation per GPU, for the runner's per-run processes
            self._send_json(200, {"gpus": b.gpu_status("util_only" not in self.path) if hasattr(b, "gpu_status") else {},
                                  "busy": b.busy_devices() if hasattr(b, "busy_devices") else []})
        elif self.path.startswith("/b2/stats"):
            self._send_json(200, {"requests": self.server.stats_log[-int(self.server.keep_stats):]})
        else:
            self._error(404, f"unknown path {self.path}")

    def do_POST(self):
        if self.path not in ("/api/generate", "/api/chat"):
            self._read_body()
            return self._error(404, f"unknown path {self.path}")
        chat = self.path == "/api/chat"
        raw = self._read_body()
        try:
            req = json.loads(raw.decode("utf-8"))  # content-type deliberately ignored (curl -d sends form-urlencoded)
            if not isinstance(req, dict):
                raise ValueError("body must be a JSON object")
        except Exception as ex:
            return self._error(400, f"invalid JSON body: {ex}")
        model, prompt = req.get("model"), req.get("prompt", "")
        if not isinstance(model, str) or not model:
            return self._error(400, "missing field: model")
        backend = self.server.backend
        if model not in backend.list_models():
            return self._error(404, f"model '{model}' not found")
        raw_prompt = bool(req.get("raw", False))
        if chat:
            try:
                fam = backend.family(model) if hasattr(backend, "family") else (ARCHS[model].family if model in ARCHS else "")
                tmpl = backend.chat_template(model) if hasattr(backend, "chat_template") else ""
                prompt = chat_prompt(req.get("messages"), fam, tmpl)
                raw_prompt = chat_family(fam, tmpl) in CHAT_FORMS   # the conversation already carries the model's chat form
            except ValueError as ex:
                return self._error(400, str(ex))
        if not isinstance(prompt, str):
            return self._error(400, "prompt must be a string")
        stream = req.get("stream", True)
        options = dict(req.get("options") or {})
        options["_stream"] = bool(stream)
        options["_prompt"] = prompt
        if "keep_alive" in req:
            options["_keep_alive"] = req.get("keep_alive")
        headers_sent = False
        try:
            if "_keep_alive" in options:
                parse_keep_alive(options["_keep_alive"], 0.0)
            gen = backend.generate(model, prompt, options, raw=raw_prompt)
            if stream:
                first = next(gen)            # a request the engine rejects still gets a plain 4xx/5xx answer
                self.send_response(200)
                self.send_header("Content-Type", "application/x-ndjson")
                self.send_header("Transfer-Encoding", "chunked")
                self.end_headers()
                headers_sent = True
                import itertools
                for part in itertools.chain([first], gen):
                    if part.get("done"):
                        line = self._final(model, part, "", chat)
                    elif chat:
                        line = {"model": model, "created_at": _now(), "message": {"role": "assistant", "content": part.get("text", tokenizer.decode(part["tokens"]))},
                                "done": False}
                    else:
                        line = {"model": model, "created_at": _now(), "response": part.get("text", tokenizer.decode(part["tokens"])), "done": False}
                    data = (json.dumps(line) + "\n").encode()
                    self.wfile.write(b"%x\r\n%s\r\n" % (len(data), data))
                    self.wfile.flush()
                self.wfile.write(b"0\r\n\r\n")
            else:
                final = None
                for part in gen:
                    if part.get("done"):
                        final = part
                self._send_json(200, self._final(model, final, final.get("text", tokenizer.decode(final["tokens"])), chat))
        except Exception as ex:  # engine failure: Ollama convention is {"error": ...} with a 5xx
            code, msg = (400, str(ex)) if isinstance(ex, ValueError) else (500, f"{type(ex).__name__}: {ex}")
            if not headers_sent:
                return self._error(code, msg)
            # the chunked 200 answer is already on the wire: one NDJSON error record, then the terminating chunk
            self.server.log(f"error {code} in mid-stream: {msg}")
            try:
                data = (json.dumps({"error": msg}) + "\n").encode()
                self.wfile.write(b"%x\r\n%s\r\n0\r\n\r\n" % (len(data), data))
                self.wfile.flush()
            except OSError:
                pass

    def _final(self, model: str, part: dict, text: str, chat: bool = False) -> dict:
        rec = {"model": model, "created_at": _now(), "response": text, "done": True, "done_reason": part["done_reason"],
               "context": part["prompt_tokens"] + part["tokens"], "total_duration": part["total_duration"],
               "load_duration": part["load_duration"], "prompt_eval_count": part["prompt_eval_count"],
               "prompt_eval_duration": part["prompt_eval_duration"], "eval_count": part["eval_count"],
               "eval_duration": part["eval_duration"]}
        if chat:   # /api/chat answers with a message object instead of response/context
            rec["message"] = {"role": "assistant", "content": rec.pop("response")}
            rec.pop("context")
        self.server.record({"t": time.time(), "model": model, "eval_count": part["eval_count"], "eval_duration": part["eval_duration"],
                            "prompt_eval_count": part["prompt_eval_count"], "prompt_eval_duration": part["prompt_eval_duration"],
                            "total_duration": part["total_duration"], "load_duration": part["load_duration"],
                            "energy_mj": part.get("energy_mj"), "device": part.get("device"), "prompt": part.get("prompt")})
        return rec


class OllamaServer(ThreadingHTTPServer):
    daemon_threads = True
    allow_reuse_address = True

    def __init__(self, addr, backend, stats_file: Optional[str] = None, verbose: bool = True):
        super().__init__(addr, Handler)
        self.backend, self.stats_file, self.verbose = backend, stats_file, verbose
        self.stats_log: List[dict] = []
        self.keep_stats = 256
        self._lock = threading.Lock()

    def log(self, msg: str):
        if self.verbose:
            print(f"[b2llm-server] {msg}", flush=True)

    def record(self, rec: dict):
        """Side channel for the runner's energy plugin: the reference discards curl's stdout
        (RunnerConfig.py:131), so token counts reach run_table.csv through this log instead."""
        with self._lock:
            self.stats_log.append(rec)
            del self.stats_log[:-1024]
            if self.stats_file:
                with open(self.stats_file, "a") as f:
                    f.write(json.dumps(rec) + "\n")


def serve(backend, host: str = "0.0.0.0", port: int = DEFAULT_PORT, stats_file: Optional[str] = None, verbose: bool = True) -> OllamaServer:
    srv = OllamaServer((host, port), backend, stats_file, verbose)
    t = threading.Thread(target=srv.serve_forever, daemon=True)
    t.start()
    return srv


def main():
    ap = argparse.ArgumentParser(description="Ollama-compatible /api/generate endpoint on the B200 engine")
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=DEFAULT_PORT)
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--devices", default=None, help='several GPUs behind this one endpoint, e.g. "0-7" or "0,2,3": one backend per GPU, '
                    "requests go to an idle GPU that holds the model, else to the least-loaded one")
    ap.add_argument("--keep-alive", default="5m", help="how long a model stays resident after its last request (Ollama default 5m; -1: for good)")
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--stats-file", default=os.environ.get("B2_STATS_FILE"))
    ap.add_argument("--preload", default="", help="comma-separated model tags to load before serving")
    ap.add_argument("--ollama-sampling", action="store_true",
                    help="requests without sampling options use Ollama's stochastic defaults (top_k 40, top_p 0.9, temperature 0.8) instead of greedy")
    ap.add_argument("--gguf", action="append", default=[], metavar="TAG=PATH",
                    help="serve TAG from a GGUF file (Q4_0 matrices; real tokenizer and EOS) instead of random-init weights")
    ap.add_argument("--requantize", action="store_true", help="convert non-Q4_0 matrices of a GGUF file to Q4_0 at load (lossy)")
    a = ap.parse_args()
    devs = parse_devices(a.devices) if a.devices else [a.device]
    backs = [B2Backend(d, a.seed, default_sampling=a.ollama_sampling, gguf=dict(x.split("=", 1) for x in a.gguf), requantize=a.requantize,
                       keep_alive=parse_keep_alive(a.keep_alive, 300.0)) for d in devs]
    for tag in filter(None, a.preload.split(",")):
        for b in backs:
            b._load(tag)
    backend = MultiBackend(backs) if len(backs) > 1 else backs[0]
    srv = OllamaServer((a.host, a.port), backend, a.stats_file)
    srv.log(f"listening on {a.host}:{a.port} (device{'s' if len(devs) > 1 else ''} {','.join(map(str, devs))})")
    try:
        srv.serve_forever()
    except KeyboardInterrupt:
        pass


if __name__ == "__main__":
    main()
