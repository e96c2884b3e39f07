"""B200-native generation engine behind an Ollama-compatible endpoint.

Drop-in for the one path the reference replication package spends its time on: the autoregressive
generation that happens inside `POST /api/generate` (reference call site
`experiment/RunnerConfig.py:128-131`).  The compute lives in `libb2llm.so` (hand-written sm_100a CUDA,
C ABI in `include/b2llm.h`); this package is the thin host side: ctypes binding (`_lib`, `engine`),
the seven architecture records (`archs`), the byte-level stand-in tokenizer (`tokenizer`), the HTTP
server (`server`) and the experiment-runner plugin and shims (`runner/`).
"""
from .archs import ARCHS, TINY, Arch  # noqa: F401

__version__ = "0.1.0"
