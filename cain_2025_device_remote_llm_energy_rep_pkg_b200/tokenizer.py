"""Byte-level stand-in tokenizer and prompt template.

Ollama applies the model's chat template and its BPE/SPM vocabulary inside /api/generate;
neither vocabularies nor templates are available offline, so prompts are encoded byte-per-token (ids 3..258;
0 pad, 1 bos, 2 eos) and generated ids — which range over the architecture's true vocabulary because the LM
head keeps its real size — are rendered as printable ASCII so the HTTP response stays valid JSON text."""
from __future__ import annotations

from typing import List

PAD, BOS, EOS = 0, 1, 2
BYTE0 = 3

TEMPLATES = {
    "llama": "<|start_header_id|>user<|end_header_id|>\n\n{prompt}<|eot_id|><|start_header_id|>assistant<|end_header_id|>\n\n",
    "mistral": "[INST] {prompt} [/INST]",
    "qwen2": "<|im_start|>user\n{prompt}<|im_end|>\n<|im_start|>assistant\n",
    "gemma": "<start_of_turn>user\n{prompt}<end_of_turn>\n<start_of_turn>model\n",
    "phi3": "<|user|>\n{prompt}<|end|>\n<|assistant|>\n",
}


def encode(text: str, vocab: int, family: str = "", template: bool = True) -> List[int]:
    if template and family in TEMPLATES:
        text = TEMPLATES[family].format(prompt=text)
    ids = [BOS] + [BYTE0 + b for b in text.encode("utf-8")]
    return [i % vocab for i in ids]


def decode(ids) -> str:
    return "".join(chr(32 + (int(i) % 95)) for i in ids)
