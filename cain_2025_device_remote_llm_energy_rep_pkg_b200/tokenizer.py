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


# ---------------------------------------------------------------------------------------------------------------------
# Tokenizers from GGUF metadata (gguf_io.TokenizerData): what Ollama applies inside /api/generate when it has the model file.
# "gpt2": byte-level BPE (llama 3, qwen2) — bytes mapped to printable code points, split by a pre-tokenizer regex, pairs merged
#         in rank order.  "llama": SentencePiece-style BPE (llama 2, mistral, gemma, phi3) — spaces become U+2581, adjacent
#         symbols are merged greedily by the highest score of the merged piece, unknown characters fall back to <0xXX> bytes.
def _bytes_to_unicode():
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(0xA1, 0xAD)) + list(range(0xAE, 0x100))
    cs, n = bs[:], 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    return dict(zip(bs, map(chr, cs)))


class GGUFTokenizer:
    def __init__(self, data, family: str = ""):
        import re
        self.data, self.family = data, family
        self.id = {t: i for i, t in enumerate(data.tokens)}
        self.bpe = data.model == "gpt2"
        self.b2u = _bytes_to_unicode()
        self.u2b = {v: k for k, v in self.b2u.items()}
        self.rank = {tuple(m.split(" ", 1)): i for i, m in enumerate(data.merges)}
        # llama-3 style pre-tokenizer without \p classes (Python's re has none): contractions, letter runs, digit groups of
        # up to three, punctuation runs, whitespace
        self.pat = re.compile(r"'(?:[sdmt]|ll|ve|re)| ?[^\W\d_]+|\d{1,3}| ?[^\s\w]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+")
        self.special = sorted((t for t, ty in zip(data.tokens, data.token_types or []) if ty in (3, 4) and len(t) > 1), key=len, reverse=True)

    # -- byte-level BPE
    def _bpe_word(self, word: str):
        parts = [self.b2u[b] for b in word.encode("utf-8")]
        while len(parts) > 1:
            best, bi = None, -1
            for i in range(len(parts) - 1):
                r = self.rank.get((parts[i], parts[i + 1]))
                if r is not None and (best is None or r < best):
                    best, bi = r, i
            if best is None:
                break
            parts[bi:bi + 2] = [parts[bi] + parts[bi + 1]]
        return parts

    # -- SentencePiece-style BPE
    def _spm(self, text: str):
        sym = list(text.replace(" ", "▁"))
        score = lambda p: self.data.scores[self.id[p]] if p in self.id and self.data.scores else (0.0 if p in self.id else None)
        while len(sym) > 1:
            best, bi = None, -1
            for i in range(len(sym) - 1):
                s = score(sym[i] + sym[i + 1])
                if s is not None and (best is None or s > best):
                    best, bi = s, i
            if best is None:
                break
            sym[bi:bi + 2] = [sym[bi] + sym[bi + 1]]
        out = []
        for p in sym:
            if p in self.id:
                out.append(self.id[p])
            else:
                out += [self.id.get("<0x%02X>" % b, 0) for b in p.encode("utf-8")]
        return out

    def _plain(self, text: str):
        if not text:
            return []
        if self.bpe:
            ids = []
            for w in self.pat.findall(text):
                for p in self._bpe_word(w):
                    if p in self.id:
                        ids.append(self.id[p])
                    else:                      # not in the vocabulary as a whole: its bytes are (byte-level BPE covers all 256)
                        ids += [self.id[c] for c in p if c in self.id]
            return ids
        return self._spm(text)

    def encode(self, text: str, template: bool = True):
        if template and self.family in TEMPLATES:
            text = TEMPLATES[self.family].format(prompt=text)
        ids = [self.data.bos_id] if (self.data.add_bos and self.data.bos_id is not None) else []
        if not self.bpe:
            text = " " + text if not text.startswith(" ") else text      # SentencePiece's dummy prefix
        chunk, i = "", 0
        while i < len(text):                   # special tokens (<|eot_id|>, <start_of_turn>, ...) are matched verbatim
            sp = next((s for s in self.special if text.startswith(s, i)), None)
            if sp is None:
                chunk += text[i]
                i += 1
            else:
                ids += self._plain(chunk) + [self.id[sp]]
                chunk, i = "", i + len(sp)
        return ids + self._plain(chunk)

    def decode(self, ids) -> str:
        toks = [self.data.tokens[int(i)] for i in ids if 0 <= int(i) < len(self.data.tokens)]
        if self.bpe:
            raw = bytearray()
            for t in toks:
                for c in t:
                    raw += bytes([self.u2b[c]]) if c in self.u2b else c.encode("utf-8")
            return raw.decode("utf-8", "replace")
        raw = bytearray()
        for t in toks:
            if len(t) == 6 and t.startswith("<0x") and t.endswith(">"):
                raw.append(int(t[3:5], 16))
            else:
                raw += t.replace("▁", " ").encode("utf-8")
        return raw.decode("utf-8", "replace")

    def is_eos(self, tok: int) -> bool:
        return self.data.eos_id is not None and int(tok) == int(self.data.eos_id)
