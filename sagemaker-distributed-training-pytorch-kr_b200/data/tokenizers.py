"""Tokenizers behind `--tokenizer-type` (reference: Megatron's tokenizer package selected at arguments.py:1145-1153,
default injected by pretrain_gpt.py:149 = GPT2BPETokenizer with --vocab-file / --merge-file, log 3_...ipynb:1211-1212).

All expose the small Megatron surface: vocab_size, vocab, inv_vocab, tokenize(text)->ids, detokenize(ids)->text, eod
(+ cls/sep/pad/mask for the BERT ones).  Written from the published algorithm descriptions (byte-level BPE: GPT-2
paper 2.2; WordPiece greedy longest-match) - pure Python, no network, no downloaded files.
"""
from __future__ import annotations

import json
import unicodedata
from functools import lru_cache
from typing import Dict, List, Tuple

try:                                    # `regex` understands \p{L}; it is in the image, `re` is the degraded fallback
    import regex as _re
    _PAT = _re.compile(r"""'s|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+""")
except ImportError:                     # pragma: no cover
    import re as _re
    _PAT = _re.compile(r"""'s|'t|'re|'ve|'m|'ll|'d| ?[^\W\d_]+| ?\d+| ?[^\s\w]+|\s+(?!\S)|\s+""")


@lru_cache()
def byte_unicode_table() -> Dict[int, str]:
    """reversible byte -> printable unicode char map: printable latin-1 bytes map to themselves, the remaining 68
    bytes to code points 256.., so BPE merges never see whitespace/control characters."""
    keep = list(range(ord("!"), ord("~") + 1)) + list(range(0xA1, 0xAD)) + list(range(0xAE, 0x100))
    table, extra = {}, 0
    for b in range(256):
        if b in keep:
            table[b] = chr(b)
        else:
            table[b] = chr(256 + extra)
            extra += 1
    return table


class GPT2BPETokenizer:
    def __init__(self, vocab_file: str, merge_file: str, errors: str = "replace"):
        with open(vocab_file, encoding="utf-8") as f:
            self.vocab: Dict[str, int] = json.load(f)
        self.inv_vocab = {v: k for k, v in self.vocab.items()}
        with open(merge_file, encoding="utf-8") as f:
            lines = f.read().split("\n")
        merges = [tuple(l.split()) for l in lines if l and not l.startswith("#version")]
        self.ranks: Dict[Tuple[str, str], int] = {m: i for i, m in enumerate(m for m in merges if len(m) == 2)}
        self.b2u = byte_unicode_table()
        self.u2b = {v: k for k, v in self.b2u.items()}
        self.errors = errors
        self._cache: Dict[str, List[str]] = {}
        self.eod_id = self.vocab.get("<|endoftext|>", len(self.vocab) - 1)

    @property
    def vocab_size(self) -> int:
        return len(self.vocab)

    @property
    def eod(self) -> int:
        return self.eod_id

    def _bpe(self, token: str) -> List[str]:
        hit = self._cache.get(token)
        if hit is not None:
            return hit
        word = list(token)
        while len(word) > 1:
            best, best_rank = None, None
            for i in range(len(word) - 1):
                r = self.ranks.get((word[i], word[i + 1]))
                if r is not None and (best_rank is None or r < best_rank):
                    best, best_rank = (word[i], word[i + 1]), r
            if best is None:
                break
            a, b = best
            out, i = [], 0
            while i < len(word):
                if i < len(word) - 1 and word[i] == a and word[i + 1] == b:
                    out.append(a + b)
                    i += 2
                else:
                    out.append(word[i])
                    i += 1
            word = out
        self._cache[token] = word
        return word

    def tokenize(self, text: str) -> List[int]:
        ids: List[int] = []
        for piece in _PAT.findall(text):
            mapped = "".join(self.b2u[b] for b in piece.encode("utf-8"))
            ids.extend(self.vocab[t] for t in self._bpe(mapped))
        return ids

    def detokenize(self, ids) -> str:
        text = "".join(self.inv_vocab[int(i)] for i in ids)
        return bytearray(self.u2b[c] for c in text).decode("utf-8", errors=self.errors)


class BertWordPieceTokenizer:
    """basic (whitespace/punctuation, optional lower-casing + accent stripping) then greedy longest-match WordPiece."""

    def __init__(self, vocab_file: str, lower_case: bool = True, max_chars_per_word: int = 100):
        self.vocab: Dict[str, int] = {}
        with open(vocab_file, encoding="utf-8") as f:
            for i, line in enumerate(f):
                self.vocab[line.rstrip("\n")] = i
        self.inv_vocab = {v: k for k, v in self.vocab.items()}
        self.lower, self.maxc = lower_case, max_chars_per_word
        self.cls, self.sep = self.vocab.get("[CLS]"), self.vocab.get("[SEP]")
        self.pad, self.mask = self.vocab.get("[PAD]"), self.vocab.get("[MASK]")
        self.unk = self.vocab.get("[UNK]", 0)

    @property
    def vocab_size(self):
        return len(self.vocab)

    @property
    def eod(self):
        return self.sep

    @staticmethod
    def _is_punct(ch: str) -> bool:
        cp = ord(ch)
        return (33 <= cp <= 47 or 58 <= cp <= 64 or 91 <= cp <= 96 or 123 <= cp <= 126) or unicodedata.category(ch).startswith("P")

    def _basic(self, text: str) -> List[str]:
        if self.lower:
            text = "".join(c for c in unicodedata.normalize("NFD", text.lower()) if unicodedata.category(c) != "Mn")
        words, cur = [], []
        for ch in text:
            if ch.isspace():
                if cur:
                    words.append("".join(cur)); cur = []
            elif self._is_punct(ch):
                if cur:
                    words.append("".join(cur)); cur = []
                words.append(ch)
            else:
                cur.append(ch)
        if cur:
            words.append("".join(cur))
        return words

    def tokenize(self, text: str) -> List[int]:
        ids = []
        for w in self._basic(text):
            if len(w) > self.maxc:
                ids.append(self.unk); continue
            start, sub = 0, []
            while start < len(w):
                end, found = len(w), None
                while start < end:
                    piece = ("##" if start else "") + w[start:end]
                    if piece in self.vocab:
                        found = piece; break
                    end -= 1
                if found is None:
                    sub = None; break
                sub.append(self.vocab[found]); start = end
            ids.extend(sub if sub is not None else [self.unk])
        return ids

    def detokenize(self, ids) -> str:
        out = []
        for i in ids:
            t = self.inv_vocab[int(i)]
            if t.startswith("##") and out:
                out[-1] += t[2:]
            else:
                out.append(t)
        return " ".join(out)


class SentencePieceTokenizer:
    """thin wrapper over the `sentencepiece` wheel (--tokenizer-model)."""

    def __init__(self, model_file: str):
        import sentencepiece
        self.sp = sentencepiece.SentencePieceProcessor(model_file=model_file)
        self.vocab = {self.sp.id_to_piece(i): i for i in range(self.sp.get_piece_size())}
        self.inv_vocab = {v: k for k, v in self.vocab.items()}

    @property
    def vocab_size(self):
        return self.sp.get_piece_size()

    @property
    def eod(self):
        return self.sp.eos_id()

    def tokenize(self, text):
        return self.sp.encode(text)

    def detokenize(self, ids):
        return self.sp.decode([int(i) for i in ids])


def build(args):
    t = args.tokenizer_type
    if t == "GPT2BPETokenizer":
        return GPT2BPETokenizer(args.vocab_file, args.merge_file)
    if t in ("BertWordPieceLowerCase", "BertWordPieceCase"):
        return BertWordPieceTokenizer(args.vocab_file, lower_case=t.endswith("LowerCase"))
    if t in ("SentencePieceTokenizer", "GPTSentencePieceTokenizer"):
        return SentencePieceTokenizer(args.tokenizer_model)
    raise NotImplementedError(f"--tokenizer-type {t}")
