"""Language-model data plumbing of the example app, without torchtext.

The reference's driver tokenises WikiText-2 with torchtext (``main.py:76-113``); neither is
available offline, so this module offers the same three steps on any text — or on a synthetic
Zipfian corpus of the same vocabulary size when there is no text at all:

``build_corpus``   text file -> ``basic_english``-style tokens -> vocabulary with ``<unk>`` -> id tensor
``batchify``       1-D id stream -> ``(len // bsz, bsz)`` columns (reference ``main.py:92-105``)
``get_batch``      ``(N, S)`` batch-first inputs + flattened next-token targets (``main.py:107-113``)
"""
from __future__ import annotations

import re
from collections import Counter
from typing import Dict, Iterable, List, Optional, Tuple

import torch
from torch import Tensor

__all__ = ["basic_english", "Vocab", "build_corpus", "synthetic_corpus", "batchify", "get_batch"]

_PUNCT = re.compile(r"([.,!?()\"';:])")


def basic_english(line: str) -> List[str]:
    """Lower-case, space out punctuation, split on whitespace."""
    return _PUNCT.sub(r" \1 ", line.lower()).split()


class Vocab:
    def __init__(self, counter: Counter, specials: Iterable[str] = ("<unk>",), min_freq: int = 1) -> None:
        self.itos: List[str] = list(specials)
        self.itos += [w for w, c in sorted(counter.items(), key=lambda kv: (-kv[1], kv[0])) if c >= min_freq
                      and w not in self.itos]
        self.stoi: Dict[str, int] = {w: i for i, w in enumerate(self.itos)}
        self.default_index = self.stoi.get("<unk>", 0)

    def __len__(self) -> int:
        return len(self.itos)

    def __call__(self, tokens: Iterable[str]) -> List[int]:
        return [self.stoi.get(t, self.default_index) for t in tokens]


def build_corpus(path: str, vocab: Optional[Vocab] = None) -> Tuple[Tensor, Vocab]:
    """Token-id stream of a text file (empty lines dropped) and its vocabulary."""
    with open(path, encoding="utf-8") as f:
        lines = [basic_english(line) for line in f]
    if vocab is None:
        counter: Counter = Counter()
        for toks in lines:
            counter.update(toks)
        vocab = Vocab(counter)
    ids = [torch.tensor(vocab(toks), dtype=torch.long) for toks in lines if toks]
    return (torch.cat(ids) if ids else torch.zeros(0, dtype=torch.long)), vocab


def synthetic_corpus(n_tokens: int, vocab_size: int = 28782, seed: int = 0) -> Tensor:
    """A Zipf-distributed id stream with WikiText-2's vocabulary size (28 782)."""
    g = torch.Generator().manual_seed(seed)
    ranks = torch.arange(1, vocab_size + 1, dtype=torch.float64)
    probs = (1.0 / ranks) / (1.0 / ranks).sum()
    return torch.multinomial(probs, n_tokens, replacement=True, generator=g)


def batchify(data: Tensor, bsz: int, device: Optional[torch.device] = None) -> Tensor:
    """Trim the stream to a multiple of ``bsz`` and lay it out as ``(len // bsz, bsz)``: column ``b``
    is the ``b``-th contiguous slice of the stream."""
    nbatch = data.size(0) // bsz
    data = data.narrow(0, 0, nbatch * bsz).view(bsz, -1).t().contiguous()
    return data.to(device) if device is not None else data


def get_batch(source: Tensor, i: int, bptt: int) -> Tuple[Tensor, Tensor]:
    """Batch-first ``(N, S)`` inputs starting at row ``i`` and the flattened next tokens.  (Targets
    are flattened sequence-major while the model's output is batch-major — the reference's tutorial
    quirk, SURVEY.md Appendix B #13, kept as is.)"""
    seq_len = min(bptt, len(source) - 1 - i)
    data = source[i:i + seq_len]
    target = source[i + 1:i + 1 + seq_len].reshape(-1)
    return data.t(), target
