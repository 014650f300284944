"""The tutorial Transformer language model, laid out for pipelining.

Same architecture as the reference's example (``main.py:24-73,139-157``): token
embedding scaled by sqrt(E) + sinusoidal positions + dropout, ``nlayers``
post-norm encoder layers (ReLU FFN, sequence-first, no attention mask), and a
vocabulary projection that hands back batch-first logits.  Micro-batching splits
dim 0, so the model takes ``(N, S)`` tokens, works ``(S, N, E)`` inside, and
returns ``(N, S, V)``.

Two bodies are offered for every block:

``impl="eager"``  stock ``torch.nn`` modules (``nn.Embedding``, ``nn.TransformerEncoderLayer``,
                  ``nn.Linear``) — semantically the reference, used as oracle and as
                  the on-box baseline.
``impl="fused"``  the sm_100a kernels of :mod:`pipe_b200.ops` (requires a B200).

``build_stages`` returns the "one nested ``nn.Sequential`` per device" shape that
``Pipe`` splits by device, mirroring ``main.py:139-157`` (``ceil(nlayers / n)``
layers per stage, encoder on the first device, decoder on the last).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Sequence, Union

import torch
from torch import Tensor, nn

__all__ = ["TransformerConfig", "PositionalEncoding", "Encoder", "Decoder", "build_layers",
           "build_layer", "build_stage", "build_stages", "count_parameters", "CONFIGS"]

Device = Union[torch.device, int, str]


@dataclass
class TransformerConfig:
    ntokens: int = 28782
    emsize: int = 4096
    nhid: int = 4096
    nlayers: int = 8
    nhead: int = 16
    dropout: float = 0.2
    bptt: int = 25
    batch: int = 20
    chunks: int = 8
    causal: bool = False   # the reference passes no mask (main.py:148)

    @property
    def tokens_per_step(self) -> int:
        return self.batch * self.bptt


#: BASELINE.json configs by short name.
CONFIGS = {
    "tiny": TransformerConfig(ntokens=1000, emsize=64, nhid=64, nlayers=2, nhead=2, chunks=4),
    "shipped": TransformerConfig(emsize=2048, nhid=2048, nlayers=16, nhead=32, bptt=128, batch=32, chunks=4),
    "cfg2": TransformerConfig(emsize=4096, nhid=4096, nlayers=8, nhead=16, chunks=8),
    "cfg3": TransformerConfig(emsize=4096, nhid=4096, nlayers=16, nhead=16, chunks=8),
    "cfg5": TransformerConfig(emsize=8192, nhid=8192, nlayers=32, nhead=64, chunks=32),
}


class PositionalEncoding(nn.Module):
    """Adds the fixed sin/cos table ``pe[:S]`` to a ``(S, N, E)`` input, then dropout."""

    def __init__(self, d_model: int, dropout: float = 0.1, max_len: int = 5000):
        super().__init__()
        self.dropout = nn.Dropout(p=dropout)
        position = torch.arange(max_len, dtype=torch.float32).unsqueeze(1)
        freq = torch.exp(torch.arange(0, d_model, 2, dtype=torch.float32) * (-math.log(10000.0) / d_model))
        pe = torch.zeros(max_len, 1, d_model)
        pe[:, 0, 0::2] = torch.sin(position * freq)
        pe[:, 0, 1::2] = torch.cos(position * freq)
        self.register_buffer("pe", pe)

    def forward(self, x: Tensor) -> Tensor:
        return self.dropout(x + self.pe[: x.size(0)])


class Encoder(nn.Module):
    """``(N, S)`` int64 tokens -> ``(S, N, E)`` embeddings."""

    def __init__(self, ntoken: int, ninp: int, dropout: float = 0.5):
        super().__init__()
        self.pos_encoder = PositionalEncoding(ninp, dropout)
        self.encoder = nn.Embedding(ntoken, ninp)
        self.ninp = ninp
        nn.init.uniform_(self.encoder.weight, -0.1, 0.1)

    def forward(self, src: Tensor) -> Tensor:
        src = src.t()                                   # sequence-first inside the model
        return self.pos_encoder(self.encoder(src) * math.sqrt(self.ninp))


class Decoder(nn.Module):
    """``(S, N, E)`` -> batch-first logits ``(N, S, V)``."""

    def __init__(self, ntoken: int, ninp: int):
        super().__init__()
        self.decoder = nn.Linear(ninp, ntoken)
        nn.init.zeros_(self.decoder.bias)
        nn.init.uniform_(self.decoder.weight, -0.1, 0.1)

    def forward(self, inp: Tensor) -> Tensor:
        return self.decoder(inp).permute(1, 0, 2)


def build_layer(cfg: TransformerConfig, index: int, impl: str = "eager",
                device: Optional[Device] = None, seed: int = 0) -> nn.Module:
    """Layer ``index`` of ``[Encoder, block x nlayers, Decoder]``, created directly on
    ``device``.  The RNG is re-seeded per layer, so a layer's initial weights do not
    depend on which stage (or rank) builds it."""
    dev = torch.device("cuda", device) if isinstance(device, int) else torch.device(device or "cpu")
    if dev.type == "cuda":
        with torch.cuda.device(dev):
            torch.manual_seed(seed * 1000 + index)
    else:
        torch.manual_seed(seed * 1000 + index)
    last = cfg.nlayers + 1
    with torch.device(dev):
        if impl == "eager":
            if index == 0:
                return Encoder(cfg.ntokens, cfg.emsize, cfg.dropout)
            if index == last:
                return Decoder(cfg.ntokens, cfg.emsize)
            return nn.TransformerEncoderLayer(cfg.emsize, cfg.nhead, cfg.nhid, cfg.dropout)
        if impl == "fused":
            from .fused import FusedDecoder, FusedEncoder, FusedEncoderLayer
            if index == 0:
                return FusedEncoder(cfg.ntokens, cfg.emsize, cfg.dropout)
            if index == last:
                return FusedDecoder(cfg.ntokens, cfg.emsize)
            return FusedEncoderLayer(cfg.emsize, cfg.nhead, cfg.nhid, cfg.dropout, causal=cfg.causal)
    raise ValueError(f"unknown impl '{impl}'")


def build_layers(cfg: TransformerConfig, impl: str = "eager", device: Optional[Device] = None,
                 seed: int = 0) -> List[nn.Module]:
    """Flat layer list ``[Encoder, block x nlayers, Decoder]``."""
    return [build_layer(cfg, k, impl, device, seed) for k in range(cfg.nlayers + 2)]


def stage_balance(nlayers: int, n_stages: int) -> List[int]:
    """Layer counts per stage over ``[Encoder, blocks..., Decoder]`` with
    ``ceil(nlayers / n)`` blocks per stage (reference ``main.py:140,147-157``)."""
    per = (nlayers - 1) // n_stages + 1
    counts = []
    left = nlayers
    for _ in range(n_stages):
        take = min(per, left)
        counts.append(take)
        left -= take
    counts = [c for c in counts if c > 0] or [0]
    counts[0] += 1      # encoder rides with the first stage
    counts[-1] += 1     # decoder with the last
    return counts


def build_stages(cfg: TransformerConfig, devices: Sequence[Device], impl: str = "eager",
                 dtype: Optional[torch.dtype] = None, seed: int = 0) -> nn.Sequential:
    """One nested ``nn.Sequential`` per device, ready for ``Pipe(...)``."""
    balance = stage_balance(cfg.nlayers, len(devices))
    groups = []
    at = 0
    for count, device in zip(balance, devices):
        group = nn.Sequential(*[build_layer(cfg, k, impl, device, seed) for k in range(at, at + count)])
        if dtype is not None:
            group.to(dtype)
        groups.append(group)
        at += count
    return nn.Sequential(*groups)


def build_stage(cfg: TransformerConfig, stage: int, n_stages: int, device: Device, impl: str = "eager",
                seed: int = 0) -> nn.Sequential:
    """Only stage ``stage`` of ``n_stages`` (for one-rank-per-stage layouts)."""
    balance = stage_balance(cfg.nlayers, n_stages)
    at = sum(balance[:stage])
    return nn.Sequential(*[build_layer(cfg, k, impl, device, seed) for k in range(at, at + balance[stage])])


def count_parameters(module: nn.Module) -> int:
    return sum(p.numel() for p in module.parameters())
