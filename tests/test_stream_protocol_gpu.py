"""Allocator stress for the copy-stream protocol: every cross-stream copy records its
source and destination on the streams that use them, and Wait fences order copy and compute streams.
If any of that were missing, memory recycled by the caching allocator while a copy or kernel is still in
flight would be overwritten — here with NaNs from a side stream that allocates, poisons and frees blocks
of the same sizes between and during steps.  Reference mechanisms: README.md:204-208, 215-217, 230-234;
pipeline.py:190-192, 251-254."""
import copy

import pytest
import torch
from torch import nn

pytestmark = pytest.mark.gpu


def _poison(device, sizes, stream):
    with torch.cuda.stream(stream):
        junk = [torch.full((s,), float("nan"), device=device) for s in sizes]
    del junk


@pytest.mark.parametrize("mode", ["never", "except_last", "always"])
def test_allocator_stress_keeps_results_exact(mode):
    from pipe_b200 import Pipe
    n_dev = torch.cuda.device_count()
    devices = [torch.device("cuda", k % n_dev) for k in range(4)]
    torch.manual_seed(0)
    width = 512
    layers = []
    for k in range(8):
        layers += [nn.Linear(width, width), nn.Tanh()]
    plain = nn.Sequential(*layers).to(devices[0])
    piped_layers = copy.deepcopy(plain)
    pipe = Pipe(piped_layers, chunks=8, checkpoint=mode, balance=[4, 4, 4, 4], devices=devices)
    side = [torch.cuda.Stream(d) for d in devices]
    sizes = [width * 4, width * 8, width * 64, 4 * width * width // 8]

    for step in range(12):
        x = torch.randn(64, width, device=devices[0])
        want = plain(x)
        want.square().mean().backward()
        for d, s in zip(devices, side):
            _poison(d, sizes, s)
        out = pipe(x).local_value()
        for d, s in zip(devices, side):
            _poison(d, sizes, s)
        out.square().mean().backward()
        for d, s in zip(devices, side):
            _poison(d, sizes, s)
        torch.cuda.synchronize()
        assert torch.isfinite(out).all()
        assert torch.allclose(out.to(devices[0]), want, rtol=1e-4, atol=1e-5), step
        for (name, p), q in zip(plain.named_parameters(), pipe.parameters()):
            assert torch.isfinite(q.grad).all(), name
            assert torch.allclose(q.grad.to(devices[0]), p.grad, rtol=1e-3, atol=1e-6), (step, name)
        plain.zero_grad(set_to_none=True)
        pipe.zero_grad(set_to_none=True)
        if step % 4 == 3:
            torch.cuda.empty_cache()
    pipe.close()
