"""DeferredBatchNorm."""
import pytest
import torch
from torch import nn

from pipe_b200 import Pipe
from pipe_b200.batchnorm import DeferredBatchNorm

CHUNKS = 4


def tilt(x):
    # make the chunks statistically different from one another
    return x + torch.arange(x.size(0), dtype=x.dtype).view(-1, 1, 1, 1) * 0.3


@pytest.mark.parametrize("momentum", [0.1, None])
def test_running_stats_match_plain_bn(momentum):
    bn = nn.BatchNorm2d(3, momentum=momentum)
    dbn = DeferredBatchNorm.convert_deferred_batch_norm(nn.BatchNorm2d(3, momentum=momentum), chunks=CHUNKS)
    for _ in range(3):
        x = tilt(torch.rand(8, 3, 4, 4))
        bn(x)
        for piece in x.chunk(CHUNKS):
            dbn(piece)
    assert torch.allclose(dbn.running_mean, bn.running_mean, atol=1e-5)
    # population variance vs unbiased: compare with a tolerance that covers n/(n-1)
    assert torch.allclose(dbn.running_var, bn.running_var, rtol=2e-2, atol=1e-3)
    assert int(dbn.num_batches_tracked) == 3


def test_normalises_with_micro_batch_stats_in_training_and_running_in_eval():
    dbn = DeferredBatchNorm(3, chunks=2)
    x = tilt(torch.rand(4, 3, 2, 2))
    y = dbn(x[:2])
    assert torch.allclose(y.mean(dim=(0, 2, 3)), torch.zeros(3), atol=1e-5)
    dbn.eval()
    y = dbn(x)
    want = nn.functional.batch_norm(x, dbn.running_mean, dbn.running_var, dbn.weight, dbn.bias, False, 0.0, dbn.eps)
    assert torch.equal(y, want)


def test_rejects_2d_input():
    with pytest.raises(ValueError):
        DeferredBatchNorm(3)(torch.rand(4, 3))


def test_conversion_keeps_parameters_and_is_idempotent():
    bn = nn.BatchNorm2d(3)
    seq = nn.Sequential(nn.Conv2d(3, 3, 1), bn, nn.Sequential(nn.BatchNorm2d(3)))
    out = DeferredBatchNorm.convert_deferred_batch_norm(seq, chunks=2)
    assert isinstance(out[1], DeferredBatchNorm) and isinstance(out[2][0], DeferredBatchNorm)
    assert out[1].weight is bn.weight and out[1].running_mean is bn.running_mean
    again = DeferredBatchNorm.convert_deferred_batch_norm(out, chunks=2)
    assert again[1] is out[1]


@pytest.mark.parametrize("mode", ["never", "always"])
def test_in_pipe_not_double_counted_under_recompute(mode):
    torch.manual_seed(0)
    bn = nn.BatchNorm2d(3)
    ref = nn.BatchNorm2d(3)
    pipe = Pipe(nn.Sequential(bn), chunks=CHUNKS, checkpoint=mode, deferred_batch_norm=True)
    assert isinstance(pipe[0], DeferredBatchNorm)
    x = tilt(torch.rand(8, 3, 2, 2)).requires_grad_()
    pipe(x).local_value().sum().backward()
    ref(x.detach())
    assert int(pipe[0].num_batches_tracked) == 1
    assert torch.allclose(pipe[0].running_mean, ref.running_mean, atol=1e-5)
