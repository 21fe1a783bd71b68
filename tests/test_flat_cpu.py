"""Flat parameter buffers (CPU): layout, aliasing, buckets."""
import torch

from tf_operator_b200.models.resnet import ResNet
from tf_operator_b200.parallel.flat import FlatParams, _is_dense


def test_flat_params_alias_and_layout():
    torch.manual_seed(0)
    m = ResNet(18, num_classes=10, width=16, fused=False).to(memory_format=torch.channels_last)
    ref = {n: p.detach().clone() for n, p in m.named_parameters()}
    fp = FlatParams(m, torch.device("cpu"), bucket_bytes=64 << 10)
    assert fp.total % 8 == 0 and fp.n_decay % 8 == 0 and 0 < fp.n_decay < fp.total
    for s in fp.slots:
        assert s.offset % 8 == 0
        assert s.param.dtype == torch.bfloat16
        assert s.param.shape == ref[s.name].shape
        torch.testing.assert_close(s.param.float(), ref[s.name].bfloat16().float())
        # parameter storage aliases the flat weight buffer
        assert s.param.data_ptr() == fp.weights.data_ptr() + 2 * s.offset
        assert s.decay == (s.param.ndim > 1)
    # writing the flat buffer changes the model
    fp.weights.zero_()
    assert all(float(p.abs().sum()) == 0.0 for p in m.parameters())
    # buckets tile the flat range without gaps and never mix decay / no-decay slots
    assert fp.buckets[0].lo == 0 and fp.buckets[-1].hi == fp.total
    for a, b in zip(fp.buckets, fp.buckets[1:]):
        assert a.hi == b.lo
    for b in fp.buckets:
        assert len({s.decay for s in b.slots}) == 1
    assert len(fp.buckets) >= 3


def test_is_dense():
    assert _is_dense(torch.zeros(4, 3, 2, 2))
    assert _is_dense(torch.zeros(4, 3, 2, 2).contiguous(memory_format=torch.channels_last))
    assert not _is_dense(torch.zeros(4, 6)[:, ::2])
