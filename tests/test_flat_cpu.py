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
    # buckets tile the flat range without gaps; the no-decay tail (complete only when backward
    # ends) rides in the last bucket instead of costing a collective of its own
    assert fp.buckets[0].lo == 0 and fp.buckets[-1].hi == fp.total
    for a, b in zip(fp.buckets, fp.buckets[1:]):
        assert a.hi == b.lo
    for b in fp.buckets[:-1]:
        assert all(s.decay for s in b.slots)
    assert any(not s.decay for s in fp.buckets[-1].slots) and fp.buckets[-1].slots[0].decay
    assert sum(len(b.slots) for b in fp.buckets) == len(fp.slots)
    assert len(fp.buckets) >= 3


def test_is_dense():
    assert _is_dense(torch.zeros(4, 3, 2, 2))
    assert _is_dense(torch.zeros(4, 3, 2, 2).contiguous(memory_format=torch.channels_last))
    assert not _is_dense(torch.zeros(4, 6)[:, ::2])


def test_named_export_import_roundtrip_across_bucket_layouts():
    """A checkpoint is keyed by parameter name: it loads into an engine whose flat layout
    (bucket size, memory format) differs from the writer's."""
    torch.manual_seed(1)
    m1 = ResNet(18, num_classes=10, width=16, fused=False).to(memory_format=torch.channels_last)
    fp1 = FlatParams(m1, torch.device("cpu"), bucket_bytes=64 << 10)
    mom1 = torch.randn(fp1.total)
    named_w, named_m = fp1.export_named(fp1.master), fp1.export_named(mom1)
    assert set(named_w) == {n for n, _ in m1.named_parameters()}
    assert all(t.is_contiguous() and t.device.type == "cpu" for t in named_w.values())

    torch.manual_seed(2)
    m2 = ResNet(18, num_classes=10, width=16, fused=False)      # different init, NCHW layout
    fp2 = FlatParams(m2, torch.device("cpu"), bucket_bytes=1 << 20)
    mom2 = torch.zeros(fp2.total)
    assert fp2.import_named(fp2.master, named_w) == []
    fp2.refresh_weights()
    fp2.import_named(mom2, named_m)
    for (n1, p1), (n2, p2) in zip(m1.named_parameters(), m2.named_parameters()):
        assert n1 == n2 and torch.equal(p1.float(), p2.float()), n1
    for s in fp2.slots:
        assert torch.equal(fp2.view_of(mom2, s), named_m[s.name])
        assert torch.equal(fp2.view_of(fp2.master, s), named_w[s.name])
    # the exported weights load into a plain module too
    m3 = ResNet(18, num_classes=10, width=16, fused=False)
    missing, unexpected = m3.load_state_dict(named_w, strict=False)
    assert unexpected == [] and all("running_" in k or "num_batches" in k for k in missing)


def test_import_named_reports_missing_and_shape_errors():
    import pytest
    m = ResNet(18, num_classes=10, width=16, fused=False)
    fp = FlatParams(m, torch.device("cpu"))
    named = fp.export_named(fp.master)
    first = next(iter(named))
    partial = {k: v for k, v in named.items() if k != first}
    with pytest.raises(KeyError):
        fp.import_named(fp.master, partial)
    assert fp.import_named(fp.master, partial, strict=False) == [first]
    bad = dict(named)
    bad[first] = torch.zeros(3)
    with pytest.raises(ValueError):
        fp.import_named(fp.master, bad)


def test_fused_region_comes_first_and_belongs_to_no_bucket():
    """FlatParams(fused=[...]): parameters whose gradient arrives already summed over the ranks
    (GradTarget.fused) sit at the front of the flat buffers, outside every bucket; the rest is
    bucketed as before and the decay / no-decay boundary does not move."""
    import torch.nn as nn
    m = nn.Sequential(nn.Linear(64, 128), nn.Linear(128, 32), nn.LayerNorm(32), nn.Linear(32, 8))
    fused = [m[0].weight, m[1].weight]
    fp = FlatParams(m, torch.device("cpu"), bucket_bytes=1 << 10, fused=fused)
    plain = FlatParams(nn.Sequential(nn.Linear(64, 128), nn.Linear(128, 32), nn.LayerNorm(32), nn.Linear(32, 8)),
                       torch.device("cpu"), bucket_bytes=1 << 10)
    assert fp.total == plain.total and fp.n_decay == plain.n_decay
    assert fp.n_fused == 64 * 128 + 128 * 32
    assert {id(s.param) for s in fp.slots if s.offset < fp.n_fused} == {id(p) for p in fused}
    assert all(fp.bucket_of(p) is None for p in fused)
    assert fp.buckets[0].lo == fp.n_fused and fp.buckets[-1].hi == fp.total
    assert sum(len(b.slots) for b in fp.buckets) == len(fp.slots) - 2
    assert fp.bucket_of(m[3].weight) is not None and plain.n_fused == 0
