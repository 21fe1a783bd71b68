"""tcgen05 implicit-GEMM convolutions (csrc/conv_tc.cu) against an fp32 ``F.conv2d`` of the same
bf16-rounded operands: forward (+ fused BatchNorm statistics), data gradient (plain and
accumulating), weight gradient - for every distinct convolution shape of ResNet-50 (stem
excluded: tests/test_kernels_gpu.py), odd image sizes, tiny batches, both CTA-pairing modes."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

CL = torch.channels_last

# (H = W of the input, Cin, Cout, k, stride): the distinct layers of ResNet-50 at 224x224
RESNET50 = [
    (56, 64, 64, 1, 1), (56, 64, 64, 3, 1), (56, 64, 256, 1, 1), (56, 256, 64, 1, 1),
    (56, 256, 128, 1, 1), (56, 128, 128, 3, 2), (28, 128, 512, 1, 1), (56, 256, 512, 1, 2),
    (28, 512, 128, 1, 1), (28, 128, 128, 3, 1),
    (28, 512, 256, 1, 1), (28, 256, 256, 3, 2), (14, 256, 1024, 1, 1), (28, 512, 1024, 1, 2),
    (14, 1024, 256, 1, 1), (14, 256, 256, 3, 1),
    (14, 1024, 512, 1, 1), (14, 512, 512, 3, 2), (7, 512, 2048, 1, 1), (14, 1024, 2048, 1, 2),
    (7, 2048, 512, 1, 1), (7, 512, 512, 3, 1),
]
# odd / tiny geometries: partial tiles in every dimension, 1-pixel images, batch < tile depth
ODD = [
    (15, 64, 128, 3, 2), (15, 128, 64, 1, 2), (9, 64, 64, 3, 1), (5, 128, 256, 3, 1),
    (2, 64, 64, 3, 1), (1, 64, 128, 1, 1), (4, 256, 128, 3, 2), (30, 64, 192, 3, 1),
]


def _fp32_ref():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False


def _mk(n, h, cin, cout, k, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn((n, cin, h, h), device="cuda", generator=g).bfloat16().contiguous(memory_format=CL)
    w = (torch.randn((cout, cin, k, k), device="cuda", generator=g) / (cin * k * k) ** 0.5)
    w = w.bfloat16().contiguous(memory_format=CL)
    return x, w


def _ids(cfgs):
    return [f"{h}x{h}_{ci}to{co}_k{k}s{s}" for h, ci, co, k, s in cfgs]


@pytest.mark.parametrize("cfg", RESNET50 + ODD, ids=_ids(RESNET50 + ODD))
@pytest.mark.parametrize("n", [8, 3])
def test_conv_forward_and_statistics(cfg, n):
    from tf_operator_b200.ops.conv import conv_fwd
    _fp32_ref()
    h, cin, cout, k, s = cfg
    x, w = _mk(n, h, cin, cout, k)
    stats = torch.zeros(2, cout, dtype=torch.float32, device="cuda")
    y = conv_fwd(x, w, s, stats)
    ref = F.conv2d(x.float(), w.float(), stride=s, padding=k // 2)
    assert y.shape == ref.shape and y.is_contiguous(memory_format=CL)
    torch.testing.assert_close(y.float(), ref, rtol=1e-2, atol=2e-2)
    # the statistics are those of the rounded output the kernel stored
    yf = y.float()
    m = yf.numel() // cout
    torch.testing.assert_close(stats[0], yf.sum(dim=(0, 2, 3)), rtol=1e-3, atol=1e-3 * m ** 0.5)
    torch.testing.assert_close(stats[1], (yf * yf).sum(dim=(0, 2, 3)), rtol=1e-3, atol=1e-3 * m ** 0.5)
    # a second call ADDS to the accumulator; without `stats` the output is unchanged
    y2 = conv_fwd(x, w, s, stats)
    assert torch.equal(y2, conv_fwd(x, w, s))
    torch.testing.assert_close(stats[0], 2 * yf.sum(dim=(0, 2, 3)), rtol=1e-3, atol=2e-3 * m ** 0.5)


@pytest.mark.parametrize("cfg", RESNET50 + ODD, ids=_ids(RESNET50 + ODD))
def test_conv_data_gradient(cfg):
    from tf_operator_b200.ops.conv import conv_dgrad, _out_hw
    _fp32_ref()
    h, cin, cout, k, s = cfg
    n = 5
    x, w = _mk(n, h, cin, cout, k, seed=1)
    oh, ow = _out_hw(h, h, k, s)
    dy = torch.randn((n, cout, oh, ow), device="cuda").bfloat16().contiguous(memory_format=CL)
    ref = torch.nn.grad.conv2d_input(x.shape, w.float(), dy.float(), stride=s, padding=k // 2)
    scale = (cout * k * k / (s * s) / (cin * k * k)) ** 0.5 + 1e-3
    dx = conv_dgrad(dy, w, x.shape, s)
    assert dx.is_contiguous(memory_format=CL)
    torch.testing.assert_close(dx.float(), ref, rtol=1e-2, atol=2e-2 * max(scale, 1.0))
    # accumulate: added (bf16) into what is there
    base = torch.randn_like(dx)
    acc = base.clone()
    conv_dgrad(dy, w, x.shape, s, out=acc, accumulate=True)
    want = base.float() + dx.float()
    torch.testing.assert_close(acc.float(), want, rtol=1e-2, atol=2e-2 * max(scale, 1.0))


@pytest.mark.parametrize("cfg", RESNET50 + ODD, ids=_ids(RESNET50 + ODD))
def test_conv_weight_gradient(cfg):
    from tf_operator_b200.ops.conv import conv_wgrad, _out_hw
    _fp32_ref()
    h, cin, cout, k, s = cfg
    n = 6
    x, w = _mk(n, h, cin, cout, k, seed=2)
    oh, ow = _out_hw(h, h, k, s)
    dy = torch.randn((n, cout, oh, ow), device="cuda").bfloat16().contiguous(memory_format=CL)
    ref = torch.nn.grad.conv2d_weight(x.float(), w.shape, dy.float(), stride=s, padding=k // 2)
    w32 = w.float()                      # fp32 "parameter": the gradient comes back unrounded
    dw = conv_wgrad(x, dy, w32, s)
    assert dw.shape == w.shape and dw.dtype == torch.float32
    m = n * oh * ow
    torch.testing.assert_close(dw, ref, rtol=2e-3, atol=2e-3 * m ** 0.5)
    dwb = conv_wgrad(x, dy, w, s)        # bf16 parameter: rounded once, same memory layout
    assert dwb.dtype == torch.bfloat16 and dwb.is_contiguous(memory_format=CL)
    torch.testing.assert_close(dwb.float(), ref, rtol=1e-2, atol=1e-2 * m ** 0.5)


@pytest.mark.parametrize("mode", [0, 1])
def test_conv_cta_pairing_modes_agree(mode):
    """cta_group::2 pairs and single CTAs compute the same tiles (bitwise: same MMA order)."""
    from tf_operator_b200.ops import _lib
    from tf_operator_b200.ops.conv import conv_fwd, conv_dgrad, conv_wgrad
    lib = _lib.load()
    x, w = _mk(8, 28, 128, 256, 3, seed=3)
    try:
        _lib.check(lib.tfop_conv_set_cluster_mode(-1))
        y0 = conv_fwd(x, w, 1)
        dy = torch.randn_like(y0)
        d0 = conv_dgrad(dy, w, x.shape, 1)
        g0 = conv_wgrad(x, dy, w.float(), 1)
        _lib.check(lib.tfop_conv_set_cluster_mode(mode))
        y1 = conv_fwd(x, w, 1)
        d1 = conv_dgrad(dy, w, x.shape, 1)
        g1 = conv_wgrad(x, dy, w.float(), 1)
    finally:
        _lib.check(lib.tfop_conv_set_cluster_mode(-1))
    assert torch.equal(y0, y1) and torch.equal(d0, d1)
    # the weight gradient sums fp32 partial tiles in whatever order they arrive
    torch.testing.assert_close(g0, g1, rtol=1e-4, atol=1e-3)


def test_conv_autograd_functions_match_torch():
    """conv2d / conv2d_skip / conv2d_pair: outputs and every gradient against torch autograd on
    fp32 copies, including the in-kernel accumulation of the shortcut gradient."""
    from tf_operator_b200.ops.conv import conv2d, conv2d_skip, conv2d_pair
    _fp32_ref()
    x, w1 = _mk(4, 14, 128, 64, 1, seed=4)
    _, w2 = _mk(4, 14, 128, 256, 1, seed=5)
    _, w3 = _mk(4, 14, 64, 64, 3, seed=6)

    def run(native):
        xs = (x if native else x.float()).detach().clone().requires_grad_(True)
        ws = [(w if native else w.float()).detach().clone().requires_grad_(True) for w in (w1, w2, w3)]
        if native:
            a, skip = conv2d_skip(xs, ws[0], 1)
            b, d = conv2d_pair(skip, ws[0], 1, None, ws[1], 2, None)
            c = conv2d(a + b, ws[2], 2)
        else:
            a, skip = F.conv2d(xs, ws[0]), xs
            b, d = F.conv2d(skip, ws[0]), F.conv2d(skip, ws[1], stride=2)
            c = F.conv2d(a + b, ws[2], stride=2, padding=1)
        loss = c.float().square().mean() + d.float().square().mean() + skip.float().mean()
        loss.backward()
        return [c, d, xs.grad] + [w.grad for w in ws]

    got, want = run(True), run(False)
    for g, r in zip(got, want):
        assert g.shape == r.shape
        tol = 3e-2 * float(r.abs().max()) + 1e-6
        torch.testing.assert_close(g.float(), r, rtol=3e-2, atol=tol)


def test_conv_rejects_what_the_kernels_do_not_cover():
    from tf_operator_b200.ops.conv import conv_fwd
    x, w = _mk(2, 8, 64, 64, 3)
    with pytest.raises(TypeError):
        conv_fwd(x.float(), w.float(), 1)
    with pytest.raises(ValueError):
        conv_fwd(x[:, :32], w[:, :32], 1)
    with pytest.raises(ValueError):
        conv_fwd(x, w, 3)


def test_fused_bn_takes_statistics_from_the_convolution():
    """FusedBatchNorm2d fed with the convolution's sums == FusedBatchNorm2d doing its own pass:
    same output, same running statistics, and the accumulator comes back zeroed."""
    from tf_operator_b200.ops.bn import FusedBatchNorm2d
    from tf_operator_b200.ops.conv import TcConv2d
    torch.manual_seed(7)
    conv = TcConv2d(128, 256, 3).cuda().bfloat16().to(memory_format=CL)
    bn_a, bn_b = FusedBatchNorm2d(256).cuda(), FusedBatchNorm2d(256).cuda()
    x = torch.randn(8, 128, 14, 14, device="cuda").bfloat16().contiguous(memory_format=CL)
    st = conv.sums(x)
    y = conv(x, st)
    assert float(st.abs().sum()) > 0
    za = bn_a(y, sums=st)
    zb = bn_b(y)
    assert float(st.abs().sum()) == 0.0
    torch.testing.assert_close(za.float(), zb.float(), rtol=1e-2, atol=1e-2)
    torch.testing.assert_close(bn_a.running_mean, bn_b.running_mean, rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(bn_a.running_var, bn_b.running_var, rtol=1e-3, atol=1e-4)


def test_resnet50_step_calls_no_library_contraction():
    """One training step of the bench model dispatches no ATen convolution / matmul at all: every
    contraction (53 convolutions + classifier, forward, dgrad and wgrad) goes through the
    library's own entry points, and the shortcut gradients are added inside the dgrad kernel."""
    from torch.utils._python_dispatch import TorchDispatchMode
    from tf_operator_b200.models.resnet import ResNet
    from tf_operator_b200.ops import _lib, conv as convmod
    from tf_operator_b200.parallel.engine import DataParallelEngine

    seen = set()

    class Spy(TorchDispatchMode):
        def __torch_dispatch__(self, func, types, args=(), kwargs=None):
            seen.add(str(func))
            return func(*args, **(kwargs or {}))

    torch.manual_seed(0)
    eng = DataParallelEngine(ResNet(50, num_classes=100), batch_size=8, image_size=64, lr=0.05,
                             graph=False, stem_s2d=True)
    x = torch.randint(0, 256, (8, 64, 64, 3), dtype=torch.uint8).pin_memory()
    y = torch.randint(0, 100, (8,), dtype=torch.int64).pin_memory()
    eng.step(x, y).result()
    before = dict(_lib.LAUNCHES)
    inplace0 = convmod.INPLACE_SKIP_ADDS
    with Spy():
        loss = eng.step(x, y).result()
    torch.cuda.synchronize()
    assert loss == loss
    bad = [n for n in seen if any(t in n for t in ("convolution", "cudnn", "aten.mm", "aten.addmm",
                                                   "aten.bmm", "aten.linear", "aten.matmul"))]
    assert not bad, bad
    ran = {k: _lib.LAUNCHES.get(k, 0) - before.get(k, 0) for k in _lib.LAUNCHES}
    # 52 convolutions: forward, data gradient and weight gradient each; the stem has its own kernels
    assert ran["conv_fwd"] == 52 and ran["conv_wgrad"] == 52 and ran["conv_dgrad"] == 52, ran
    assert ran["conv_stem_s2d_fwd"] == 1 and ran["conv_stem_s2d_wgrad"] == 1, ran
    assert ran["bn_finalize"] == 52 and ran.get("bn_fwd_stats", 0) == 1, ran
    # 12 identity blocks: their shortcut gradient was accumulated by the dgrad epilogue
    assert convmod.INPLACE_SKIP_ADDS - inplace0 == 12
