"""Out-of-bounds write checks without a sanitizer (compute-sanitizer is not available on the
GPU pool): every output of a kernel is carved out of a larger buffer pre-filled with a sentinel,
and the bytes in front of and behind it must be untouched afterwards. Odd sizes on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SENT = 0x5A


class Arena:
    """One byte buffer; `take` returns a 256-byte aligned tensor view with a 4 KB sentinel band on
    both sides."""

    def __init__(self, nbytes=512 << 20):
        self.buf = torch.full((nbytes,), SENT, dtype=torch.uint8, device="cuda")
        self.off = 4096
        self.regions = []

    def take(self, shape, dtype):
        n = 1
        for s in shape:
            n *= s
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        start = (self.off + 255) // 256 * 256
        self.regions.append((start, nbytes))
        self.off = start + nbytes + 4096
        assert self.off < self.buf.numel()
        return self.buf[start:start + nbytes].view(dtype).view(shape)

    def check(self):
        torch.cuda.synchronize()
        mask = torch.ones_like(self.buf, dtype=torch.bool)
        for start, nbytes in self.regions:
            mask[start:start + nbytes] = False
        bad = (self.buf != SENT) & mask
        assert not bool(bad.any()), f"out-of-bounds write at byte offsets {bad.nonzero()[:8].flatten().tolist()}"


def test_gemm_outputs_stay_in_bounds():
    from tf_operator_b200.ops.gemm import gemm
    ar = Arena()
    torch.manual_seed(0)
    for (m, n, k, a_t, b_t, dt) in [(1000, 520, 200, False, False, torch.bfloat16),
                                    (129, 1000, 72, False, True, torch.bfloat16),
                                    (264, 136, 1000, True, True, torch.float32),
                                    (77, 40, 24, False, False, torch.float32),
                                    (300, 1002, 128, False, False, torch.bfloat16)]:   # direct stores
        a = torch.randn((k, m) if a_t else (m, k), device="cuda").bfloat16()
        b = torch.randn((k, n) if b_t else (n, k), device="cuda").bfloat16()
        out = ar.take((m, n), dt)
        gemm(a, b, a_t=a_t, b_t=b_t, out=out)
        ref = (a.float().t() if a_t else a.float()) @ (b.float() if b_t else b.float().t())
        torch.testing.assert_close(out.float(), ref, rtol=3e-2, atol=3e-2 * k ** 0.5)
    ar.check()


def test_layernorm_gelu_xent_outputs_stay_in_bounds():
    lib_ops = __import__("tf_operator_b200.ops._lib", fromlist=["x"])
    lib = lib_ops.load()
    ar = Arena()
    torch.manual_seed(1)
    rows, H = 37, 1032                      # odd row count, hidden not a multiple of 256
    h = torch.randn(rows, H, device="cuda").bfloat16()
    res = torch.randn(rows, H, device="cuda").bfloat16()
    bias = torch.randn(H, device="cuda").bfloat16()
    g = torch.ones(H, device="cuda").bfloat16()
    b = torch.zeros(H, device="cuda").bfloat16()
    y, s = ar.take((rows, H), torch.bfloat16), ar.take((rows, H), torch.bfloat16)
    mean, rstd = ar.take((rows,), torch.float32), ar.take((rows,), torch.float32)
    st = torch.cuda.current_stream().cuda_stream
    assert lib.tfop_ln_fwd(h.data_ptr(), bias.data_ptr(), res.data_ptr(), g.data_ptr(), b.data_ptr(),
                           y.data_ptr(), s.data_ptr(), mean.data_ptr(), rstd.data_ptr(), rows, H, 1e-5, st) == 0
    dx = ar.take((rows, H), torch.bfloat16)
    cols = ar.take((3, H), torch.float32)
    cols.zero_()
    dy = torch.randn(rows, H, device="cuda").bfloat16()
    assert lib.tfop_ln_bwd(dy.data_ptr(), s.data_ptr(), mean.data_ptr(), rstd.data_ptr(), g.data_ptr(),
                           dx.data_ptr(), cols[0].data_ptr(), cols[1].data_ptr(), cols[2].data_ptr(),
                           rows, H, st) == 0
    # tiled backward (H % 256 == 0) with a row count that is not a multiple of the trip size
    rows2, H2 = 1003, 768
    dy2 = torch.randn(rows2, H2, device="cuda").bfloat16()
    s2 = torch.randn(rows2, H2, device="cuda").bfloat16()
    m2, r2 = torch.zeros(rows2, device="cuda"), torch.ones(rows2, device="cuda")
    g2 = torch.ones(H2, device="cuda").bfloat16()
    dx2, cols2 = ar.take((rows2, H2), torch.bfloat16), ar.take((3, H2), torch.float32)
    cols2.zero_()
    assert lib.tfop_ln_bwd(dy2.data_ptr(), s2.data_ptr(), m2.data_ptr(), r2.data_ptr(), g2.data_ptr(),
                           dx2.data_ptr(), cols2[0].data_ptr(), cols2[1].data_ptr(), cols2[2].data_ptr(),
                           rows2, H2, st) == 0
    # bias + GeLU forward / backward and the pure column sum
    rows3, H3 = 91, 520
    x3 = torch.randn(rows3, H3, device="cuda").bfloat16()
    b3 = torch.randn(H3, device="cuda").bfloat16()
    y3, dh3, db3 = ar.take((rows3, H3), torch.bfloat16), ar.take((rows3, H3), torch.bfloat16), ar.take((H3,), torch.float32)
    db3.zero_()
    assert lib.tfop_bias_gelu_fwd(x3.data_ptr(), b3.data_ptr(), y3.data_ptr(), rows3, H3, st) == 0
    assert lib.tfop_bias_act_bwd(y3.data_ptr(), x3.data_ptr(), b3.data_ptr(), dh3.data_ptr(), db3.data_ptr(),
                                 rows3, H3, 1, st) == 0
    db4 = ar.take((H3,), torch.float32)
    db4.zero_()
    assert lib.tfop_bias_act_bwd(y3.data_ptr(), 0, 0, 0, db4.data_ptr(), rows3, H3, 0, st) == 0
    # cross-entropy with a vocabulary that is not a multiple of 8
    rows5, V = 29, 1027
    logits = torch.randn(rows5, V, device="cuda").bfloat16()
    labels = torch.randint(0, V, (rows5,), device="cuda")
    labels[::3] = -100
    inv_n = torch.tensor([1.0 / 20], device="cuda")
    loss, lse = ar.take((rows5,), torch.float32), ar.take((rows5,), torch.float32)
    dlog = ar.take((rows5, V), torch.bfloat16)
    one = torch.ones(1, device="cuda")
    assert lib.tfop_xent_fwd(logits.data_ptr(), V, labels.data_ptr(), rows5, V, -100, inv_n.data_ptr(),
                             loss.data_ptr(), lse.data_ptr(), st) == 0
    assert lib.tfop_xent_bwd(logits.data_ptr(), V, labels.data_ptr(), rows5, V, -100, inv_n.data_ptr(),
                             lse.data_ptr(), one.data_ptr(), dlog.data_ptr(), V, st) == 0
    ar.check()
    assert bool(torch.isfinite(y.float()).all()) and bool(torch.isfinite(dx.float()).all())


def test_conv_stem_outputs_stay_in_bounds():
    from tf_operator_b200.ops import _lib
    from tf_operator_b200.ops.preproc import stem_space_to_depth, stem_weight_s2d
    lib = _lib.load()
    ar = Arena()
    n, img = 3, 96                                  # OH = OW = 48: 6 x 3 tiles per image
    u8 = torch.randint(0, 256, (n, img, img, 3), dtype=torch.uint8, device="cuda")
    x = stem_space_to_depth(u8)
    w4 = stem_weight_s2d((torch.randn(64, 3, 7, 7, device="cuda") * 0.05).bfloat16())
    oh = ow = img // 2
    y = ar.take((n, oh, ow, 64), torch.bfloat16)
    dw = ar.take((64, 4, 4, 16), torch.float32)
    dw.zero_()
    st = torch.cuda.current_stream().cuda_stream
    assert lib.tfop_conv_stem_s2d_fwd(x.data_ptr(), w4.data_ptr(), y.data_ptr(), n, oh, ow, st) == 0
    dy = torch.randn(n, oh, ow, 64, device="cuda").bfloat16()
    assert lib.tfop_conv_stem_s2d_wgrad(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), n, oh, ow, st) == 0
    ar.check()
    ref = torch.nn.functional.conv2d(x.float(), w4.float())
    torch.testing.assert_close(y.permute(0, 3, 1, 2).float(), ref, rtol=2e-2, atol=3e-2)
