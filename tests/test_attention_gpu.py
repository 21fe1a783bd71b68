"""tcgen05 attention (csrc/attention.cu) against fp32 softmax(QK^T/8)V of the same bf16-rounded
packed qkv: forward, log-sum-exp, and the packed gradient [dQ | dK | dV]."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(qkv32, heads):
    b, s, h3 = qkv32.shape
    d = h3 // (3 * heads)
    q, k, v = qkv32.view(b, s, 3, heads, d).unbind(2)
    q, k, v = (t.transpose(1, 2) for t in (q, k, v))             # [b, heads, s, d]
    sc = (q @ k.transpose(-1, -2)) * d ** -0.5
    p = torch.softmax(sc, dim=-1)
    o = (p @ v).transpose(1, 2).reshape(b, s, heads * d)
    return o, torch.logsumexp(sc, dim=-1)


@pytest.mark.parametrize("b,s,heads", [(2, 128, 2), (1, 512, 4), (3, 256, 16), (2, 384, 1)])
def test_attention_forward_and_backward(b, s, heads):
    from tf_operator_b200.ops.attention import attention
    torch.backends.cuda.matmul.allow_tf32 = False
    g = torch.Generator(device="cuda").manual_seed(b * 1000 + s + heads)
    qkv = (torch.randn((b, s, 3 * heads * 64), device="cuda", generator=g) * 1.5).bfloat16()
    dout = torch.randn((b, s, heads * 64), device="cuda", generator=g).bfloat16()
    x = qkv.detach().clone().requires_grad_(True)
    out = attention(x, heads)
    lse2 = out.grad_fn.saved_tensors[2].clone()      # base-2 log-sum-exp kept for the backward
    out.backward(dout)
    x32 = qkv.float().requires_grad_(True)
    ref, lse = _ref(x32, heads)
    ref.backward(dout.float())
    assert out.shape == ref.shape and out.dtype == torch.bfloat16
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(lse2 * 0.6931471805599453, lse, rtol=1e-3, atol=1e-2)
    gq, gk, gv = x.grad.float().view(b, s, 3, -1).unbind(2)
    rq, rk, rv = x32.grad.view(b, s, 3, -1).unbind(2)
    for name, got, want in (("dV", gv, rv), ("dK", gk, rk), ("dQ", gq, rq)):
        tol = 3e-2 * float(want.abs().max())
        torch.testing.assert_close(got, want, rtol=3e-2, atol=tol, msg=lambda m, n=name: f"{n}: {m}")


def test_attention_handles_large_logits_without_overflow():
    """Scores far beyond exp's fp32 range: the exact row maximum keeps every exponent <= 0."""
    from tf_operator_b200.ops.attention import attention
    qkv = torch.zeros((1, 128, 3 * 64), device="cuda")
    qkv[..., :128] = 40.0                     # q . k = 64 * 1600 -> logits 12,800
    qkv[..., 128:] = torch.randn((1, 128, 64), device="cuda")
    out = attention(qkv.bfloat16(), 1)
    assert torch.isfinite(out.float()).all()
    want = qkv[..., 128:].bfloat16().float().mean(dim=1, keepdim=True).expand(1, 128, 64)
    torch.testing.assert_close(out.float(), want, rtol=2e-2, atol=2e-2)


def test_attention_rejects_what_the_kernels_do_not_cover():
    from tf_operator_b200.ops.attention import attention
    with pytest.raises(TypeError):
        attention(torch.zeros((1, 128, 192), device="cuda"), 1)
    with pytest.raises(ValueError):
        attention(torch.zeros((1, 100, 192), device="cuda", dtype=torch.bfloat16), 1)
    with pytest.raises(ValueError):
        attention(torch.zeros((1, 128, 96 * 3), device="cuda", dtype=torch.bfloat16), 1)
