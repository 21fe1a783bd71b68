"""Fused bias + residual + LayerNorm and bias + GeLU kernels (csrc/layernorm.cu) against fp32
PyTorch references of the same math."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,H", [(515, 1024), (64, 128), (1000, 768), (33, 1536), (8192, 1024),
                                    (100, 520), (7, 256)])
@pytest.mark.parametrize("with_bias,with_res", [(True, True), (False, False), (True, False)])
@pytest.mark.parametrize("warp_rows", [False, True])
def test_add_bias_layernorm_matches_fp32(rows, H, with_bias, with_res, warp_rows):
    from tf_operator_b200.ops import _lib
    from tf_operator_b200.ops.layernorm import add_bias_layernorm
    if warp_rows and H % 256:
        pytest.skip("the warp-per-row backward is already the default for this hidden size")
    _lib.load().tfop_ln_force_warp_rows(1 if warp_rows else 0)
    try:
        _ln_case(add_bias_layernorm, rows, H, with_bias, with_res)
    finally:
        _lib.load().tfop_ln_force_warp_rows(0)


def _ln_case(add_bias_layernorm, rows, H, with_bias, with_res):
    torch.manual_seed(rows + H)
    dev = "cuda"
    h = torch.randn(rows, H, device=dev).bfloat16().requires_grad_(True)
    bias = (torch.randn(H, device=dev) * 0.5).bfloat16().requires_grad_(True) if with_bias else None
    res = torch.randn(rows, H, device=dev).bfloat16().requires_grad_(True) if with_res else None
    gamma = (1 + 0.1 * torch.randn(H, device=dev)).bfloat16().requires_grad_(True)
    beta = (0.1 * torch.randn(H, device=dev)).bfloat16().requires_grad_(True)
    y = add_bias_layernorm(h, bias, res, gamma, beta, 1e-12)
    dy = torch.randn_like(y)
    y.backward(dy)

    # reference: same bf16 rounding of the sum, everything else fp32
    hf, gf, bf = h.detach().float().requires_grad_(True), gamma.detach().float().requires_grad_(True), \
        beta.detach().float().requires_grad_(True)
    s = hf
    biasf = resf = None
    if with_bias:
        biasf = bias.detach().float().requires_grad_(True)
        s = s + biasf
    if with_res:
        resf = res.detach().float().requires_grad_(True)
        s = s + resf
    s = s + (s.detach().bfloat16().float() - s.detach())      # straight-through bf16 rounding
    yr = F.layer_norm(s, (H,), gf, bf, 1e-12)
    yr.backward(dy.float())
    torch.testing.assert_close(y.float(), yr, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(h.grad.float(), hf.grad, rtol=2e-2, atol=2e-2)
    scale = rows ** 0.5
    torch.testing.assert_close(gamma.grad.float(), gf.grad, rtol=2e-2, atol=0.05 * scale)
    torch.testing.assert_close(beta.grad.float(), bf.grad, rtol=2e-2, atol=0.05 * scale)
    if with_bias:
        torch.testing.assert_close(bias.grad.float(), biasf.grad, rtol=2e-2, atol=0.05 * scale)
    if with_res:
        assert torch.equal(res.grad, h.grad)


@pytest.mark.parametrize("rows,H", [(515, 4096), (64, 512), (8192, 4096), (77, 1000)])
def test_bias_gelu_matches_fp32(rows, H):
    from tf_operator_b200.ops.layernorm import bias_gelu
    torch.manual_seed(rows * 3 + H)
    h = torch.randn(rows, H, device="cuda").bfloat16().requires_grad_(True)
    b = torch.randn(H, device="cuda").bfloat16().requires_grad_(True)
    y = bias_gelu(h, b)
    dy = torch.randn_like(y)
    y.backward(dy)
    hf, bfl = h.detach().float().requires_grad_(True), b.detach().float().requires_grad_(True)
    yr = F.gelu(hf + bfl)
    yr.backward(dy.float())
    torch.testing.assert_close(y.float(), yr, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(h.grad.float(), hf.grad, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(b.grad.float(), bfl.grad, rtol=2e-2, atol=0.05 * rows ** 0.5)


@pytest.mark.parametrize("rows,H", [(8192, 3072), (100, 8), (1, 1024), (4097, 520)])
def test_bias_grad_column_sum(rows, H):
    from tf_operator_b200.ops.layernorm import bias_grad
    torch.manual_seed(rows + H)
    dy = torch.randn(rows, H, device="cuda").bfloat16()
    torch.testing.assert_close(bias_grad(dy).float(), dy.float().sum(0), rtol=1e-2,
                               atol=0.02 * rows ** 0.5)


def test_bert_tiny_fused_path_matches_stock_twin():
    """The native graph (tcgen05 GEMMs + fused LN / GeLU kernels) and its all-library twin share
    a state dict and must agree on loss and gradients."""
    from tf_operator_b200.models.bert import BertConfig, BertForPreTraining
    torch.manual_seed(3)
    cn, cs = BertConfig.tiny(), BertConfig.tiny()
    cs.tc_gemm = False
    mn = BertForPreTraining(cn).cuda().bfloat16()
    ms = BertForPreTraining(cs).cuda().bfloat16()
    ms.load_state_dict(mn.state_dict())
    ids = torch.randint(0, cn.vocab_size, (4, 64), device="cuda")
    lab = torch.where(torch.rand(4, 64, device="cuda") < 0.3, ids, torch.full_like(ids, -100))
    nsp = torch.randint(0, 2, (4,), device="cuda")
    ln, ls = mn(ids, lab, nsp), ms(ids, lab, nsp)
    ln.backward()
    ls.backward()
    assert abs(float(ln) - float(ls)) < 0.05
    for (name, pn), (_, ps) in zip(mn.named_parameters(), ms.named_parameters()):
        if pn.grad is None:
            continue
        gn, gs = pn.grad.float(), ps.grad.float()
        denom = gs.abs().max().item() + 1e-6
        assert (gn - gs).abs().max().item() / denom < 0.08, name


@pytest.mark.parametrize("rows,V", [(512, 30528), (64, 1000), (33, 1027), (8, 8)])
def test_softmax_cross_entropy_matches_fp32(rows, V):
    from tf_operator_b200.ops.xent import softmax_cross_entropy
    torch.manual_seed(rows + V)
    logits = (torch.randn(rows, V, device="cuda") * 3).bfloat16().requires_grad_(True)
    labels = torch.randint(0, V, (rows,), device="cuda")
    labels[torch.rand(rows, device="cuda") < 0.6] = -100
    labels[0] = 5 % V      # at least one labelled row
    w = torch.tensor(0.7, device="cuda")
    loss = softmax_cross_entropy(logits, labels, -100)
    (loss * w).backward()
    lf = logits.detach().float().requires_grad_(True)
    ref = F.cross_entropy(lf, labels, ignore_index=-100)
    (ref * w).backward()
    torch.testing.assert_close(loss, ref, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(logits.grad.float(), lf.grad, rtol=2e-2, atol=1e-5)
    assert bool((logits.grad[labels == -100] == 0).all())


def test_softmax_cross_entropy_all_ignored():
    from tf_operator_b200.ops.xent import softmax_cross_entropy
    logits = torch.randn(16, 64, device="cuda").bfloat16().requires_grad_(True)
    labels = torch.full((16,), -100, device="cuda", dtype=torch.int64)
    loss = softmax_cross_entropy(logits, labels, -100)
    loss.backward()
    assert float(loss) == 0.0 and bool((logits.grad == 0).all())
