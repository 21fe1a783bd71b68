"""CPU paths of the model code and the ops' non-CUDA fallbacks: the same modules that run the
hand-written kernels on a B200 must still compute the same function on CPU (the control-plane
e2e jobs run them there)."""
import torch
import torch.nn.functional as F

from tf_operator_b200.models.bert import BertConfig, BertForPreTraining
from tf_operator_b200.models.resnet import ResNet
from tf_operator_b200.ops.gemm import linear
from tf_operator_b200.ops.layernorm import add_bias_layernorm, bias_gelu, bias_grad
from tf_operator_b200.ops.preproc import stem_space_to_depth, stem_weight_s2d
from tf_operator_b200.ops.xent import softmax_cross_entropy


def test_bert_native_graph_equals_library_twin_on_cpu():
    torch.manual_seed(0)
    cn, cs = BertConfig.tiny(), BertConfig.tiny()
    cs.tc_gemm = False
    mn, ms = BertForPreTraining(cn), BertForPreTraining(cs)
    ms.load_state_dict(mn.state_dict())
    ids = torch.randint(0, cn.vocab_size, (2, 24))
    lab = torch.where(torch.rand(2, 24) < 0.3, ids, torch.full_like(ids, -100))
    nsp = torch.randint(0, 2, (2,))
    ln, ls = mn(ids, lab, nsp), ms(ids, lab, nsp)
    ln.backward()
    ls.backward()
    torch.testing.assert_close(ln, ls)
    for (name, pn), (_, ps) in zip(mn.named_parameters(), ms.named_parameters()):
        if pn.grad is not None:
            torch.testing.assert_close(pn.grad, ps.grad, rtol=1e-4, atol=1e-6, msg=name)
    assert set(mn.state_dict()) == set(ms.state_dict())


def test_op_fallbacks_match_torch_composition():
    torch.manual_seed(1)
    h, res = torch.randn(5, 7, 32), torch.randn(5, 7, 32)
    bias, g, b = torch.randn(32), torch.rand(32) + 0.5, torch.randn(32)
    torch.testing.assert_close(add_bias_layernorm(h, bias, res, g, b, 1e-5),
                               F.layer_norm(h + bias + res, (32,), g, b, 1e-5))
    torch.testing.assert_close(add_bias_layernorm(h, None, None, g, b, 1e-5),
                               F.layer_norm(h, (32,), g, b, 1e-5))
    torch.testing.assert_close(bias_gelu(h, bias), F.gelu(h + bias))
    torch.testing.assert_close(bias_grad(h.reshape(-1, 32)), h.reshape(-1, 32).sum(0))
    w = torch.randn(16, 32)
    torch.testing.assert_close(linear(h, w, None, "gelu"), F.gelu(F.linear(h, w)))
    torch.testing.assert_close(linear(h, w, torch.ones(16), "relu"), F.relu(F.linear(h, w, torch.ones(16))))
    logits = torch.randn(9, 11)
    labels = torch.tensor([1, -100, 3, 4, -100, 0, 10, 2, -100])
    torch.testing.assert_close(softmax_cross_entropy(logits, labels, -100),
                               F.cross_entropy(logits, labels, ignore_index=-100))


def test_stem_space_to_depth_is_the_7x7_stride2_convolution():
    """The 4x4/1 convolution over the 16-channel space-to-depth input, with the re-laid-out
    weight, equals conv7x7/2/pad3 of the normalised image - same parameters, same gradient."""
    torch.manual_seed(2)
    u8 = torch.randint(0, 256, (2, 32, 32, 3), dtype=torch.uint8)
    w7 = (torch.randn(8, 3, 7, 7) * 0.1).requires_grad_(True)
    x = stem_space_to_depth(u8).float()                          # logical [N, 16, 19, 19]
    y = F.conv2d(x, stem_weight_s2d(w7))
    mean = torch.tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1)
    std = torch.tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1)
    img = ((u8.permute(0, 3, 1, 2).float() / 255.0 - mean) / std).bfloat16().float()
    w7r = w7.detach().clone().requires_grad_(True)
    ref = F.conv2d(img, w7r, stride=2, padding=3)
    assert y.shape == ref.shape == (2, 8, 16, 16)
    torch.testing.assert_close(y, ref, rtol=1e-3, atol=1e-3)
    y.sum().backward()
    ref.sum().backward()
    torch.testing.assert_close(w7.grad, w7r.grad, rtol=1e-3, atol=1e-2)


def test_resnet_accepts_plain_and_space_to_depth_inputs_on_cpu():
    torch.manual_seed(3)
    m = ResNet(18, num_classes=10, fused=True, width=64).eval()
    u8 = torch.randint(0, 256, (2, 64, 64, 3), dtype=torch.uint8)
    mean = torch.tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1)
    std = torch.tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1)
    img = ((u8.permute(0, 3, 1, 2).float() / 255.0 - mean) / std).bfloat16().float()
    with torch.no_grad():
        a = m(img)
        b = m(stem_space_to_depth(u8).float())
    torch.testing.assert_close(a, b, rtol=1e-3, atol=1e-3)
    # the opt-in fused stem tail falls back to bn + pool off-GPU
    m2 = ResNet(18, num_classes=10, fused=True, width=64, fused_stem_tail=True)
    m2.load_state_dict(m.state_dict())
    m2.eval()
    with torch.no_grad():
        torch.testing.assert_close(m2(img), a)
