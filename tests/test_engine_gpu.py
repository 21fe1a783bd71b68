"""End-to-end replica runtime on one GPU: engine step (eager and CUDA-graphed)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _make(graph, fused=True):
    from tf_operator_b200.models.resnet import ResNet
    from tf_operator_b200.parallel.engine import DataParallelEngine
    torch.manual_seed(0)
    model = ResNet(18, num_classes=10, fused=fused, width=64)
    return DataParallelEngine(model, batch_size=16, image_size=64, lr=0.05, graph=graph)


@pytest.mark.parametrize("graph", [False, True])
def test_engine_loss_decreases(graph):
    eng = _make(graph)
    x = torch.randint(0, 256, (16, 64, 64, 3), dtype=torch.uint8).pin_memory()
    y = torch.randint(0, 10, (16,), dtype=torch.int64).pin_memory()
    losses = [eng.step(x, y).result() for _ in range(30)]
    assert all(l == l for l in losses), "NaN loss"
    assert losses[-1] < losses[0] * 0.7, losses
    assert eng.kernel_launches_per_step > 20  # fused BN + pack + SGD kernels actually ran


def test_engine_matches_unfused_reference_first_step():
    e1, e2 = _make(False, fused=True), _make(False, fused=False)
    x = torch.randint(0, 256, (16, 64, 64, 3), dtype=torch.uint8).pin_memory()
    y = torch.randint(0, 10, (16,), dtype=torch.int64).pin_memory()
    l1 = [e1.step(x, y).result() for _ in range(3)]
    l2 = [e2.step(x, y).result() for _ in range(3)]
    assert abs(l1[0] - l2[0]) < 0.05, (l1, l2)
    assert abs(l1[2] - l2[2]) < 0.3, (l1, l2)


def test_graft_smoke():
    import __graft_entry__ as g
    g.smoke()


@pytest.mark.parametrize("graph", [False, True])
def test_checkpoint_resume_reproduces_the_uninterrupted_run(tmp_path, graph):
    """Train 6 steps, save, train 4 more; a fresh engine (other seed) that loads the checkpoint
    and trains the same 4 steps reaches the same losses and weights - also through the CUDA
    graph, whose captured kernels read the reloaded flat buffers in place."""
    x = torch.randint(0, 256, (16, 64, 64, 3), dtype=torch.uint8).pin_memory()
    y = torch.randint(0, 10, (16,), dtype=torch.int64).pin_memory()
    a = _make(graph)
    for _ in range(6):
        a.step(x, y).result()
    path = str(tmp_path / "ckpt.pt")
    a.save_checkpoint(path)
    tail_a = [a.step(x, y).result() for _ in range(4)]

    torch.manual_seed(123)
    from tf_operator_b200.models.resnet import ResNet
    from tf_operator_b200.parallel.engine import DataParallelEngine
    b = DataParallelEngine(ResNet(18, num_classes=10, fused=True, width=64), batch_size=16,
                           image_size=64, lr=0.05, graph=graph)
    b.step(x, y).result()                      # captures the graph before the load, on purpose
    assert b.load_checkpoint(path) == 6
    tail_b = [b.step(x, y).result() for _ in range(4)]
    assert tail_b == pytest.approx(tail_a, rel=2e-2, abs=2e-2), (tail_a, tail_b)
    torch.testing.assert_close(b.flat.master, a.flat.master, rtol=1e-2, atol=1e-2)
    sd = torch.load(path, weights_only=False)
    assert sd["step"] == 6 and sd["optimizer"] == "sgd" and "fc.weight" in sd["model"]
    assert any("running_mean" in k for k in sd["buffers"])


def test_linear_weight_gradients_land_in_the_bucket_without_a_pack_copy():
    """TcLinear's wgrad GEMM stores dW straight into the flat gradient bucket (GradTarget);
    autograd adopts that view as .grad, the pack step skips it, and training matches a twin
    engine whose parameters have no direct target (everything packed)."""
    import torch.nn as nn
    from tf_operator_b200.ops.gemm import TcLinear
    from tf_operator_b200.parallel.engine import FlatDataParallel

    def make():
        torch.manual_seed(5)
        return nn.Sequential(TcLinear(256, 512, act="gelu"), TcLinear(512, 256), TcLinear(256, 64)).cuda().bfloat16()
    x = torch.randn(128, 256, device="cuda").bfloat16()
    t = torch.randn(128, 64, device="cuda")
    ma, mb = make(), make()
    ea = FlatDataParallel(ma, optimizer="adamw", lr=1e-3, weight_decay=0.0)
    eb = FlatDataParallel(mb, optimizer="adamw", lr=1e-3, weight_decay=0.0)
    for p in mb.parameters():
        del p._tfop_grad_target                     # twin: classic path, every gradient packed
    la = [float(ea.train_step(lambda: (ma(x).float() - t).pow(2).mean())) for _ in range(5)]
    in_place = [s.name for s in ea.flat.slots if ea.flat.grad_in_place(s)]
    assert sorted(in_place) == ["0.weight", "1.weight", "2.weight"], in_place
    lb = [float(eb.train_step(lambda: (mb(x).float() - t).pow(2).mean())) for _ in range(5)]
    assert not any(eb.flat.grad_in_place(s) for s in eb.flat.slots)
    assert la == pytest.approx(lb, rel=1e-3, abs=1e-4), (la, lb)
    torch.testing.assert_close(ea.flat.master, eb.flat.master, rtol=1e-3, atol=1e-5)
    assert la[-1] < la[0]
    # a weight used twice in one step: the second use must not overwrite the first in the bucket
    torch.manual_seed(6)
    lin = TcLinear(64, 64).cuda().bfloat16()
    ref = TcLinear(64, 64).cuda().bfloat16()
    ref.load_state_dict(lin.state_dict())
    ec = FlatDataParallel(lin, optimizer="sgd", lr=0.0, momentum=0.0, weight_decay=0.0)
    z = torch.randn(32, 64, device="cuda").bfloat16()
    ec.train_step(lambda: lin(lin(z)).float().pow(2).mean())
    ref(ref(z)).float().pow(2).mean().backward()
    got = ec.flat.view_of(ec.flat.grads, [s for s in ec.flat.slots if s.name == "weight"][0])
    torch.testing.assert_close(got.float(), ref.weight.grad.float(), rtol=2e-2, atol=2e-3)
