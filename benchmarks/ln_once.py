#!/usr/bin/env python
"""Launch each fused transformer elementwise kernel once at BERT-large shapes (for ncu) and
print device-timed microseconds (for the roofline table)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tf_operator_b200.ops.layernorm import add_bias_layernorm, bias_gelu, bias_grad  # noqa: E402
from tf_operator_b200.ops.xent import softmax_cross_entropy  # noqa: E402


def timed(fn, iters=10):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    fn()
    tot = 0.0
    for _ in range(iters):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters * 1e3


def main():
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    dev = "cuda"
    rows, H, I, V = 8192, 1024, 4096, 30528
    torch.manual_seed(0)
    h = torch.randn(rows, H, device=dev).bfloat16().requires_grad_(True)
    res = torch.randn(rows, H, device=dev).bfloat16().requires_grad_(True)
    bias = torch.randn(H, device=dev).bfloat16().requires_grad_(True)
    g = torch.ones(H, device=dev).bfloat16().requires_grad_(True)
    b = torch.zeros(H, device=dev).bfloat16().requires_grad_(True)
    out = {}
    MB = 1 << 20
    out["ln_fwd_us"] = timed(lambda: add_bias_layernorm(h, bias, res, g, b, 1e-12), iters)
    y = add_bias_layernorm(h, bias, res, g, b, 1e-12)
    dy = torch.randn_like(y)
    out["ln_bwd_us"] = timed(lambda: torch.autograd.grad(y, (h, bias, res, g, b), dy, retain_graph=True), iters)
    # the backward kernel alone (no allocations, no clears), two and four rows per thread and trip
    from tf_operator_b200.ops import _lib
    lib = _lib.load()
    s_in = (h + bias + res).detach()
    mean = s_in.float().mean(1)
    rstd = (s_in.float().var(1, unbiased=False) + 1e-12).rsqrt()
    dx = torch.empty_like(s_in)
    dg, db_, dbi = (torch.zeros(H, device=dev) for _ in range(3))
    for r in (2, 4):
        lib.tfop_ln_set_bwd_rows(r)
        out[f"ln_bwd_kernel_only_rows{r}_us"] = timed(lambda: lib.tfop_ln_bwd(
            dy.data_ptr(), s_in.data_ptr(), mean.data_ptr(), rstd.data_ptr(), g.data_ptr(), dx.data_ptr(),
            dg.data_ptr(), db_.data_ptr(), dbi.data_ptr(), rows, H, _lib.stream_handle()), iters)
    lib.tfop_ln_set_bwd_rows(2)
    x = torch.randn(rows, I, device=dev).bfloat16().requires_grad_(True)
    bi = torch.randn(I, device=dev).bfloat16().requires_grad_(True)
    out["bias_gelu_fwd_us"] = timed(lambda: bias_gelu(x, bi), iters)
    z = bias_gelu(x, bi)
    dz = torch.randn_like(z)
    out["bias_gelu_bwd_us"] = timed(lambda: torch.autograd.grad(z, (x, bi), dz, retain_graph=True), iters)
    d3 = torch.randn(rows, 3 * H, device=dev).bfloat16()
    out["bias_grad_3072_us"] = timed(lambda: bias_grad(d3), iters)
    logits = torch.randn(rows, V, device=dev).bfloat16().requires_grad_(True)
    labels = torch.randint(0, V, (rows,), device=dev)
    labels[torch.rand(rows, device=dev) > 0.15] = -100
    out["xent_fwd_us"] = timed(lambda: softmax_cross_entropy(logits, labels), iters)
    loss = softmax_cross_entropy(logits, labels)
    out["xent_bwd_us"] = timed(lambda: torch.autograd.grad(loss, logits, retain_graph=True), iters)
    # algorithmic bytes (bf16 tensors in MiB): the roofline numerators
    out["bytes_MB"] = {"ln_fwd": rows * H * 2 * 4 / MB, "ln_bwd": rows * H * 2 * 3 / MB,
                       "bias_gelu_fwd": rows * I * 2 * 2 / MB, "bias_gelu_bwd": rows * I * 2 * 3 / MB,
                       "bias_grad_3072": rows * 3 * H * 2 / MB,
                       "xent_bwd": rows * V * 2 * 1.15 / MB}
    print(json.dumps({k: (round(v, 1) if isinstance(v, float) else v) for k, v in out.items()}))


if __name__ == "__main__":
    main()
