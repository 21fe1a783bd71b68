#!/usr/bin/env python
"""Probe: cost of the ResNet stem conv (7x7/2, 64 out) for input channels 3 / 4 / 8 (NHWC bf16),
and torch's max-pool forward/backward next to ours."""
import os
import sys
import json

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
torch.backends.cudnn.benchmark = True


def timed(fn, iters=10, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


res = {}
for c in (3, 4, 8):
    x = torch.randn(256, c, 224, 224, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    w = torch.randn(64, c, 7, 7, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    y = F.conv2d(x, w, stride=2, padding=3)
    dy = torch.randn_like(y)
    fwd = timed(lambda: F.conv2d(x, w, stride=2, padding=3))
    bwd = timed(lambda: torch.autograd.grad(F.conv2d(x, w, stride=2, padding=3), w, dy))
    res[f"stem_c{c}"] = {"fwd_us": round(fwd, 1), "fwd_plus_wgrad_us": round(bwd, 1)}

x = torch.randn(256, 64, 112, 112, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
y = F.max_pool2d(x, 3, 2, 1)
dy = torch.randn_like(y)
res["torch_maxpool"] = {"fwd_us": round(timed(lambda: F.max_pool2d(x, 3, 2, 1)), 1),
                        "bwd_us": round(timed(lambda: torch.autograd.grad(y, x, dy, retain_graph=True)), 1)}
try:
    from tf_operator_b200.ops.pool import max_pool_3x3_s2
    y2 = max_pool_3x3_s2(x)
    res["tfop_maxpool"] = {"fwd_us": round(timed(lambda: max_pool_3x3_s2(x)), 1),
                           "bwd_us": round(timed(lambda: torch.autograd.grad(y2, x, dy, retain_graph=True)), 1),
                           "fwd_equal": bool(torch.equal(y2, y)),
                           "bwd_max_err": float((torch.autograd.grad(y2, x, dy, retain_graph=True)[0].float()
                                                 - torch.autograd.grad(y, x, dy, retain_graph=True)[0].float()).abs().max())}
except Exception as e:  # noqa: BLE001
    res["tfop_maxpool"] = {"error": repr(e)}
print(json.dumps(res))
