#!/usr/bin/env python
"""Run one convolution layer a few times (ncu target): --what fwd|dgrad|wgrad."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tf_operator_b200.ops import conv as C  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--shape", default="56,64,256,1,1", help="H,Cin,Cout,k,stride")
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--what", default="fwd")
ap.add_argument("--iters", type=int, default=4)
ap.add_argument("--no-stats", action="store_true")
a = ap.parse_args()
h, cin, cout, k, s = map(int, a.shape.split(","))
CL = torch.channels_last
x = torch.randn((a.batch, cin, h, h), device="cuda").bfloat16().contiguous(memory_format=CL)
w = (torch.randn((cout, cin, k, k), device="cuda") / (cin * k * k) ** 0.5).bfloat16().contiguous(memory_format=CL)
oh, ow = C._out_hw(h, h, k, s)
dy = torch.randn((a.batch, cout, oh, ow), device="cuda").bfloat16().contiguous(memory_format=CL)
stats = None if a.no_stats else torch.zeros(2, cout, dtype=torch.float32, device="cuda")
for _ in range(a.iters):
    if a.what == "fwd":
        C.conv_fwd(x, w, s, stats)
    elif a.what == "dgrad":
        C.conv_dgrad(dy, w, x.shape, s)
    else:
        C.conv_wgrad(x, dy, w, s)
torch.cuda.synchronize()
print("ok")
