#!/usr/bin/env python
"""One tcgen05 GEMM launch sequence for ncu captures (8192^3 by default)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tf_operator_b200.ops.gemm import gemm_tn  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
tile = int(sys.argv[2]) if len(sys.argv) > 2 else 256
a = torch.randn(m, m, device="cuda").bfloat16()
b = torch.randn(m, m, device="cuda").bfloat16()
for _ in range(4):
    c = gemm_tn(a, b, tile_n=tile)
torch.cuda.synchronize()
print("ok", float(c.float().abs().mean()))
