#!/usr/bin/env python
"""Band-height sweep of the tcgen05 GEMM tile order (DRAM re-reads vs L2 residency)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tf_operator_b200.ops import _lib  # noqa: E402
from tf_operator_b200.ops.gemm import gemm_tn  # noqa: E402
from kernel_bench import timed  # noqa: E402

lib = _lib.load()
res = []
for (m, n, k) in [(8192, 8192, 8192), (4096, 4096, 4096), (16384, 4096, 1024), (8192, 4096, 1024),
                  (8192, 1024, 4096), (8192, 30528, 1024), (30528, 1024, 8192)]:
    a = torch.randn(m, k, device="cuda").bfloat16()
    b = torch.randn(n, k, device="cuda").bfloat16()
    row = {"m": m, "n": n, "k": k}
    for g in (2, 4, 8, 16, 32):
        lib.tfop_gemm_set_group_m(g)
        us = timed(lambda: gemm_tn(a, b, tile_n=256))
        row[f"g{g}"] = round(2 * m * n * k / us / 1e6, 1)
    lib.tfop_gemm_set_group_m(0)
    res.append(row)
    print(json.dumps(row), flush=True)
