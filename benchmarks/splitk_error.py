#!/usr/bin/env python
"""How much does split-K into a bf16 C cost in accuracy? The weight-gradient GEMMs of BERT-large
(dW = dY^T X, K = 8192 tokens) run with 2-4 K-splits whose partial tiles are rounded to bf16 and
added in bf16 by TMA reduce-add (csrc/gemm_tcgen05.cu). This measures, per shape, the error of
splits = 1 (one rounding) and of the heuristic's split count against an fp32 product of the same
bf16 operands, in units of the bf16 spacing at each output value (ulp)."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tf_operator_b200.ops import _lib  # noqa: E402
from tf_operator_b200.ops.gemm import gemm  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--out", default="")
a = ap.parse_args()
torch.backends.cuda.matmul.allow_tf32 = False
lib = _lib.load()
rows = []
for m, n, k in [(1024, 1024, 8192), (1024, 4096, 8192), (4096, 1024, 8192), (3072, 1024, 8192)]:
    torch.manual_seed(m + n)
    dy = (torch.randn(k, m, device="cuda") * 0.02).bfloat16()      # gradient-sized values
    x = torch.randn(k, n, device="cuda").bfloat16()
    ref = dy.float().t() @ x.float()
    row = {"shape": f"{m}x{n}x{k}"}
    for name, s in (("splits_1", -1), ("heuristic", 0), ("splits_4", 4)):
        _lib.check(lib.tfop_gemm_set_splits(s))
        c = gemm(dy, x, a_t=True, b_t=True).float()
        # bf16 spacing at |ref|: 2^(floor(log2|ref|) - 7)
        ulp = torch.exp2(torch.floor(torch.log2(ref.abs().clamp_min(1e-30))) - 7)
        e = ((c - ref).abs() / ulp)
        row[name] = {"max_ulp": round(e.max().item(), 2), "mean_ulp": round(e.mean().item(), 3),
                     "p999_ulp": round(torch.quantile(e.flatten()[:4_000_000], 0.999).item(), 2),
                     "rel_fro": float((c - ref).norm() / ref.norm())}
    _lib.check(lib.tfop_gemm_set_splits(0))
    rows.append(row)
    print(json.dumps(row), flush=True)
if a.out:
    json.dump(rows, open(a.out, "w"), indent=1)
