#!/usr/bin/env python
"""Run the attention kernels a few times on the BERT-large shape (ncu target) and print their
CUDA-event times next to cuDNN SDPA on the same tensors."""
import argparse
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tf_operator_b200.ops.attention import attention, attention_reference  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=16)
ap.add_argument("--seq", type=int, default=512)
ap.add_argument("--heads", type=int, default=16)
ap.add_argument("--iters", type=int, default=5)
a = ap.parse_args()
qkv = torch.randn((a.batch, a.seq, 3 * a.heads * 64), device="cuda").bfloat16().requires_grad_(True)
dout = torch.randn((a.batch, a.seq, a.heads * 64), device="cuda").bfloat16()


def timed(fn):
    fn()
    ts = []
    for _ in range(a.iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return sorted(ts)[len(ts) // 2], r


flops = 4.0 * a.batch * a.heads * a.seq * a.seq * 64
for name, f in (("tcgen05", attention), ("cudnn_sdpa", attention_reference)):
    t_f, out = timed(lambda: f(qkv, a.heads))
    t_b, _ = timed(lambda: torch.autograd.grad(f(qkv, a.heads), qkv, dout))
    print(f"{name}: fwd {t_f:.1f} us ({flops / t_f / 1e6:.0f} TFLOP/s)  fwd+bwd {t_b:.1f} us "
          f"(bwd ~{t_b - t_f:.1f} us, {2.5 * flops / max(t_b - t_f, 1) / 1e6:.0f} TFLOP/s)", flush=True)
# kernel-only: the C entry points on preallocated buffers, 20 launches per event pair
from tf_operator_b200.ops import _lib  # noqa: E402
lib = _lib.load()
b, s_, h = a.batch, a.seq, a.heads * 64
q = qkv.detach()
out = torch.empty((b, s_, h), dtype=torch.bfloat16, device="cuda")
lse = torch.empty((b, a.heads, s_), dtype=torch.float32, device="cuda")
dqkv = torch.empty_like(q)
delta = torch.empty_like(lse)
dq32 = torch.zeros((b * s_, h), dtype=torch.float32, device="cuda")
st = torch.cuda.current_stream().cuda_stream


def many(fn, n=20):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


tf = many(lambda: lib.tfop_attn_fwd(q.data_ptr(), out.data_ptr(), lse.data_ptr(), b, s_, a.heads, 64, st))
tb = many(lambda: lib.tfop_attn_bwd(q.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(),
                                    dqkv.data_ptr(), delta.data_ptr(), dq32.data_ptr(), b, s_, a.heads, 64, st))
print(f"kernel-only: fwd {tf:.1f} us ({flops / tf / 1e6:.0f} TFLOP/s)  bwd (memset + delta + main + convert) "
      f"{tb:.1f} us ({2.5 * flops / tb / 1e6:.0f} TFLOP/s)")
from torch.profiler import ProfilerActivity, profile  # noqa: E402
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(10):
        lib.tfop_attn_bwd(q.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(),
                          dqkv.data_ptr(), delta.data_ptr(), dq32.data_ptr(), b, s_, a.heads, 64, st)
    torch.cuda.synchronize()
for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:6]:
    print(f"  {e.device_time_total / e.count:8.1f} us x{e.count}  {e.key[:90]}")
print("ok")
