#!/usr/bin/env python
"""Launch every hand-written kernel that does not already have its own ncu script once, at a
realistic size, on ONE GPU (multi-rank kernels through their emulated single-GPU launch), so a
single `ncu --set full` run yields a capture per kernel. Not a benchmark."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tf_operator_b200.ops import optim as fo  # noqa: E402
from tf_operator_b200.ops import preproc  # noqa: E402
from tf_operator_b200.ops.conv_stem import stem_conv_s2d  # noqa: E402
from tf_operator_b200.ops.gemm import gemm, transpose  # noqa: E402
from tf_operator_b200.ops.hist import build_histogram  # noqa: E402
from tf_operator_b200.ops.pool import max_pool_3x3_s2  # noqa: E402
from tf_operator_b200.ops.ps_embedding import ShardedEmbedding  # noqa: E402
from tf_operator_b200.ops.xent import softmax_cross_entropy  # noqa: E402
from tf_operator_b200.runtime.symm import (ONE_SHOT, TWO_SHOT, SymmetricComm, all_reduce_emulated,  # noqa: E402
                                           gemm_all_reduce_emulated)

dev = torch.device("cuda", 0)
torch.manual_seed(0)

# preprocessing + stem conv (forward and weight gradient) + max-pool
u8 = torch.randint(0, 256, (64, 224, 224, 3), dtype=torch.uint8, device=dev)
x = preproc.stem_space_to_depth(u8)
preproc.u8_to_bf16_normalize(u8)
w4 = preproc.stem_weight_s2d((torch.randn(64, 3, 7, 7, device=dev) * 0.05).bfloat16()).requires_grad_(True)
y = stem_conv_s2d(x, w4)
y.backward(torch.randn_like(y))
p = y.detach().requires_grad_(True)
q = max_pool_3x3_s2(p)
q.backward(torch.randn_like(q))

# optimizers + pack
n = 25_600_000
master, mom, mom2 = torch.randn(n, device=dev), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
grad = torch.randn(n, device=dev).bfloat16()
w16 = torch.empty(n, device=dev, dtype=torch.bfloat16)
hp = torch.tensor([0.1, 0.9, 1e-4, 1.0, 0.9, 0.999, 1e-8, 1.0], device=dev)
fo.sgd_momentum_(master, mom, grad, w16, n, hp)
fo.adamw_(master, mom, mom2, grad, w16, n, hp)
srcs = [torch.randn(1024, 1024, device=dev).bfloat16() for _ in range(16)]
flat = torch.zeros(16 << 20, device=dev, dtype=torch.bfloat16)
fo.multi_pack_bf16(srcs, [i << 20 for i in range(16)], flat)
fo.zero_(flat)

# transposed-operand GEMMs (dgrad / wgrad forms) and the transpose kernel
a = torch.randn(8192, 4096, device=dev).bfloat16()
b = torch.randn(4096, 1024, device=dev).bfloat16()
gemm(a, b, b_t=True)
gemm(a, torch.randn(8192, 1024, device=dev).bfloat16(), a_t=True, b_t=True)
transpose(a)

# cross-entropy
logits = torch.randn(8192, 30528, device=dev).bfloat16().requires_grad_(True)
labels = torch.randint(0, 30528, (8192,), device=dev)
labels[torch.rand(8192, device=dev) > 0.15] = -100
softmax_cross_entropy(logits, labels).backward()

# parameter-server embedding: gather + fused Adagrad push (2 shards on one GPU)
emb = ShardedEmbedding.local(1_000_000, 64, num_shards=2, device=dev, lr=0.1)
ids = torch.randint(0, 1_000_000, (262144,), device=dev)
rows = emb.gather(ids, torch.bfloat16)
emb.push(ids, torch.randn(ids.numel(), 64, device=dev).bfloat16())

# histogram build
bins = torch.randint(0, 256, (1_000_000, 50), dtype=torch.uint8, device=dev)
build_histogram(bins, torch.randn(1_000_000, device=dev), torch.rand(1_000_000, device=dev))

# P2P all-reduce kernels, 4 emulated ranks in one cooperative launch
comms = SymmetricComm.emulated(4, 64 << 20)
for c in comms:
    c.view(0, 16 << 20, torch.bfloat16).normal_()
all_reduce_emulated(comms, 16 << 20, sym_off=0, scale=0.25, algo=TWO_SHOT)
ins = [torch.randn(4096, device=dev).bfloat16() for _ in range(4)]
outs = [torch.zeros(4096, device=dev, dtype=torch.bfloat16) for _ in range(4)]
all_reduce_emulated(comms, 4096, inps=ins, outs=outs, scale=0.25, algo=ONE_SHOT)
for c in comms:
    c.close()

# fused GEMM -> reduce-scatter -> all-gather, 2 emulated ranks
W, M, N, K = 2, 2048, 1024, 2048
ws = SymmetricComm.gemm_allreduce_workspace_bytes(M, N, W)
comms = SymmetricComm.emulated(W, ws + M * N * 2 + 4096)
al = [torch.randn(M, K, device=dev).bfloat16() for _ in range(W)]
bl = [torch.randn(N, K, device=dev).bfloat16() for _ in range(W)]
gemm_all_reduce_emulated(comms, al, bl, ws_off=0, out_off=ws, scale=0.5)
torch.cuda.synchronize()
assert not comms[0].check_error()
print("all kernels launched")
