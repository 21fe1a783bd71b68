"""Wide-and-deep recommender for the TFJob PS/Worker config (BASELINE.json config 4).

Sparse side: one sharded embedding table on the PS replicas (ops/ps_embedding.py): the deep
part looks up ``num_fields`` ids per sample (dim ``embed_dim``), the wide part is a linear model
over the same ids (a second table of dim 4, column 0 used). Dense side: an MLP tower whose
GEMMs run on the hand-written tcgen05 kernel (ops/gemm.py), replicated on the workers and
all-reduced among them.

The reference has no such model; its PS example is the dense MLP of
examples/v1/dist-mnist/dist_mnist.py:166-194. Shapes come from BASELINE.json.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch
import torch.nn as nn

from ..ops.gemm import TcLinear


class DeepTower(nn.Module):
    def __init__(self, in_features: int, hidden: Sequence[int] = (1024, 512, 256)):
        super().__init__()
        layers, d = [], in_features
        for h in hidden:
            layers.append(TcLinear(d, h, act="relu"))
            d = h
        self.layers = nn.ModuleList(layers)
        self.out = TcLinear(d, 8)      # logit in column 0 (N padded to 8 for 16-byte rows)

    def forward(self, x):
        for l in self.layers:
            x = l(x)
        return self.out(x)[..., 0]


class WideAndDeep(nn.Module):
    def __init__(self, deep_table, wide_table, num_fields: int = 26, embed_dim: int = 64,
                 hidden: Sequence[int] = (1024, 512, 256), dense_features: int = 16,
                 compute_dtype: torch.dtype = torch.bfloat16):
        super().__init__()
        self.deep_table, self.wide_table = deep_table, wide_table
        self.num_fields, self.embed_dim = num_fields, embed_dim
        self.dense_features = dense_features
        in_f = num_fields * embed_dim + dense_features
        in_f_padded = (in_f + 7) // 8 * 8
        self.pad = in_f_padded - in_f
        self.tower = DeepTower(in_f_padded, hidden)
        self.wide_bias = nn.Parameter(torch.zeros(1))
        self.compute_dtype = compute_dtype

    def forward(self, ids: torch.Tensor, dense: torch.Tensor) -> torch.Tensor:
        """ids: [B, num_fields] int64 in [0, rows); dense: [B, dense_features]. Returns logits [B]."""
        b = ids.shape[0]
        emb = self.deep_table.lookup(ids, self.compute_dtype)            # [B, F, D]
        x = torch.cat([emb.reshape(b, -1), dense.to(self.compute_dtype)], dim=1)
        if self.pad:
            x = torch.nn.functional.pad(x, (0, self.pad))
        deep = self.tower(x).float()
        wide = self.wide_table.lookup(ids, torch.float32)[..., 0].sum(dim=1)
        return deep + wide + self.wide_bias.float()
