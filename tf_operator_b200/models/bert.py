"""BERT encoder (base / large) for the TFJob Chief+Workers data-parallel pre-training config
(BASELINE.json config 3). Every projection / MLP / decoder GEMM - forward, dgrad and wgrad - runs
on the hand-written tcgen05 kernel (ops/gemm); bias + residual + LayerNorm, bias + GeLU and the
masked-LM cross-entropy are fused hand-written kernels (ops/layernorm, ops/xent); attention uses
torch's fused SDPA (a library op, like the convolutions in the ResNet path). ``tc_gemm=False``
builds the all-library twin (same parameters) that serves as the measured baseline.

The reference has no BERT; the config is defined by BASELINE.json (bf16, seq 512, bucketed
fused all-reduce).
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops.attention import attention
from ..ops.gemm import TcLinear, linear
from ..ops.layernorm import add_bias_layernorm, bias_gelu
from ..ops.xent import softmax_cross_entropy


@dataclass
class BertConfig:
    vocab_size: int = 30528          # 30522 padded to a multiple of 64
    hidden: int = 1024
    layers: int = 24
    heads: int = 16
    intermediate: int = 4096
    max_position: int = 512
    type_vocab: int = 2
    dropout: float = 0.0
    eps: float = 1e-12
    tc_gemm: bool = True      # False: stock nn.Linear (cuBLAS) - the measured baseline graph

    @classmethod
    def large(cls):
        return cls()

    @classmethod
    def base(cls):
        return cls(hidden=768, layers=12, heads=12, intermediate=3072)

    @classmethod
    def tiny(cls):
        return cls(vocab_size=1024, hidden=128, layers=2, heads=2, intermediate=512, max_position=128)


class _StockLinear(nn.Linear):
    """nn.Linear + activation from torch ops (library GEMM): the baseline twin of TcLinear."""

    def __init__(self, i, o, act: str = "none"):
        super().__init__(i, o)
        self.act = act

    def forward(self, x):
        y = F.linear(x, self.weight, self.bias)
        return F.gelu(y) if self.act == "gelu" else y


def _lin(c: "BertConfig", i, o, act="none"):
    return TcLinear(i, o, act=act) if c.tc_gemm else _StockLinear(i, o, act)


class BertLayer(nn.Module):
    def __init__(self, c: BertConfig):
        super().__init__()
        self.heads = c.heads
        self.fused = c.tc_gemm
        self.qkv = _lin(c, c.hidden, 3 * c.hidden)
        self.proj = _lin(c, c.hidden, c.hidden)
        self.ln1 = nn.LayerNorm(c.hidden, eps=c.eps)
        self.fc1 = _lin(c, c.hidden, c.intermediate, "gelu")
        self.fc2 = _lin(c, c.intermediate, c.hidden)
        self.ln2 = nn.LayerNorm(c.hidden, eps=c.eps)

    def forward(self, x, mask=None):
        b, s, h = x.shape
        if self.fused and mask is None and x.is_cuda and h // self.heads == 64 and s % 128 == 0:
            # tcgen05 attention straight on the packed projection output: no unbind / transpose
            # copies forward, no cat backward (ops/attention.py, csrc/attention.cu)
            a = attention(self.qkv(x), self.heads)
        else:
            # library path: the stock twin, masks, other head sizes, CPU
            # unbind (not permute + index): its backward is one stack copy instead of three
            # zero-filled [3, ...] tensors that are then added
            q, k, v = self.qkv(x).view(b, s, 3, self.heads, h // self.heads).unbind(2)
            a = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2),
                                               v.transpose(1, 2), attn_mask=mask)
            a = a.transpose(1, 2).reshape(b, s, h)
        if not self.fused:
            x = self.ln1(x + self.proj(a))
            return self.ln2(x + self.fc2(self.fc1(x)))
        # tcgen05 GEMMs without epilogue; bias + residual + LayerNorm and bias + GeLU each run
        # as one fused pass (ops/layernorm.py), whose backward also yields the bias gradients
        x = add_bias_layernorm(linear(a, self.proj.weight), self.proj.bias, x,
                               self.ln1.weight, self.ln1.bias, self.ln1.eps)
        m = bias_gelu(linear(x, self.fc1.weight), self.fc1.bias)
        return add_bias_layernorm(linear(m, self.fc2.weight), self.fc2.bias, x,
                                  self.ln2.weight, self.ln2.bias, self.ln2.eps)


class BertForPreTraining(nn.Module):
    """Masked-LM + next-sentence heads over the encoder; ``forward`` returns the summed loss."""

    def __init__(self, c: BertConfig):
        super().__init__()
        self.cfg = c
        self.tok = nn.Embedding(c.vocab_size, c.hidden)
        self.pos = nn.Embedding(c.max_position, c.hidden)
        self.typ = nn.Embedding(c.type_vocab, c.hidden)
        self.ln = nn.LayerNorm(c.hidden, eps=c.eps)
        self.layers = nn.ModuleList(BertLayer(c) for _ in range(c.layers))
        self.mlm_dense = _lin(c, c.hidden, c.hidden, "gelu")
        self.mlm_ln = nn.LayerNorm(c.hidden, eps=c.eps)
        self.mlm_bias = nn.Parameter(torch.zeros(c.vocab_size))
        self.pooler = _lin(c, c.hidden, c.hidden)
        self.nsp = _lin(c, c.hidden, 8)       # 2 classes, padded to 8 columns
        self.apply(self._init)

    @staticmethod
    def _init(m):
        if isinstance(m, (nn.Linear, nn.Embedding)):
            nn.init.normal_(m.weight, std=0.02)
        if isinstance(m, nn.Linear) and m.bias is not None:
            nn.init.zeros_(m.bias)

    def encode(self, ids, token_type=None):
        b, s = ids.shape
        pos = torch.arange(s, device=ids.device)
        x = self.tok(ids) + self.pos(pos)[None] + self.typ(
            token_type if token_type is not None else torch.zeros_like(ids))
        if self.cfg.tc_gemm:
            x = add_bias_layernorm(x, None, None, self.ln.weight, self.ln.bias, self.ln.eps)
        else:
            x = self.ln(x)
        for l in self.layers:
            x = l(x)
        return x

    def forward(self, ids, mlm_labels, nsp_labels, token_type=None):
        x = self.encode(ids, token_type)
        if self.cfg.tc_gemm:
            h = bias_gelu(linear(x, self.mlm_dense.weight), self.mlm_dense.bias)
            h = add_bias_layernorm(h, None, None, self.mlm_ln.weight, self.mlm_ln.bias,
                                   self.mlm_ln.eps)
            logits = linear(h, self.tok.weight, self.mlm_bias)  # tied decoder, tcgen05 GEMM
        else:
            h = self.mlm_ln(self.mlm_dense(x))
            logits = F.linear(h, self.tok.weight, self.mlm_bias)
        if self.cfg.tc_gemm:
            # fused kernel: reads only the labelled rows, writes bf16 dlogits (ops/xent.py)
            mlm = softmax_cross_entropy(logits.view(-1, logits.shape[-1]), mlm_labels.view(-1), -100)
        else:
            mlm = F.cross_entropy(logits.float().view(-1, logits.shape[-1]), mlm_labels.view(-1),
                                  ignore_index=-100)
        pooled = torch.tanh(self.pooler(x[:, 0]))
        nsp = F.cross_entropy(self.nsp(pooled)[..., :2].float(), nsp_labels)
        return mlm + nsp
