"""ResNet family (18/34/50/101/152) for the PyTorchJob ResNet-50 DDP config
(BASELINE.json config 2). Every convolution runs on the hand-written tcgen05 implicit-GEMM
kernels (ops/conv.py for the 1x1 / 3x3 layers, ops/conv_stem.py for the stem); their epilogues
also reduce the BatchNorm statistics, and their data gradients are added into the shortcut
branch's gradient inside the kernel. Every BatchNorm+ReLU(+residual add) goes through the fused
NHWC kernels in ops/bn.py. ``fused=False`` builds the same network from stock torch modules
(cuDNN convolutions) and is the "reference mode" compute graph for the measured NCCL baseline
(BASELINE.md section 3). The reference itself has no model code; its CNN example is
examples/v1/distribution_strategy/keras-API/multi_worker_strategy-with-keras.py:42-60.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops.bn import FusedBatchNorm2d, bn_relu_maxpool
from ..ops.pool import MaxPool3x3s2
from ..ops.conv import TcConv2d, conv2d_pair, conv2d_skip
from ..ops.conv_stem import stem_conv_s2d
from ..ops.gemm import TcLinear
from ..ops.preproc import stem_weight_s2d


class _StockBNAct(nn.Module):
    def __init__(self, c, relu=True, zero_init=False):
        super().__init__()
        self.bn = nn.BatchNorm2d(c)
        if zero_init:
            nn.init.zeros_(self.bn.weight)
        self.relu = relu

    def forward(self, x, residual=None):
        y = self.bn(x)
        if residual is not None:
            y = y + residual
        return F.relu(y, inplace=True) if self.relu else y


def _bn(c, relu=True, fused=True, zero_init=False):
    return FusedBatchNorm2d(c, relu=relu, zero_init=zero_init) if fused \
        else _StockBNAct(c, relu, zero_init)


def _conv(cin, cout, k, stride=1, fused=True):
    # the tcgen05 kernels tile channels in 64s; other widths (toy models of the CPU tests) are
    # built from the stock module, decided here at construction, never per call
    if fused and cin % 64 == 0 and cout % 64 == 0:
        return TcConv2d(cin, cout, k, stride)
    return nn.Conv2d(cin, cout, k, stride=stride, padding=k // 2, bias=False)


class _Block(nn.Module):
    """Shared forward of the residual blocks. On the native path the first convolution and the
    shortcut are issued together (conv2d_skip / conv2d_pair), so their data gradients meet inside
    the dgrad kernel, and each convolution hands its BatchNorm the statistics it reduced."""

    def _native(self):
        return isinstance(self.conv1, TcConv2d) and (
            self.down is None or isinstance(self.down[0], TcConv2d))

    def _head(self, x):
        """(bn1(conv1(x)), shortcut)"""
        if not self._native():
            idt = x if self.down is None else self.down[1](self.down[0](x))
            return self.bn1(self.conv1(x)), idt
        tr = self.training
        st1 = self.conv1.sums(x) if tr else None
        if self.down is None:
            y, idt = conv2d_skip(x, self.conv1.weight, self.conv1.stride, st1)
        else:
            dconv, dbn = self.down
            std = dconv.sums(x) if tr else None
            y, d = conv2d_pair(x, self.conv1.weight, self.conv1.stride, st1,
                               dconv.weight, dconv.stride, std)
            idt = dbn(d, sums=std)
        return self.bn1(y, sums=st1), idt

    def _conv_bn(self, conv, bn, x, residual=None):
        if isinstance(conv, TcConv2d):
            st = conv.sums(x) if self.training else None
            return bn(conv(x, st), residual, sums=st)
        return bn(conv(x), residual)


class Bottleneck(_Block):
    expansion = 4

    def __init__(self, cin, width, stride, fused, zero_init_residual):
        super().__init__()
        cout = width * self.expansion
        self.conv1, self.bn1 = _conv(cin, width, 1, 1, fused), _bn(width, True, fused)
        self.conv2, self.bn2 = _conv(width, width, 3, stride, fused), _bn(width, True, fused)
        self.conv3 = _conv(width, cout, 1, 1, fused)
        self.bn3 = _bn(cout, True, fused, zero_init=zero_init_residual)
        self.down = None
        if stride != 1 or cin != cout:
            self.down = nn.ModuleList([_conv(cin, cout, 1, stride, fused), _bn(cout, False, fused)])

    def forward(self, x):
        y, idt = self._head(x)
        y = self._conv_bn(self.conv2, self.bn2, y)
        return self._conv_bn(self.conv3, self.bn3, y, idt)


class BasicBlock(_Block):
    expansion = 1

    def __init__(self, cin, width, stride, fused, zero_init_residual):
        super().__init__()
        self.conv1, self.bn1 = _conv(cin, width, 3, stride, fused), _bn(width, True, fused)
        self.conv2 = _conv(width, width, 3, 1, fused)
        self.bn2 = _bn(width, True, fused, zero_init=zero_init_residual)
        self.down = None
        if stride != 1 or cin != width:
            self.down = nn.ModuleList([_conv(cin, width, 1, stride, fused), _bn(width, False, fused)])

    def forward(self, x):
        y, idt = self._head(x)
        return self._conv_bn(self.conv2, self.bn2, y, idt)


_CFG = {
    18: (BasicBlock, (2, 2, 2, 2)), 34: (BasicBlock, (3, 4, 6, 3)),
    50: (Bottleneck, (3, 4, 6, 3)), 101: (Bottleneck, (3, 4, 23, 3)),
    152: (Bottleneck, (3, 8, 36, 3)),
}


class ResNet(nn.Module):
    def __init__(self, depth: int = 50, num_classes: int = 1000, fused: bool = True,
                 zero_init_residual: bool = True, width: int = 64, fused_stem_tail: bool = False):
        super().__init__()
        block, layers = _CFG[depth]
        self.conv1 = nn.Conv2d(3, width, 7, stride=2, padding=3, bias=False)
        self.bn1 = _bn(width, True, fused)
        self.pool = MaxPool3x3s2() if fused else nn.MaxPool2d(3, stride=2, padding=1)
        # opt-in: BN + ReLU + max-pool of the stem in one pass. Measured a wash (the pooling-type
        # kernels are instruction-bound, not bandwidth-bound: 21.55 vs 21.52 ms per step), so off.
        self.fused_stem_tail = fused and fused_stem_tail
        stages, cin = [], width
        for i, nblk in enumerate(layers):
            w = width * (2 ** i)
            for j in range(nblk):
                stages.append(block(cin, w, (2 if i > 0 and j == 0 else 1), fused,
                                    zero_init_residual))
                cin = w * block.expansion
        self.blocks = nn.Sequential(*stages)
        # classifier GEMM (forward, dgrad, wgrad) on the tcgen05 kernel; plain nn.Linear off-GPU
        self.fc = TcLinear(cin, num_classes) if fused else nn.Linear(cin, num_classes)
        for m in self.modules():
            if isinstance(m, (nn.Conv2d, TcConv2d)):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def forward(self, x):
        if x.shape[1] == 16:
            # input already in the stem's space-to-depth layout (ops/preproc.py): the 7x7/2
            # convolution runs as a 4x4/1 convolution over 16 channels, same parameters
            # tcgen05 implicit GEMM when the shape allows it (ops/conv_stem.py), else cuDNN
            x = stem_conv_s2d(x, stem_weight_s2d(self.conv1.weight))
        else:
            x = self.conv1(x)
        if self.fused_stem_tail:
            # BN + ReLU + max-pool in one pass, BN output never written (ops/bn.bn_relu_maxpool)
            x = bn_relu_maxpool(self.bn1, x)
        else:
            x = self.pool(self.bn1(x))
        x = self.blocks(x)
        x = F.adaptive_avg_pool2d(x, 1).flatten(1)
        return self.fc(x)


def resnet50(**kw) -> ResNet:
    return ResNet(50, **kw)


def resnet18(**kw) -> ResNet:
    return ResNet(18, **kw)
