"""YOLO11 detector in plain PyTorch -- CALLER-SIDE code for the examples and the DDP step test, not part of the hot path.

BASELINE.json configs[2] needs a full train step (PyTorch backbone + the CUDA loss path); the reference model
(/root/reference/yolo_scratch_cuda/models/yolo11_model.py, yolo11_modules.py, configs/yolo11{n,s}_crater.yaml) does not
exist on the GPU box, so this file states the same architecture directly, as a channel plan instead of a YAML that is
eval()-ed: stem, four C3k2 stages, SPPF, C2PSA, a PAN neck and the decoupled Detect head with 4*16 DFL channels.
Width/depth scaling follows the reference's parse_model (models/yolo11_model.py:121-133): channels =
make_divisible(c * width, 8) with NO max_channels clamp, repeats = max(round(n * depth), 1).
tests/test_example_model.py pins the parameter counts to the reference's (n: 3.22 M, s: 11.40 M at nc=80, ch=3).
"""
from __future__ import annotations

import math
from typing import List, Sequence

import torch
import torch.nn as nn

SCALES = {"n": (0.50, 0.25), "s": (0.50, 0.50), "m": (0.50, 1.00), "l": (1.00, 1.00), "x": (1.00, 1.50)}


def _ch(c: int, width: float) -> int:
    return int(math.ceil(c * width / 8) * 8)


def _rep(n: int, depth: float) -> int:
    return max(round(n * depth), 1) if n > 1 else n


class ConvBNAct(nn.Module):
    """conv (no bias, 'same' padding) -> BatchNorm -> SiLU."""

    def __init__(self, cin: int, cout: int, k: int = 1, s: int = 1, groups: int = 1, act: bool = True):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, k, s, k // 2, groups=groups, bias=False)
        self.bn = nn.BatchNorm2d(cout)
        self.act = nn.SiLU() if act else nn.Identity()

    def forward(self, x):
        return self.act(self.bn(self.conv(x)))


class Residual3x3(nn.Module):
    """Two 3x3 conv blocks with an optional skip (the bottleneck of every CSP stage)."""

    def __init__(self, c: int, skip: bool = True):
        super().__init__()
        self.a = ConvBNAct(c, c, 3)
        self.b = ConvBNAct(c, c, 3)
        self.skip = skip

    def forward(self, x):
        y = self.b(self.a(x))
        return x + y if self.skip else y


class CSP3(nn.Module):
    """Cross-stage block with three 1x1 convs around a chain of bottlenecks (the inner block of the deep stages)."""

    def __init__(self, c: int, n: int = 2, skip: bool = True):
        super().__init__()
        h = c // 2
        self.left = ConvBNAct(c, h)
        self.right = ConvBNAct(c, h)
        self.out = ConvBNAct(2 * h, c)
        self.chain = nn.Sequential(*(Residual3x3(h, skip) for _ in range(n)))

    def forward(self, x):
        return self.out(torch.cat((self.chain(self.left(x)), self.right(x)), 1))


class CSPStage(nn.Module):
    """Split -> n inner blocks, every intermediate kept -> concat -> 1x1 (YOLO11's C3k2)."""

    def __init__(self, cin: int, cout: int, n: int, deep: bool, ratio: float = 0.5, skip: bool = True):
        super().__init__()
        self.h = int(cout * ratio)
        self.split = ConvBNAct(cin, 2 * self.h)
        self.merge = ConvBNAct((2 + n) * self.h, cout)
        self.blocks = nn.ModuleList(CSP3(self.h, 2, skip) if deep else Residual3x3(self.h, skip) for _ in range(n))

    def forward(self, x):
        parts = list(self.split(x).split((self.h, self.h), 1))
        for blk in self.blocks:
            parts.append(blk(parts[-1]))
        return self.merge(torch.cat(parts, 1))


class PyramidPool(nn.Module):
    """SPPF: three chained 5x5 max-pools, concatenated."""

    def __init__(self, c: int, k: int = 5):
        super().__init__()
        self.reduce = ConvBNAct(c, c // 2)
        self.expand = ConvBNAct(2 * c, c)
        self.pool = nn.MaxPool2d(k, 1, k // 2)

    def forward(self, x):
        y = [self.reduce(x)]
        for _ in range(3):
            y.append(self.pool(y[-1]))
        return self.expand(torch.cat(y, 1))


class AreaAttention(nn.Module):
    """Multi-head self-attention over the h*w positions of a feature map, with a depth-wise positional conv."""

    def __init__(self, dim: int, heads: int, key_ratio: float = 0.5):
        super().__init__()
        self.heads = heads
        self.dh = dim // heads
        self.dk = int(self.dh * key_ratio)
        self.scale = self.dk ** -0.5
        self.qkv = ConvBNAct(dim, dim + 2 * self.dk * heads, 1, act=False)
        self.proj = ConvBNAct(dim, dim, 1, act=False)
        self.pos = ConvBNAct(dim, dim, 3, groups=dim, act=False)

    def forward(self, x):
        b, c, h, w = x.shape
        q, k, v = self.qkv(x).view(b, self.heads, 2 * self.dk + self.dh, h * w).split((self.dk, self.dk, self.dh), 2)
        att = ((q.transpose(-2, -1) @ k) * self.scale).softmax(-1)
        y = (v @ att.transpose(-2, -1)).view(b, c, h, w) + self.pos(v.reshape(b, c, h, w))
        return self.proj(y)


class AttnBlock(nn.Module):
    """Split -> (attention + feed-forward, both residual) on one half -> concat -> 1x1."""

    def __init__(self, c: int, ratio: float):
        super().__init__()
        self.h = int(c * ratio)
        self.split = ConvBNAct(c, 2 * self.h)
        self.merge = ConvBNAct(2 * self.h, c)
        self.attn = AreaAttention(self.h, heads=max(self.h // 64, 1))
        self.ffn = nn.Sequential(ConvBNAct(self.h, 2 * self.h), ConvBNAct(2 * self.h, self.h, act=False))

    def forward(self, x):
        a, b = self.split(x).split((self.h, self.h), 1)
        b = b + self.attn(b)
        b = b + self.ffn(b)
        return self.merge(torch.cat((a, b), 1))


class AttnStage(nn.Module):
    """C2PSA: split, n attention blocks on one half, concat, 1x1."""

    def __init__(self, c: int, n: int):
        super().__init__()
        self.h = c // 2
        self.split = ConvBNAct(c, 2 * self.h)
        self.merge = ConvBNAct(2 * self.h, c)
        self.blocks = nn.Sequential(*(AttnBlock(self.h, 1.0) for _ in range(n)))

    def forward(self, x):
        a, b = self.split(x).split((self.h, self.h), 1)
        return self.merge(torch.cat((a, self.blocks(b)), 1))


class Detect(nn.Module):
    """Decoupled head: per level 4*reg_max box-distribution channels followed by nc class logits.
    The class is NAMED Detect and carries nc / reg_max / stride because v8DetectionLoss finds its head by that name
    (reference losses/yolo_v8_loss.py:348-351)."""

    def __init__(self, nc: int, chans: Sequence[int], strides: Sequence[float] = (8.0, 16.0, 32.0), reg_max: int = 16):
        super().__init__()
        self.nc, self.reg_max, self.nl = nc, reg_max, len(chans)
        self.no = nc + 4 * reg_max
        self.stride = torch.tensor([float(s) for s in strides])
        cb = max(16, chans[0] // 4, 4 * reg_max)
        cc = max(chans[0], min(nc, 100))
        self.box = nn.ModuleList(nn.Sequential(ConvBNAct(c, cb, 3), ConvBNAct(cb, cb, 3), nn.Conv2d(cb, 4 * reg_max, 1)) for c in chans)
        self.cls = nn.ModuleList(nn.Sequential(ConvBNAct(c, cc, 3), ConvBNAct(cc, cc, 3), nn.Conv2d(cc, nc, 1)) for c in chans)
        for seq_b, seq_c, s in zip(self.box, self.cls, strides):        # the usual detection-head bias init
            seq_b[-1].bias.data.fill_(1.0)
            seq_c[-1].bias.data[:nc] = math.log(5 / nc / (640 / s) ** 2)

    def forward(self, feats: List[torch.Tensor]) -> List[torch.Tensor]:
        return [torch.cat((self.box[i](f), self.cls[i](f)), 1) for i, f in enumerate(feats)]


class YOLO11(nn.Module):
    """forward(img) -> list of 3 NCHW head tensors (B, 4*16+nc, H/8.., W/8..): the training-mode output of the
    reference model (models/yolo11_modules.py:237-243), which is what the loss path consumes."""

    def __init__(self, scale: str = "s", ch: int = 3, nc: int = 80):
        super().__init__()
        depth, width = SCALES[scale]
        c = lambda v: _ch(v, width)
        r = lambda n: _rep(n, depth)
        # backbone
        self.stem = nn.Sequential(ConvBNAct(ch, c(64), 3, 2), ConvBNAct(c(64), c(128), 3, 2))                       # /4
        self.s2 = CSPStage(c(128), c(256), r(2), deep=False, ratio=0.25)
        self.d3 = ConvBNAct(c(256), c(256), 3, 2)                                                                   # /8
        self.s3 = CSPStage(c(256), c(512), r(2), deep=False, ratio=0.25)
        self.d4 = ConvBNAct(c(512), c(512), 3, 2)                                                                   # /16
        self.s4 = CSPStage(c(512), c(512), r(2), deep=True)
        self.d5 = ConvBNAct(c(512), c(1024), 3, 2)                                                                  # /32
        self.s5 = CSPStage(c(1024), c(1024), r(2), deep=True)
        self.spp = PyramidPool(c(1024))
        self.psa = AttnStage(c(1024), r(2))
        # neck (top-down then bottom-up)
        self.up = nn.Upsample(scale_factor=2, mode="nearest")
        self.n4 = CSPStage(c(1024) + c(512), c(512), r(2), deep=False)
        self.n3 = CSPStage(c(512) + c(512), c(256), r(2), deep=False)
        self.dn3 = ConvBNAct(c(256), c(256), 3, 2)
        self.m4 = CSPStage(c(256) + c(512), c(512), r(2), deep=False)
        self.dn4 = ConvBNAct(c(512), c(512), 3, 2)
        self.m5 = CSPStage(c(512) + c(1024), c(1024), r(2), deep=True)
        self.head = Detect(nc, (c(256), c(512), c(1024)))
        self.nc, self.stride = nc, self.head.stride

    def forward(self, x):
        p3 = self.s3(self.d3(self.s2(self.stem(x))))
        p4 = self.s4(self.d4(p3))
        p5 = self.psa(self.spp(self.s5(self.d5(p4))))
        t4 = self.n4(torch.cat((self.up(p5), p4), 1))
        o3 = self.n3(torch.cat((self.up(t4), p3), 1))
        o4 = self.m4(torch.cat((self.dn3(o3), t4), 1))
        o5 = self.m5(torch.cat((self.dn4(o4), p5), 1))
        return self.head([o3, o4, o5])
