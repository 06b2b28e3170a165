"""ResNet + FPN backbone with frozen batch norm folded into the convolutions.

tensorpack's and aws-samples' Mask R-CNN train with `BACKBONE.NORM=FreezeBN`
(examples/maskrcnn/train-maskrcnn-tensorpack.yaml:33): the batch-norm statistics and affine terms are constants, so
`bn(conv(x, W)) = conv(x, W * s) + t` with `s = gamma / sqrt(var + eps)` and `t = beta - mean * s` per output channel.
This module applies that identity instead of running a normalisation pass over every activation: per convolution one
multiply on the *weight* (cached for the frozen stem and first stage), the shift rides in the convolution's bias, and
ReLU is applied in place.  Convolutions stay on cuDNN (channels_last, bf16).

Parameter and buffer names follow the torchvision layout (`body.layer1.0.conv1.weight`, `body.layer1.0.bn1.running_mean`,
`fpn.inner_blocks.0.0.weight`, ...), so tensorpack `.npz` backbones (`models/maskrcnn.py::load_tensorpack_backbone`) and
checkpoints written by earlier versions load unchanged.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn

from ..ops.detection import bias_act_, conv_bias_act


# Random initialisation only (a backbone checkpoint overwrites every buffer): with the statistics frozen at identity
# nothing re-normalises the residual stream, so the last scale of every block starts small and the activation variance
# grows by (1 + RESIDUAL_GAIN^2) per block instead of doubling.
RESIDUAL_GAIN = 0.25


class ConvW(nn.Module):
    """A convolution's weight plus its geometry (the normalisation that follows it owns the bias)."""

    def __init__(self, cin, cout, k, stride=1, padding=0):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(cout, cin, k, k))
        nn.init.kaiming_normal_(self.weight, mode="fan_in", nonlinearity="relu")   # variance-preserving under ReLU
        self.stride, self.padding = stride, padding


class FrozenBN(nn.Module):
    """Constant batch-norm terms (buffers, as torchvision's FrozenBatchNorm2d keeps them)."""

    def __init__(self, c, eps=1e-5):
        super().__init__()
        self.eps = eps
        self.register_buffer("weight", torch.ones(c))
        self.register_buffer("bias", torch.zeros(c))
        self.register_buffer("running_mean", torch.zeros(c))
        self.register_buffer("running_var", torch.ones(c))
        self._cache = None

    def scale_shift(self, dtype):
        """(s [C, 1, 1, 1], t [C]) in `dtype`; recomputed when the buffers were replaced (load_state_dict, .to())."""
        key = (self.weight._version, self.running_var._version, self.running_mean._version, self.bias._version,
               self.weight.data_ptr(), dtype)
        if self._cache is None or self._cache[0] != key:
            s = self.weight.float() * torch.rsqrt(self.running_var.float() + self.eps)
            t = self.bias.float() - self.running_mean.float() * s
            self._cache = (key, s.view(-1, 1, 1, 1).to(dtype), t.to(dtype))
        return self._cache[1], self._cache[2]


def conv_bn(x, conv: ConvW, bn: FrozenBN, relu: bool, cache: dict | None = None, residual=None):
    """relu?(bn(conv(x)) (+ residual)) as one cuDNN convolution with a scaled weight, followed by ONE in-place pass
    over the output for shift, residual and ReLU (`ops.detection.bias_act_`; cuDNN would add the bias with a separate
    broadcast kernel and leave the other two to two more passes)."""
    s, t = bn.scale_shift(x.dtype)
    w = conv.weight
    if not w.requires_grad and cache is not None:          # frozen layers: fold once per weight version
        key = (id(conv), w._version, w.data_ptr(), x.dtype)
        hit = cache.get(id(conv))
        if hit is None or hit[0] != key:
            hit = (key, (w.detach().to(x.dtype) * s).contiguous(memory_format=torch.channels_last))
            cache[id(conv)] = hit
        wf = hit[1]
    else:
        wf = w.to(x.dtype) * s
    y = F.conv2d(x, wf, None, conv.stride, conv.padding)
    return bias_act_(y, t, residual, relu)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin, planes, stride=1, downsample=False):
        super().__init__()
        self.conv1, self.bn1 = ConvW(cin, planes, 1), FrozenBN(planes)
        self.conv2, self.bn2 = ConvW(planes, planes, 3, stride, 1), FrozenBN(planes)       # stride on the 3x3 (v1.5)
        self.conv3, self.bn3 = ConvW(planes, planes * 4, 1), FrozenBN(planes * 4)
        self.bn3.weight.fill_(RESIDUAL_GAIN)
        self.downsample = None
        if downsample:
            self.downsample = nn.Sequential(ConvW(cin, planes * 4, 1, stride), FrozenBN(planes * 4))

    def forward(self, x, cache):
        idt = x if self.downsample is None else conv_bn(x, self.downsample[0], self.downsample[1], False, cache)
        y = conv_bn(x, self.conv1, self.bn1, True, cache)
        y = conv_bn(y, self.conv2, self.bn2, True, cache)
        return conv_bn(y, self.conv3, self.bn3, True, cache, residual=idt)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, cin, planes, stride=1, downsample=False):
        super().__init__()
        self.conv1, self.bn1 = ConvW(cin, planes, 3, stride, 1), FrozenBN(planes)
        self.conv2, self.bn2 = ConvW(planes, planes, 3, 1, 1), FrozenBN(planes)
        self.bn2.weight.fill_(RESIDUAL_GAIN)
        self.downsample = None
        if downsample:
            self.downsample = nn.Sequential(ConvW(cin, planes, 1, stride), FrozenBN(planes))

    def forward(self, x, cache):
        idt = x if self.downsample is None else conv_bn(x, self.downsample[0], self.downsample[1], False, cache)
        y = conv_bn(x, self.conv1, self.bn1, True, cache)
        return conv_bn(y, self.conv2, self.bn2, True, cache, residual=idt)


class _Stage(nn.ModuleList):
    def forward(self, x, cache):
        for blk in self:
            x = blk(x, cache)
        return x


class ResNetBody(nn.Module):
    """conv1 .. layer4 of a ResNet; returns the four stage outputs (strides 4, 8, 16, 32)."""

    def __init__(self, block, layers):
        super().__init__()
        self.conv1, self.bn1 = ConvW(3, 64, 7, 2, 3), FrozenBN(64)
        cin, stages = 64, []
        for i, (planes, n) in enumerate(zip((64, 128, 256, 512), layers)):
            blocks = []
            for b in range(n):
                stride = 2 if (b == 0 and i > 0) else 1
                blocks.append(block(cin, planes, stride, downsample=(b == 0 and (stride != 1 or cin != planes * block.expansion))))
                cin = planes * block.expansion
            stages.append(_Stage(blocks))
        self.layer1, self.layer2, self.layer3, self.layer4 = stages
        self.out_channels = [64 * block.expansion, 128 * block.expansion, 256 * block.expansion, 512 * block.expansion]
        self._cache: dict = {}

    def forward(self, x):
        c = self._cache
        x = conv_bn(x, self.conv1, self.bn1, True, c)
        x = F.max_pool2d(x, 3, 2, 1)
        c2 = self.layer1(x, c)
        c3 = self.layer2(c2, c)
        c4 = self.layer3(c3, c)
        c5 = self.layer4(c4, c)
        return [c2, c3, c4, c5]


def resnet50_body() -> ResNetBody:
    return ResNetBody(Bottleneck, (3, 4, 6, 3))


def resnet18_body() -> ResNetBody:
    return ResNetBody(BasicBlock, (2, 2, 2, 2))


class FPN(nn.Module):
    """Top-down pathway: 1x1 lateral convolutions, nearest 2x upsampling, 3x3 output convolutions, and a stride-2
    max-pool of the coarsest map as the extra level (tensorpack's `fpn_model`; names as torchvision's
    FeaturePyramidNetwork)."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.inner_blocks = nn.ModuleList(nn.Sequential(nn.Conv2d(c, out_channels, 1)) for c in in_channels)
        self.layer_blocks = nn.ModuleList(nn.Sequential(nn.Conv2d(out_channels, out_channels, 3, padding=1))
                                          for _ in in_channels)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_uniform_(m.weight, a=1)
                nn.init.zeros_(m.bias)

    def forward(self, feats):
        last = conv_bias_act(feats[-1], self.inner_blocks[-1][0])
        outs = [conv_bias_act(last, self.layer_blocks[-1][0])]
        for i in range(len(feats) - 2, -1, -1):
            lat = conv_bias_act(feats[i], self.inner_blocks[i][0])
            last = lat + F.interpolate(last, size=lat.shape[-2:], mode="nearest")
            outs.insert(0, conv_bias_act(last, self.layer_blocks[i][0]))
        outs.append(F.max_pool2d(outs[-1], 1, 2, 0))
        return outs


class ResNetFPN(nn.Module):
    """`body` + `fpn`; forward returns {"0", "1", "2", "3", "pool"} like torchvision's BackboneWithFPN."""

    def __init__(self, depth: int, out_channels: int, frozen_prefixes=("conv1", "layer1")):
        super().__init__()
        self.body = resnet50_body() if depth == 50 else resnet18_body()
        for name, p in self.body.named_parameters():           # tensorpack freezes the stem and the first stage
            if name.startswith(frozen_prefixes):
                p.requires_grad_(False)
        self.fpn = FPN(self.body.out_channels, out_channels)
        self.out_channels = out_channels

    def forward(self, x):
        outs = self.fpn(self.body(x))
        return dict(zip(("0", "1", "2", "3", "pool"), outs))
