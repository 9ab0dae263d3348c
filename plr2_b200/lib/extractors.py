"""Dilated, BatchNorm-free ResNet feature extractor used by the PSPNet backbone.  Plain PyTorch
(the backbone is outside the CUDA hot path); attribute names follow the reference's
lib/extractors.py:78-124 so that its checkpoints load (``feats.layer1.0.conv1.weight`` ...).
Only the BasicBlock variants the reference instantiates by default (resnet18/34) plus the
bottleneck variants are provided."""
import math

import torch.nn as nn


def _conv3x3(cin, cout, stride=1, dilation=1):
    return nn.Conv2d(cin, cout, 3, stride=stride, padding=dilation, dilation=dilation, bias=False)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, downsample=None, dilation=1):
        super().__init__()
        self.conv1 = _conv3x3(inplanes, planes, stride, dilation)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = _conv3x3(planes, planes, 1, dilation)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        skip = x if self.downsample is None else self.downsample(x)
        y = self.conv2(self.relu(self.conv1(x)))
        return self.relu(y + skip)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None, dilation=1):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 1, bias=False)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride=stride, dilation=dilation, padding=dilation, bias=False)
        self.conv3 = nn.Conv2d(planes, planes * 4, 1, bias=False)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        skip = x if self.downsample is None else self.downsample(x)
        y = self.relu(self.conv1(x))
        y = self.relu(self.conv2(y))
        return self.relu(self.conv3(y) + skip)


class ResNet(nn.Module):
    """conv7x7/2 -> maxpool/2 -> 4 stages; stages 3 and 4 keep stride 1 and dilate by 2 and 4, so
    the output is 1/8 of the input resolution.  Returns (stage-4 features, stage-3 features)."""

    def __init__(self, block, layers=(3, 4, 23, 3)):
        super().__init__()
        self.inplanes = 64
        self.conv1 = nn.Conv2d(3, 64, 7, stride=2, padding=3, bias=False)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(3, stride=2, padding=1)
        self.layer1 = self._stage(block, 64, layers[0])
        self.layer2 = self._stage(block, 128, layers[1], stride=2)
        self.layer3 = self._stage(block, 256, layers[2], dilation=2)
        self.layer4 = self._stage(block, 512, layers[3], dilation=4)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):   # He init on fan-out, as the reference does (:94-97)
                fan = m.kernel_size[0] * m.kernel_size[1] * m.out_channels
                m.weight.data.normal_(0, math.sqrt(2.0 / fan))

    def _stage(self, block, planes, blocks, stride=1, dilation=1):
        down = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            down = nn.Sequential(nn.Conv2d(self.inplanes, planes * block.expansion, 1, stride=stride, bias=False))
        mods = [block(self.inplanes, planes, stride, down)]      # first block is never dilated (:111)
        self.inplanes = planes * block.expansion
        mods += [block(self.inplanes, planes, dilation=dilation) for _ in range(1, blocks)]
        return nn.Sequential(*mods)

    def forward(self, x):
        x = self.maxpool(self.relu(self.conv1(x)))
        x = self.layer2(self.layer1(x))
        x3 = self.layer3(x)
        return self.layer4(x3), x3


def resnet18(pretrained=False):
    return ResNet(BasicBlock, [2, 2, 2, 2])


def resnet34(pretrained=False):
    return ResNet(BasicBlock, [3, 4, 6, 3])


def resnet50(pretrained=False):
    return ResNet(Bottleneck, [3, 4, 6, 3])


def resnet101(pretrained=False):
    return ResNet(Bottleneck, [3, 4, 23, 3])


def resnet152(pretrained=False):
    return ResNet(Bottleneck, [3, 8, 36, 3])
