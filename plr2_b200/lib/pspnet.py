"""PSPNet head over the dilated ResNet (plain PyTorch; outside the CUDA hot path).  Produces the
32-channel per-pixel log-softmax map that PoseNet gathers at ``choose``.  Attribute names follow the
reference's lib/pspnet.py:7-77 so checkpoints load (``psp.stages.0.1.weight``, ``up_1.conv.1.weight``,
``final.0.weight``, ``classifier.0.weight`` ...)."""
import torch
from torch import nn
from torch.nn import functional as F

from . import extractors


class PSPModule(nn.Module):
    def __init__(self, features, out_features=1024, sizes=(1, 2, 3, 6)):
        super().__init__()
        self.stages = nn.ModuleList(
            nn.Sequential(nn.AdaptiveAvgPool2d((s, s)), nn.Conv2d(features, features, 1, bias=False)) for s in sizes)
        self.bottleneck = nn.Conv2d(features * (len(sizes) + 1), out_features, 1)
        self.relu = nn.ReLU()

    def forward(self, feats):
        size = feats.shape[2:]
        pyramid = [F.interpolate(stage(feats), size=size, mode='bilinear') for stage in self.stages]
        return self.relu(self.bottleneck(torch.cat(pyramid + [feats], 1)))


class PSPUpsample(nn.Module):
    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.conv = nn.Sequential(nn.Upsample(scale_factor=2, mode='bilinear', align_corners=True),
                                  nn.Conv2d(in_channels, out_channels, 3, padding=1), nn.PReLU())

    def forward(self, x):
        return self.conv(x)


class PSPNet(nn.Module):
    def __init__(self, n_classes=21, sizes=(1, 2, 3, 6), psp_size=2048, deep_features_size=1024,
                 backend='resnet18', pretrained=False):
        super().__init__()
        self.feats = getattr(extractors, backend)(pretrained)
        self.psp = PSPModule(psp_size, 1024, sizes)
        self.drop_1 = nn.Dropout2d(p=0.3)
        self.up_1 = PSPUpsample(1024, 256)
        self.up_2 = PSPUpsample(256, 64)
        self.up_3 = PSPUpsample(64, 64)
        self.drop_2 = nn.Dropout2d(p=0.15)
        self.final = nn.Sequential(nn.Conv2d(64, 32, 1), nn.LogSoftmax(dim=1))
        # unused by forward, present in the reference's checkpoints (lib/pspnet.py:58-62)
        self.classifier = nn.Sequential(nn.Linear(deep_features_size, 256), nn.ReLU(), nn.Linear(256, n_classes))

    def forward(self, x):
        f, _ = self.feats(x)
        p = self.drop_1(self.psp(f))
        p = self.drop_2(self.up_1(p))
        p = self.drop_2(self.up_2(p))
        return self.final(self.up_3(p))
