"""Stand-ins for the 2D backbone / FPN when mmdet is not installed.

The 2D part of ImVoxelNet is mmdet's (third-party ``ResNet`` + ``FPN``,
configs/imvoxelnet/*.py ``model.backbone`` / ``model.neck``) and is OUT OF
SCOPE of this library: with mmdet importable the detector builds mmdet's own
modules.  Without it (the build and GPU machines have no mmcv/mmdet) these
minimal modules accept the same config keys so that the reference's ``model``
dicts still build and ``extract_feat`` can run end to end on synthetic
images.  They are plain PyTorch (torchvision ResNet, a textbook FPN) and make
no attempt at being fast.
"""
import torch.nn.functional as F
from torch import nn


class ResNet(nn.Module):
    """mmdet-style config keys: depth, num_stages, out_indices, frozen_stages,
    norm_cfg, norm_eval, style (the last four only affect training)."""

    def __init__(self, depth=50, num_stages=4, out_indices=(0, 1, 2, 3), frozen_stages=-1,
                 norm_cfg=None, norm_eval=True, style='pytorch', **_):
        super().__init__()
        import torchvision
        net = getattr(torchvision.models, f'resnet{depth}')(weights=None)
        self.conv1, self.bn1, self.relu, self.maxpool = net.conv1, net.bn1, net.relu, net.maxpool
        for i in range(num_stages):
            setattr(self, f'layer{i + 1}', getattr(net, f'layer{i + 1}'))
        self.num_stages, self.out_indices = num_stages, tuple(out_indices)

    def init_weights(self, pretrained=None):
        pass

    def forward(self, x):
        x = self.maxpool(self.relu(self.bn1(self.conv1(x))))
        outs = []
        for i in range(self.num_stages):
            x = getattr(self, f'layer{i + 1}')(x)
            if i in self.out_indices:
                outs.append(x)
        return tuple(outs)


class _Conv(nn.Module):
    """Holder with mmcv ConvModule's key layout (``conv.weight`` / ``conv.bias``)."""

    def __init__(self, cin, cout, k):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, k, padding=k // 2)

    def forward(self, x):
        return self.conv(x)


class FPN(nn.Module):
    """mmdet-style config keys: in_channels, out_channels, num_outs."""

    def __init__(self, in_channels, out_channels, num_outs, **_):
        super().__init__()
        self.lateral_convs = nn.ModuleList(_Conv(c, out_channels, 1) for c in in_channels)
        self.fpn_convs = nn.ModuleList(_Conv(out_channels, out_channels, 3) for _ in in_channels)
        self.num_outs = num_outs

    def init_weights(self):
        pass

    def forward(self, inputs):
        lat = [l(x) for l, x in zip(self.lateral_convs, inputs)]
        for i in range(len(lat) - 1, 0, -1):
            lat[i - 1] = lat[i - 1] + F.interpolate(lat[i], size=lat[i - 1].shape[2:], mode='nearest')
        outs = [c(x) for c, x in zip(self.fpn_convs, lat)]
        while len(outs) < self.num_outs:
            outs.append(F.max_pool2d(outs[-1], 1, stride=2))
        return tuple(outs[:self.num_outs])
