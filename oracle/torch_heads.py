"""Plain-PyTorch restatement of the dense heads' convolution tail.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): the fp32 reference the head
kernels (imvoxelnet_b200/heads.py, csrc/ivx_head.cu) are compared with.  Module
and parameter names follow the reference so state dicts are interchangeable
(checked against the live reference classes and against committed goldens in
tests/test_head_oracle.py):

  HeadV2(box_angle=False)  ScanNetImVoxelHeadV2.forward_single   imvoxel_head_v2.py:444-449
  HeadV2(box_angle=True)   SunRgbdImVoxelHeadV2.forward_single   imvoxel_head_v2.py:305-313
  HeadV1(box_angle=False)  ScanNetImVoxelHead.forward_single     imvoxel_head.py:454-461
  HeadV1(box_angle=True)   SunRgbdImVoxelHead.forward_single     imvoxel_head.py:326-336
  resize_valid             imvoxel_head_v2.py:91-93, 224-226
  scores                   imvoxel_head_v2.py:262-270

``mmcv.cnn.Scale`` (mmcv-full 1.2.7, absent from /root/reference) is a learnable
scalar: ``x * self.scale``.
"""
import torch
from torch import nn


class Scale(nn.Module):
    def __init__(self, scale=1.0):
        super().__init__()
        self.scale = nn.Parameter(torch.tensor(scale, dtype=torch.float))

    def forward(self, x):
        return x * self.scale


def _box(reg_final, scale, box_angle):
    if box_angle:       # six distances through exp, the heading angle untouched
        return torch.cat((torch.exp(scale(reg_final[:, :6])), reg_final[:, 6:]), dim=1)
    return torch.exp(scale(reg_final))


class HeadV2(nn.Module):
    def __init__(self, n_classes, n_channels, n_reg_outs, n_scales, box_angle=False):
        super().__init__()
        self.box_angle = box_angle
        self.centerness_conv = nn.Conv3d(n_channels, 1, 3, padding=1, bias=False)
        self.reg_conv = nn.Conv3d(n_channels, n_reg_outs, 3, padding=1, bias=False)
        self.cls_conv = nn.Conv3d(n_channels, n_classes, 3, padding=1)
        self.scales = nn.ModuleList([Scale(1.) for _ in range(n_scales)])

    def forward(self, xs):
        outs = [(self.centerness_conv(x), _box(self.reg_conv(x), s, self.box_angle), self.cls_conv(x))
                for x, s in zip(xs, self.scales)]
        return tuple(map(list, zip(*outs)))


class HeadV1(nn.Module):
    def __init__(self, n_classes, n_channels, n_convs, n_reg_outs, n_scales=3, box_angle=False):
        super().__init__()
        self.box_angle = box_angle

        def stack():
            return nn.Sequential(*[nn.Sequential(nn.Conv3d(n_channels, n_channels, 3, padding=1, bias=False),
                                                 nn.BatchNorm3d(n_channels), nn.ReLU(inplace=True))
                                   for _ in range(n_convs)])

        self.reg_convs, self.cls_convs = stack(), stack()
        self.centerness_conv = nn.Conv3d(n_channels, 1, 3, padding=1, bias=False)
        self.reg_conv = nn.Conv3d(n_channels, n_reg_outs, 3, padding=1, bias=False)
        self.cls_conv = nn.Conv3d(n_channels, n_classes, 3, padding=1)
        self.scales = nn.ModuleList([Scale(1.) for _ in range(n_scales)])

    def forward(self, xs):
        outs = []
        for x, s in zip(xs, self.scales):
            reg, cls = self.reg_convs(x), self.cls_convs(x)
            outs.append((self.centerness_conv(reg), _box(self.reg_conv(reg), s, self.box_angle),
                         self.cls_conv(cls)))
        return tuple(map(list, zip(*outs)))


def resize_valid(valid, size):
    """valid [B,1,X,Y,Z] float 0/1 -> bool [B,1,*size]."""
    return nn.Upsample(size=tuple(size), mode='trilinear')(valid.float()).round().bool()


def scores(centerness, cls_score, valid):
    """One scene: centerness [1,X,Y,Z], cls_score [n,X,Y,Z], valid bool [1,X,Y,Z] -> (scores [S,n], max [S])."""
    n = cls_score.shape[0]
    c = centerness.permute(1, 2, 3, 0).reshape(-1).sigmoid()
    s = cls_score.permute(1, 2, 3, 0).reshape(-1, n).sigmoid()
    v = valid.permute(1, 2, 3, 0).reshape(-1)
    s = s * c[:, None] * v[:, None]
    return s, s.max(dim=1)[0]


def randomise(head, seed):
    """Seeded non-trivial parameters (conv weights ~N(0, .05), BN statistics, Scale != 1)."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for m in head.modules():
            if isinstance(m, nn.Conv3d):
                m.weight.copy_(torch.randn(m.weight.shape, generator=g) * 0.05)
                if m.bias is not None:
                    m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
            elif isinstance(m, nn.BatchNorm3d):
                m.running_mean.copy_(torch.randn(m.running_mean.shape, generator=g) * 0.1)
                m.running_var.copy_(torch.rand(m.running_var.shape, generator=g) + 0.5)
                m.weight.copy_(torch.rand(m.weight.shape, generator=g) + 0.5)
                m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
        for i, s in enumerate(head.scales):
            s.scale.fill_(0.8 + 0.15 * i)
    return head
