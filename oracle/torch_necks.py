"""Plain-PyTorch restatement of the reference's four 3D necks.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): the fp32 reference the CUDA
neck kernels are compared with.  Module/parameter names follow
/root/reference/mmdet3d/models/necks/imvoxelnet.py so that state dicts are
interchangeable with the reference modules (checked in
tests/test_neck_oracle.py against the live reference and against goldens):

  FastIndoorImVoxelNeck :8-67     ResBlockV2 (= BasicBlock3dV2 :233-260)
  ImVoxelNeck           :70-91    EncoderDecoder :297-372, ResBlock (= BasicBlock3d :191-230),
                                  SkipProjection (= ConditionalProjection :263-294)
  KittiImVoxelNeck      :94-123
  NuScenesImVoxelNeck   :126-154
"""
import torch
import torch.nn.functional as F
from torch import nn


def _conv_bn_relu(cin, cout, stride=1, padding=1, bias=False):
    return nn.Sequential(nn.Conv3d(cin, cout, 3, stride, padding, bias=bias), nn.BatchNorm3d(cout),
                         nn.ReLU(inplace=True))


class ResBlockV2(nn.Module):
    """conv-bn-relu-conv-bn (+ 1x1x1 strided projection of the input) -> add -> relu."""

    def __init__(self, cin, cout, stride=1):
        super().__init__()
        self.conv1 = nn.Conv3d(cin, cout, 3, stride, 1, bias=False)
        self.norm1 = nn.BatchNorm3d(cout)
        self.conv2 = nn.Conv3d(cout, cout, 3, 1, 1, bias=False)
        self.norm2 = nn.BatchNorm3d(cout)
        self.stride = stride
        if stride != 1:
            self.downsample = nn.Sequential(nn.Conv3d(cin, cout, 1, stride, bias=False), nn.BatchNorm3d(cout))

    def forward(self, x):
        y = F.relu(self.norm1(self.conv1(x)))
        y = self.norm2(self.conv2(y))
        skip = self.downsample(x) if self.stride != 1 else x
        return F.relu(y + skip)


class FastIndoorImVoxelNeck(nn.Module):
    def __init__(self, in_channels, n_blocks, out_channels):
        super().__init__()
        self.n_scales = len(n_blocks)
        c = in_channels
        for i, n in enumerate(n_blocks):
            stride = 1 if i == 0 else 2
            blocks = []
            for k in range(n):
                if k == 0 and stride != 1:
                    blocks.append(ResBlockV2(c, c * 2, stride))
                    c *= 2
                else:
                    blocks.append(ResBlockV2(c, c))
            setattr(self, f'down_layer_{i}', nn.Sequential(*blocks))
            if i > 0:
                setattr(self, f'up_block_{i}', nn.Sequential(
                    nn.ConvTranspose3d(c, c // 2, 2, 2, bias=False), nn.BatchNorm3d(c // 2),
                    nn.ReLU(inplace=True),
                    nn.Conv3d(c // 2, c // 2, 3, 1, 1, bias=False), nn.BatchNorm3d(c // 2),
                    nn.ReLU(inplace=True)))
            setattr(self, f'out_block_{i}', _conv_bn_relu(c, out_channels))

    def forward(self, x):
        skips = []
        for i in range(self.n_scales):
            x = getattr(self, f'down_layer_{i}')(x)
            skips.append(x)
        outs = []
        for i in reversed(range(self.n_scales)):
            if i < self.n_scales - 1:
                x = skips[i] + getattr(self, f'up_block_{i + 1}')(x)
            outs.append(getattr(self, f'out_block_{i}')(x))
        return outs[::-1]


class ResBlock(nn.Module):
    """Atlas basic block: conv-bn-relu-conv-bn -> add input -> relu (dropout p=0)."""

    def __init__(self, c):
        super().__init__()
        self.conv1 = nn.Conv3d(c, c, 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm3d(c)
        self.drop1 = nn.Dropout(0, True)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = nn.Conv3d(c, c, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm3d(c)
        self.drop2 = nn.Dropout(0, True)

    def forward(self, x):
        y = F.relu(self.bn1(self.conv1(x)))
        y = self.bn2(self.conv2(y))
        return F.relu(y + x)


class SkipProjection(nn.Module):
    """necks/imvoxelnet.py:263-294: relu(bn(conv1x1x1(skip))); with `condition` the decoder tensor
    replaces the projected skip where the mask is off, before the norm (:290-291)."""

    def __init__(self, c, condition=False):
        super().__init__()
        self.conv = nn.Conv3d(c, c, 1, bias=False)
        self.norm = nn.BatchNorm3d(c)
        self.condition = condition

    def forward(self, x, decoder=None, mask=None):
        x = self.conv(x)
        if self.condition:
            x = torch.where(mask, x, decoder)
        return F.relu(self.norm(x))


class EncoderDecoder(nn.Module):
    def __init__(self, channels, layers_down, layers_up, cond_proj=False):
        super().__init__()
        self.cond_proj = cond_proj
        self.layers_down = nn.ModuleList()
        proj = []
        self.layers_down.append(nn.Sequential(*[ResBlock(channels[0]) for _ in range(layers_down[0])]))
        proj.append(SkipProjection(channels[0], cond_proj))
        for i in range(1, len(channels)):
            layer = [nn.Conv3d(channels[i - 1], channels[i], 3, 2, 1, bias=False),
                     nn.BatchNorm3d(channels[i]), nn.Dropout(0, True), nn.ReLU(inplace=True)]
            layer += [ResBlock(channels[i]) for _ in range(layers_down[i])]
            self.layers_down.append(nn.Sequential(*layer))
            if i < len(channels) - 1:
                proj.append(SkipProjection(channels[i], cond_proj))
        self.proj = nn.ModuleList(proj[::-1])
        rev = channels[::-1]
        self.layers_up_conv = nn.ModuleList()
        self.layers_up_res = nn.ModuleList()
        for i in range(1, len(rev)):
            self.layers_up_conv.append(nn.Conv3d(rev[i - 1], rev[i], 1, bias=False))
            self.layers_up_res.append(nn.Sequential(*[ResBlock(rev[i]) for _ in range(layers_up[i - 1])]))
        for m in self.modules():                      # zero-init of the residual branches (:340-343)
            if isinstance(m, ResBlock):
                nn.init.constant_(m.bn2.weight, 0)

    def forward(self, x):
        if self.cond_proj:                            # :347-348: voxels with any non-zero input channel
            valid_mask = (x != 0).any(1, keepdim=True).float()
        feats = []
        for layer in self.layers_down:
            x = layer(x)
            feats.append(x)
        feats = feats[::-1]
        outs = []
        for i in range(len(self.layers_up_conv)):
            x = F.interpolate(x, scale_factor=2, mode='trilinear', align_corners=False)
            x = self.layers_up_conv[i](x)
            mask = None
            if self.cond_proj:                        # :361-363: nearest-neighbour resampling of the mask
                scale = 1 / 2 ** (len(self.layers_up_conv) - i - 1)
                mask = F.interpolate(valid_mask, scale_factor=scale) != 0
            x = (x + self.proj[i](feats[i + 1], x, mask)) / 2
            x = self.layers_up_res[i](x)
            outs.append(x)
        return outs


class ImVoxelNeck(nn.Module):
    def __init__(self, channels, out_channels, down_layers, up_layers, conditional):
        super().__init__()
        self.model = EncoderDecoder(list(channels), list(down_layers), list(up_layers), bool(conditional))
        self.conv_blocks = nn.ModuleList([_conv_bn_relu(c, out_channels, bias=True) for c in channels])

    def forward(self, x):
        x = self.model(x)[::-1]
        return [self.conv_blocks[i](x[i]) for i in range(len(x))]


class _BevNeck(nn.Module):
    def forward(self, x):
        x = self.model(x)
        assert x.shape[-1] == 1
        return [x[..., 0].transpose(-1, -2)]


class KittiImVoxelNeck(_BevNeck):
    def __init__(self, in_channels, out_channels):
        super().__init__()
        c = in_channels
        self.model = nn.Sequential(
            ResBlock(c), _conv_bn_relu(c, c * 2, (1, 1, 2), 1, bias=True),
            ResBlock(c * 2), _conv_bn_relu(c * 2, c * 4, (1, 1, 2), 1, bias=True),
            ResBlock(c * 4), _conv_bn_relu(c * 4, out_channels, 1, 0, bias=True))


class NuScenesImVoxelNeck(_BevNeck):
    def __init__(self, in_channels, out_channels):
        super().__init__()
        c = in_channels
        self.model = nn.Sequential(
            ResBlock(c), _conv_bn_relu(c, c * 2, 2, 1, bias=True),
            ResBlock(c * 2), _conv_bn_relu(c * 2, c * 4, (1, 1, 2), 1, bias=True),
            ResBlock(c * 4), _conv_bn_relu(c * 4, out_channels, 1, (1, 1, 0), bias=True))


NECKS = {cls.__name__: cls for cls in
         (FastIndoorImVoxelNeck, ImVoxelNeck, KittiImVoxelNeck, NuScenesImVoxelNeck)}


def randomize_norms(module, seed=0):
    """Random BN affine and running statistics: with fresh BN
    (and the zero-initialised bn2 of the Atlas blocks) residual branches would
    be trivial and a wrong kernel could pass."""
    g = torch.Generator().manual_seed(seed)
    for m in module.modules():
        if isinstance(m, nn.BatchNorm3d):
            with torch.no_grad():
                m.running_mean.copy_(torch.randn(m.num_features, generator=g) * 0.1)
                m.running_var.copy_(torch.rand(m.num_features, generator=g) + 0.5)
                m.weight.copy_(torch.rand(m.num_features, generator=g) + 0.5)
                m.bias.copy_(torch.randn(m.num_features, generator=g) * 0.1)
    return module
