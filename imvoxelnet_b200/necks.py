"""Drop-in 3D necks: same registry names, constructor arguments, state-dict
keys and ``forward(x) -> list[Tensor]`` as mmdet3d/models/necks/imvoxelnet.py,
executed by the fused CUDA layers of ``conv.py`` in eval mode.

The module tree mirrors the reference's attribute names exactly
(``down_layer_i``, ``up_block_i``, ``out_block_i``; ``model.layers_down`` ...;
``conv_blocks``; ``model.{0..5}``) so released checkpoints load unchanged; the
``nn.Conv3d`` / ``nn.BatchNorm3d`` children only hold parameters.

In ``.train()`` mode (fp32 only) the same layers run through ``train.py``:
BatchNorm with batch statistics (running statistics updated as nn.BatchNorm3d
does) and autograd, every forward and backward step a kernel of this library
(SURVEY.md section 8(f)-1).
"""
import torch
from torch import nn

from .conv import FusedConv, from_ndhwc, to_ndhwc, upsample2x_trilinear
from .conv import fp32x3 as conv_fp32x3
from .conv import tf32 as conv_tf32
from .registry import NECKS, register

__all__ = ["FastIndoorImVoxelNeck", "ImVoxelNeck", "KittiImVoxelNeck", "NuScenesImVoxelNeck"]

PRECISIONS = ("fp32", "fp32_simt", "tf32", "bf16")


class _NeckBase(nn.Module):
    #: "fp32": fp32 activations, matches the reference to 1e-5: layers with c_in % 32 == 0 and
    #:         c_out % 64 == 0 run on the tensor cores as three tf32 products (3xTF32), the rest on
    #:         the CUDA-core kernel.
    #: "fp32_simt": fp32 activations, every layer on the CUDA-core kernel (1e-5; the slow cross-check).
    #: "tf32": fp32 activations, tf32 tensor-core math (what cuDNN does for fp32 by default; 1e-3).
    #: "bf16": bf16 activations, tensor-core kernels where available (1e-2).
    precision = "fp32"

    def __call__(self, *args, **kwargs):
        with conv_tf32(self.precision == "tf32"), conv_fp32x3(self.precision != "fp32_simt"):
            return super().__call__(*args, **kwargs)

    def init_weights(self):
        pass

    def _enter(self, x):
        if self.precision not in PRECISIONS:
            raise ValueError(f"precision must be one of {PRECISIONS}")
        if self.training and torch.is_grad_enabled() and self.precision == "bf16":
            raise NotImplementedError(
                f"{type(self).__name__}: training keeps fp32 tensors (batch-statistics BatchNorm and autograd); "
                "set precision='fp32' or 'tf32'")
        y = to_ndhwc(x)
        return y.to(torch.bfloat16) if self.precision == "bf16" else y.float()

    @staticmethod
    def _leave(y):
        return from_ndhwc(y)


# ----------------------------- blocks (parameter holders + fused forward) -----------------------------
class BasicBlock3dV2(nn.Module):
    """reference necks/imvoxelnet.py:233-260."""

    def __init__(self, in_channels, out_channels, stride=1):
        super().__init__()
        self.stride = stride
        self.conv1 = nn.Conv3d(in_channels, out_channels, 3, stride, 1, bias=False)
        self.norm1 = nn.BatchNorm3d(out_channels)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = nn.Conv3d(out_channels, out_channels, 3, 1, 1, bias=False)
        self.norm2 = nn.BatchNorm3d(out_channels)
        if stride != 1:
            self.downsample = nn.Sequential(nn.Conv3d(in_channels, out_channels, 1, stride, bias=False),
                                            nn.BatchNorm3d(out_channels))
        self._f1 = FusedConv(self.conv1, self.norm1, relu=True)
        self._f2 = FusedConv(self.conv2, self.norm2, relu=True)
        self._fd = FusedConv(self.downsample[0], self.downsample[1]) if stride != 1 else None

    def forward(self, x):
        skip = self._fd(x) if self._fd is not None else x
        return self._f2(self._f1(x), residual=skip)


class BasicBlock3d(nn.Module):
    """reference necks/imvoxelnet.py:191-230 (norm='BN', drop=0, no downsample)."""

    def __init__(self, inplanes, planes):
        super().__init__()
        self.conv1 = nn.Conv3d(inplanes, planes, 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm3d(planes)
        self.drop1 = nn.Dropout(0, True)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = nn.Conv3d(planes, planes, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm3d(planes)
        self.drop2 = nn.Dropout(0, True)
        self._f1 = FusedConv(self.conv1, self.bn1, relu=True)
        self._f2 = FusedConv(self.conv2, self.bn2, relu=True)

    def forward(self, x):
        return self._f2(self._f1(x), residual=x)


class _ConvBnRelu(nn.Sequential):
    """nn.Sequential(conv, bn, relu) holder (keys ``0.*``, ``1.*``) with a fused forward."""

    def __init__(self, conv):
        super().__init__(conv, nn.BatchNorm3d(conv.out_channels), nn.ReLU(inplace=True))
        self._f = FusedConv(self[0], self[1], relu=True)

    def forward(self, x, **kw):
        return self._f(x, **kw)


class _UpBlock(nn.Sequential):
    """reference _make_up_block (:56-64): ConvT(k2,s2)+BN+ReLU, Conv3+BN+ReLU."""

    def __init__(self, in_channels, out_channels):
        super().__init__(nn.ConvTranspose3d(in_channels, out_channels, 2, 2, bias=False),
                         nn.BatchNorm3d(out_channels), nn.ReLU(inplace=True),
                         nn.Conv3d(out_channels, out_channels, 3, 1, 1, bias=False),
                         nn.BatchNorm3d(out_channels), nn.ReLU(inplace=True))
        self._up = FusedConv(self[0], self[1], relu=True)
        self._conv = FusedConv(self[3], self[4], relu=True)

    def forward(self, x, skip):
        # the reference adds the encoder feature after the block's last ReLU (:31)
        return self._conv(self._up(x), residual=skip, residual_after_act=True)


class _Stack(nn.Sequential):
    def forward(self, x):
        for m in self:
            if not isinstance(m, (nn.Dropout, nn.ReLU, nn.BatchNorm3d)):
                x = m(x)
        return x


# ----------------------------------------------- necks -----------------------------------------------
@register(NECKS)
class FastIndoorImVoxelNeck(_NeckBase):
    """reference necks/imvoxelnet.py:8-67."""

    def __init__(self, in_channels, n_blocks, out_channels):
        super().__init__()
        self.n_scales = len(n_blocks)
        c = in_channels
        for i, n in enumerate(n_blocks):
            stride = 1 if i == 0 else 2
            blocks = []
            for k in range(n):
                if k == 0 and stride != 1:
                    blocks.append(BasicBlock3dV2(c, c * 2, stride))
                    c *= 2
                else:
                    blocks.append(BasicBlock3dV2(c, c))
            setattr(self, f'down_layer_{i}', _Stack(*blocks))
            if i > 0:
                setattr(self, f'up_block_{i}', _UpBlock(c, c // 2))
            setattr(self, f'out_block_{i}', _ConvBnRelu(nn.Conv3d(c, out_channels, 3, 1, 1, bias=False)))

    def forward(self, x):
        x = self._enter(x)
        skips = []
        for i in range(self.n_scales):
            x = getattr(self, f'down_layer_{i}')(x)
            skips.append(x)
        outs = []
        for i in reversed(range(self.n_scales)):
            if i < self.n_scales - 1:
                x = getattr(self, f'up_block_{i + 1}')(x, skips[i])
            outs.append(self._leave(getattr(self, f'out_block_{i}')(x)))
        return outs[::-1]


class _StridedStage(nn.Sequential):
    """Encoder stage of the Atlas net (:314-321): conv s2, BN, Dropout, ReLU, blocks."""

    def __init__(self, cin, cout, n_blocks):
        super().__init__(nn.Conv3d(cin, cout, 3, 2, 1, bias=False), nn.BatchNorm3d(cout),
                         nn.Dropout(0, True), nn.ReLU(inplace=True),
                         *[BasicBlock3d(cout, cout) for _ in range(n_blocks)])
        self._f = FusedConv(self[0], self[1], relu=True)

    def forward(self, x):
        x = self._f(x)
        for m in list(self)[4:]:
            x = m(x)
        return x


class ConditionalProjection(nn.Module):
    """reference :263-294: relu(bn(conv1x1(skip))) as one fused layer.  With ``condition=True``
    (no reference config sets it) the decoder tensor replaces the projected skip where the mask is
    off *before* the norm (:290-291), so the convolution runs alone and the select / affine / ReLU
    are element-wise torch ops on the device; inference only."""

    def __init__(self, n, condition=False):
        super().__init__()
        self.conv = nn.Conv3d(n, n, 1, bias=False)
        self.norm = nn.BatchNorm3d(n)
        self.relu = nn.ReLU(True)
        self.condition = bool(condition)
        self._f = FusedConv(self.conv, self.norm, relu=True)
        self._raw = FusedConv(self.conv, None, relu=False)

    def conditional_half_sum(self, skip, decoder, mask):
        """(decoder + relu(bn(where(mask, conv(skip), decoder)))) / 2 on [B,X,Y,Z,C] tensors (:289-294, :367)."""
        if self.norm.training and torch.is_grad_enabled():
            raise NotImplementedError("ConditionalProjection(condition=True) is supported in eval mode only")
        z = torch.where(mask, self._raw(skip).float(), decoder.float())
        bn = self.norm
        scale = bn.weight / torch.sqrt(bn.running_var + bn.eps)
        y = torch.relu(z * scale + (bn.bias - bn.running_mean * scale))
        return ((decoder.float() + y) / 2).to(decoder.dtype)


class EncoderDecoder(nn.Module):
    """reference :297-372 (norm='BN', drop=0, cond_proj=False, zero_init_residual=True)."""

    def __init__(self, channels, layers_down, layers_up, cond_proj=False):
        super().__init__()
        self.cond_proj = bool(cond_proj)
        self.layers_down = nn.ModuleList()
        proj = []
        self.layers_down.append(_Stack(*[BasicBlock3d(channels[0], channels[0])
                                         for _ in range(layers_down[0])]))
        proj.append(ConditionalProjection(channels[0], cond_proj))
        for i in range(1, len(channels)):
            self.layers_down.append(_StridedStage(channels[i - 1], channels[i], layers_down[i]))
            if i < len(channels) - 1:
                proj.append(ConditionalProjection(channels[i], cond_proj))
        self.proj = nn.ModuleList(proj[::-1])
        rev = channels[::-1]
        self.layers_up_conv = nn.ModuleList()
        self.layers_up_res = nn.ModuleList()
        for i in range(1, len(rev)):
            self.layers_up_conv.append(nn.Conv3d(rev[i - 1], rev[i], 1, bias=False))
            self.layers_up_res.append(_Stack(*[BasicBlock3d(rev[i], rev[i])
                                               for _ in range(layers_up[i - 1])]))
        self._up = [FusedConv(c) for c in self.layers_up_conv]
        for m in self.modules():
            if isinstance(m, BasicBlock3d):
                nn.init.constant_(m.bn2.weight, 0)

    def forward(self, x):
        if self.cond_proj:      # :347-348, on the channels-last tensor: [B,X,Y,Z,1]
            valid_mask = (x != 0).any(-1, keepdim=True)
        feats = []
        for layer in self.layers_down:
            x = layer(x)
            feats.append(x)
        feats = feats[::-1]
        outs = []
        for i in range(len(self.layers_up_conv)):
            if self.cond_proj:
                # :359-367 with the mask resampled as F.interpolate(scale_factor=1/k) does (nearest: index k*j)
                x = upsample2x_trilinear(self._up[i](x))
                k = 2 ** (len(self.layers_up_conv) - i - 1)
                x = self.proj[i].conditional_half_sum(feats[i + 1], x, valid_mask[:, ::k, ::k, ::k])
                x = self.layers_up_res[i](x)
                outs.append(x)
                continue
            # :359-360 computes conv1x1x1(upsample(x)); a bias-free pointwise conv and trilinear
            # interpolation (weights summing to 1) commute, so the conv runs on the 8x smaller
            # tensor and the up-sampling on the narrower one
            x = upsample2x_trilinear(self._up[i](x))
            # (x + relu(bn(conv(skip)))) / 2  (:366-367) in the projection's epilogue
            x = self.proj[i]._f(feats[i + 1], residual=x, residual_after_act=True, post_scale=0.5)
            x = self.layers_up_res[i](x)
            outs.append(x)
        return outs


@register(NECKS)
class ImVoxelNeck(_NeckBase):
    """reference necks/imvoxelnet.py:70-91."""

    def __init__(self, channels, out_channels, down_layers, up_layers, conditional):
        super().__init__()
        self.model = EncoderDecoder(list(channels), list(down_layers), list(up_layers), bool(conditional))
        self.conv_blocks = nn.ModuleList([
            _ConvBnRelu(nn.Conv3d(c, out_channels, 3, padding=1)) for c in channels])

    def forward(self, x):
        x = self.model(self._enter(x))[::-1]
        return [self._leave(self.conv_blocks[i](x[i])) for i in range(len(x))]


class _BevNeck(_NeckBase):
    def forward(self, x):
        x = self.model(self._enter(x))                 # [B, X', Y', Z', C]
        assert x.shape[3] == 1, "the voxel grid must collapse to one z slice (reference :119, :150)"
        return [self._leave(x)[..., 0].transpose(-1, -2)]


def _strided(cin, cout, stride=(1, 1, 2), padding=(1, 1, 1)):
    return _ConvBnRelu(nn.Conv3d(cin, cout, 3, stride=stride, padding=padding))


@register(NECKS)
class KittiImVoxelNeck(_BevNeck):
    """reference necks/imvoxelnet.py:94-123 (last conv keeps the reference's padding 0)."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        c = in_channels
        self.model = _Stack(BasicBlock3d(c, c), _strided(c, c * 2), BasicBlock3d(c * 2, c * 2),
                            _strided(c * 2, c * 4), BasicBlock3d(c * 4, c * 4),
                            _strided(c * 4, out_channels, 1, 0))


@register(NECKS)
class NuScenesImVoxelNeck(_BevNeck):
    """reference necks/imvoxelnet.py:126-154."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        c = in_channels
        self.model = _Stack(BasicBlock3d(c, c), _strided(c, c * 2, 2, 1), BasicBlock3d(c * 2, c * 2),
                            _strided(c * 2, c * 4), BasicBlock3d(c * 4, c * 4),
                            _strided(c * 4, out_channels, 1, (1, 1, 0)))
