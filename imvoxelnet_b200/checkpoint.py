"""Checkpoint transform: fold BatchNorm3d into the preceding 3D convolution.

The reference's ``tools/fuse_conv_bn.py`` (:9-43) folds ``Conv2d`` + ``BatchNorm2d``
pairs of a loaded detector and saves the result; its walker skips every 3D
layer, i.e. the whole ``neck_3d`` and the v1 heads' conv stacks.  This module
does the same transform for ``Conv3d`` / ``ConvTranspose3d`` + ``BatchNorm3d``:

* ``fold_conv_bn3d(module)`` rewrites the pairs in place with the reference
  tool's convention: the convolution gets the folded weight and a bias,
  the norm becomes ``nn.Identity`` (so the folded state dict has no BN keys);
* it works on any module tree whose children are registered conv-then-norm
  (the reference's necks, the oracle's, this package's) and keeps this
  package's fused layers in step (their BN reference is dropped, packed
  weights are rebuilt on the next call);
* ``fold_state_dict(module_factory, state_dict)`` is the file-level transform
  behind ``tools/fuse_conv_bn3d.py``.

Folding is exact up to one fp32 rounding of ``w * gamma / sqrt(var + eps)``;
the kernels apply the same affine in their epilogue otherwise, so outputs agree
to ~1e-6 (checked in tests/test_checkpoint.py and tests/test_head_gpu.py).
"""
import torch
from torch import nn

_CONVS = (nn.Conv3d, nn.ConvTranspose3d)


def fold_pair(conv, bn):
    """Fold eval-mode ``bn`` into ``conv`` in place (fuse_conv_bn, tools/fuse_conv_bn.py:9-22, for 3D)."""
    with torch.no_grad():
        factor = bn.weight / torch.sqrt(bn.running_var + bn.eps)
        bias = conv.bias if conv.bias is not None else torch.zeros_like(bn.running_mean)
        if isinstance(conv, nn.ConvTranspose3d):      # weight [Cin, Cout, k, k, k]
            weight = conv.weight * factor.reshape(1, -1, 1, 1, 1)
        else:                                         # weight [Cout, Cin, k, k, k]
            weight = conv.weight * factor.reshape(-1, 1, 1, 1, 1)
        conv.weight = nn.Parameter(weight)
        conv.bias = nn.Parameter((bias - bn.running_mean) * factor + bn.bias)
    return conv


def _fused_layers(root):
    """Every conv.FusedConv held (directly or in a list) by a module of the tree."""
    from .conv import FusedConv
    for m in root.modules():
        for v in vars(m).values():
            if isinstance(v, FusedConv):
                yield v
            elif isinstance(v, (list, tuple)):
                for item in v:
                    if isinstance(item, FusedConv):
                        yield item


def fold_conv_bn3d(module):
    """Fold every Conv3d/ConvTranspose3d + BatchNorm3d pair below ``module``; returns the pair count."""
    folded = {}

    def walk(m):
        last, n = None, 0
        if getattr(m, "condition", False):
            # ConditionalProjection(condition=True) (necks/imvoxelnet.py:289-291) selects between the
            # convolution and the decoder tensor BEFORE the norm: folding would rescale one branch only
            raise ValueError("fold_conv_bn3d: a module with condition=True applies its norm to a select of "
                             "two tensors, not to its convolution; it cannot be folded")
        for name, child in list(m.named_children()):
            if isinstance(child, nn.BatchNorm3d):
                if last is not None:                  # only a norm that directly follows a convolution
                    fold_pair(last, child)
                    folded[id(child)] = True
                    m._modules[name] = nn.Identity()
                    n += 1
                last = None
            elif isinstance(child, _CONVS):
                last = child
            elif isinstance(child, (nn.Dropout, nn.Dropout3d)) and child.p == 0:
                pass                                  # BasicBlock3d keeps its p=0 dropouts between layers
            else:
                last = None
                n += walk(child)
        return n

    count = walk(module)
    for fc in _fused_layers(module):
        if fc.bn is not None and id(fc.bn) in folded:
            fc.bn = None
            fc._packed = fc._packed_tc = fc._packed_tf32 = fc._packed_x3 = None
    return count


def fold_state_dict(module_factory, state_dict, prefix=""):
    """Folded counterpart of the ``prefix`` part of a checkpoint's state dict.

    ``module_factory()`` builds the (unfolded) module the keys belong to, e.g.
    ``lambda: necks.KittiImVoxelNeck(64, 256)`` with ``prefix='neck_3d.'``.  Keys outside the
    prefix are passed through untouched."""
    module = module_factory().eval()
    own = {k[len(prefix):]: v for k, v in state_dict.items() if k.startswith(prefix)}
    module.load_state_dict(own)
    fold_conv_bn3d(module)
    out = {k: v for k, v in state_dict.items() if not k.startswith(prefix)}
    out.update({prefix + k: v.detach().clone() for k, v in module.state_dict().items()})
    return out
