"""Dense-head tail on the neck outputs.

Mirrors what the reference's ``ImVoxelHead`` / ``ImVoxelHeadV2`` families do
between the neck and top-k / NMS
(mmdet3d/models/dense_heads/imvoxel_head_v2.py, imvoxel_head.py):

* ``head_forward(head, xs)``   = ``head.forward(xs)``  (:57-58, ``forward_single`` :305-313, :444-449;
  v1 :326-336, :454-461): the final convolutions that share an input run as ONE
  fused launch with their output channels concatenated, then
  ``ivx_head_split_ndhwc`` writes the reference's ``[B,c,X,Y,Z]`` fp32 tensors
  (``exp(scale * distance)`` applied there);
* ``resize_valid(valid, size)`` = ``nn.Upsample(size, mode='trilinear')(valid).round().bool()`` (:91-93, :224-226);
* ``head_scores(...)``          = ``sigmoid(cls) * sigmoid(centerness) * valid`` and its per-voxel max (:262-270).

``head`` is any module with the reference's attribute names
(``centerness_conv``, ``reg_conv``, ``cls_conv``, ``scales`` and, for the v1
heads, ``reg_convs`` / ``cls_convs``); ``HeadConvsV2`` / ``HeadConvsV1`` are
parameter holders with exactly those names (same state-dict keys as the
reference heads' convolution part) for use without mmdet.  Losses, target
assignment, box decoding and NMS stay the reference's PyTorch code.
"""
import types

import torch
from torch import nn

from . import _lib
from .conv import FusedConv, to_ndhwc, _DT, _stream

__all__ = ["head_forward", "resize_valid", "head_scores", "supports", "accelerate", "HeadConvsV2", "HeadConvsV1",
           "launch_count"]

_launches = 0
N_DISTANCES = 6     # reg_final[:, :6] are the exp()-ed face distances, the rest (angle) stays raw


def launch_count():
    return _launches


def _call(name, *args):
    global _launches
    _lib.check(getattr(_lib.lib(), name)(*args, _stream()), name)
    _launches += 1


class _FusedFinal:
    """Several final Conv3d modules with one input, run as one layer with concatenated outputs."""

    def __init__(self, convs):
        self.convs = list(convs)
        first = self.convs[0]
        for c in self.convs:
            assert (c.in_channels, tuple(c.kernel_size), tuple(c.stride), tuple(c.padding)) == \
                   (first.in_channels, tuple(first.kernel_size), tuple(first.stride), tuple(first.padding))
        self.widths = [c.out_channels for c in self.convs]
        self.offsets = [sum(self.widths[:i]) for i in range(len(self.widths))]
        self._templates = {}
        self._key = None

    def _source_key(self):
        ts = [t for c in self.convs for t in (c.weight, c.bias) if t is not None]
        return tuple((t.data_ptr(), t._version, str(t.device)) for t in ts)

    def _cat(self, pad_to, device=None):
        """Concatenated (differentiable) weight and bias, zero-padded to a multiple of pad_to."""
        total = sum(self.widths)
        cpad = -(-total // pad_to) * pad_to
        ref = self.convs[0].weight
        ws = [c.weight for c in self.convs]
        bs = [c.bias if c.bias is not None else ref.new_zeros(c.out_channels) for c in self.convs]
        if cpad > total:
            ws.append(ref.new_zeros((cpad - total,) + tuple(ref.shape[1:])))
            bs.append(ref.new_zeros(cpad - total))
        return torch.cat(ws), torch.cat(bs), cpad

    def template(self, pad_to):
        """An un-registered nn.Conv3d holding the concatenated parameters (refreshed when they change)."""
        key = self._source_key()
        if key != self._key:
            self._templates, self._key = {}, key
        fc = self._templates.get(pad_to)
        if fc is None:
            first = self.convs[0]
            with torch.no_grad():
                w, b, cpad = self._cat(pad_to)
                conv = nn.Conv3d(first.in_channels, cpad, first.kernel_size, first.stride, first.padding,
                                 bias=True).to(w.device)
                conv.weight.copy_(w), conv.bias.copy_(b)
            conv.requires_grad_(False).eval()
            fc = self._templates[pad_to] = FusedConv(conv)
        return fc

    def __call__(self, x):
        """x [B,X,Y,Z,Cin] -> [B,X,Y,Z,Cpad] (bf16 in -> bf16 out on tcgen05, else fp32)."""
        needs_grad = torch.is_grad_enabled() and (x.requires_grad or any(c.weight.requires_grad and c.training
                                                                         for c in self.convs))
        if needs_grad:
            from . import train
            w, b, _ = self._cat(4)
            fc = self.template(4)
            return train._FusedLayer.apply(x.float().contiguous(), w, b, None, None, None, fc, False, 1.0)
        return self.template(64 if x.dtype == torch.bfloat16 else 4)(x)


def _plan(head):
    plan = head.__dict__.get("_ivx_head_plan")
    if plan is None:
        v1 = hasattr(head, "reg_convs")
        plan = {"v1": v1}
        if v1:
            plan["reg_stack"] = [FusedConv(seq[0], seq[1], relu=True) for seq in head.reg_convs]
            plan["cls_stack"] = [FusedConv(seq[0], seq[1], relu=True) for seq in head.cls_convs]
            plan["reg_final"] = _FusedFinal([head.centerness_conv, head.reg_conv])
            plan["cls_final"] = _FusedFinal([head.cls_conv])
        else:
            plan["final"] = _FusedFinal([head.centerness_conv, head.reg_conv, head.cls_conv])
        object.__setattr__(head, "_ivx_head_plan", plan)
    return plan


def _split(y, c_off, c_n, n_exp=0, scale=None):
    """Channels [c_off, c_off+c_n) of y [B,X,Y,Z,Ct] as a [B,c_n,X,Y,Z] fp32 tensor."""
    if y.requires_grad:                                   # training: plain differentiable torch ops
        v = y[..., c_off:c_off + c_n].permute(0, 4, 1, 2, 3)
        if n_exp:
            v = torch.cat((torch.exp(v[:, :n_exp] * scale), v[:, n_exp:]), dim=1)
        return v
    b, dims, ct = y.shape[0], tuple(y.shape[1:4]), y.shape[4]
    out = torch.empty((b, c_n) + dims, dtype=torch.float32, device=y.device)
    sp = None
    if n_exp:
        sp = scale.detach().to(device=y.device, dtype=torch.float32).reshape(1).contiguous()
    _call("ivx_head_split_ndhwc", y.data_ptr(), _DT[y.dtype], b, dims[0] * dims[1] * dims[2], ct, c_off, c_n,
          n_exp, sp.data_ptr() if sp is not None else None, out.data_ptr())
    return out


def _forward_single(head, plan, x, scale):
    n_reg, n_cls = head.reg_conv.out_channels, head.cls_conv.out_channels
    n_exp = min(N_DISTANCES, n_reg)
    x = to_ndhwc(x)
    if plan["v1"]:
        reg = cls = x
        for layer in plan["reg_stack"]:
            reg = layer(reg)
        for layer in plan["cls_stack"]:
            cls = layer(cls)
        y_reg, y_cls = plan["reg_final"](reg), plan["cls_final"](cls)
        return (_split(y_reg, 0, 1), _split(y_reg, 1, n_reg, n_exp, scale.scale), _split(y_cls, 0, n_cls))
    y = plan["final"](x)
    return (_split(y, 0, 1), _split(y, 1, n_reg, n_exp, scale.scale), _split(y, 1 + n_reg, n_cls))


def head_forward(head, xs):
    """``head.forward(xs)``: per-level lists (centernesses, bbox_preds, cls_scores), reference layout."""
    if not all(x.is_cuda for x in xs):
        raise _lib.IvxError("the head kernels need CUDA tensors: there is no CPU implementation")
    plan = _plan(head)
    with torch.cuda.device(xs[0].device):
        outs = [_forward_single(head, plan, x, s) for x, s in zip(xs, head.scales)]
    return tuple(map(list, zip(*outs)))


def supports(head):
    """Does ``head`` look like one of the reference's ImVoxelHead / ImVoxelHeadV2 classes?"""
    convs = [getattr(head, n, None) for n in ("centerness_conv", "reg_conv", "cls_conv")]
    if not all(isinstance(c, nn.Conv3d) for c in convs) or not hasattr(head, "scales"):
        return False
    if not all(tuple(c.kernel_size) == (3, 3, 3) and tuple(c.padding) == (1, 1, 1) and tuple(c.stride) == (1, 1, 1)
               for c in convs):
        return False
    return hasattr(head, "reg_convs") == hasattr(head, "cls_convs")


def accelerate(head):
    """Route ``head(xs)`` (and with it ``forward_train`` / ``get_bboxes`` callers) through ``head_forward``.
    Parameters, state-dict keys and every other method of the head stay as they are."""
    if not supports(head):
        raise TypeError(f"{type(head).__name__} does not have the ImVoxelHead convolution layout")
    object.__setattr__(head, "forward", types.MethodType(head_forward, head))   # survives copy.deepcopy
    return head


def resize_valid(valid, size):
    """``nn.Upsample(size=size, mode='trilinear')(valid).round().bool()`` for valid [B,1,X,Y,Z] (float or bool)."""
    if not valid.is_cuda:
        raise _lib.IvxError("resize_valid needs a CUDA tensor: there is no CPU implementation")
    assert valid.dim() == 5 and valid.shape[1] == 1
    if valid.dtype == torch.bool:
        src, dt = valid.contiguous().view(torch.uint8), _lib.IVX_U8
    else:
        src, dt = valid.float().contiguous(), _lib.IVX_F32
    size = tuple(int(s) for s in size)
    out = torch.empty((valid.shape[0], 1) + size, dtype=torch.bool, device=valid.device)
    with torch.cuda.device(valid.device):
        _call("ivx_resize_mask_trilinear", src.data_ptr(), dt, valid.shape[0], _lib.i32(list(valid.shape[2:])),
              _lib.i32(list(size)), out.data_ptr())
    return out


def head_scores(centerness, cls_score, valid):
    """scores [B,S,n] = sigmoid(cls) * sigmoid(centerness) * valid and their max over classes [B,S]
    from [B,1,X,Y,Z], [B,n,X,Y,Z], bool [B,1,X,Y,Z] (voxel order of ``permute(1,2,3,0).reshape(-1)``)."""
    if not centerness.is_cuda:
        raise _lib.IvxError("head_scores needs CUDA tensors: there is no CPU implementation")
    b, n = cls_score.shape[0], cls_score.shape[1]
    spatial = cls_score[0, 0].numel()
    assert centerness.shape == (b, 1) + tuple(cls_score.shape[2:]) == valid.shape and valid.dtype == torch.bool
    ctr, cls, ok = centerness.float().contiguous(), cls_score.float().contiguous(), valid.contiguous()
    scores = torch.empty((b, spatial, n), dtype=torch.float32, device=cls.device)
    best = torch.empty((b, spatial), dtype=torch.float32, device=cls.device)
    with torch.cuda.device(cls.device):
        _call("ivx_head_scores", ctr.data_ptr(), cls.data_ptr(), ok.data_ptr(), b, spatial, n, scores.data_ptr(),
              best.data_ptr())
    return scores, best


class _Scale(nn.Module):
    """mmcv.cnn.Scale: a learnable scalar factor (state-dict key ``scale``)."""

    def __init__(self, scale=1.0):
        super().__init__()
        self.scale = nn.Parameter(torch.tensor(scale, dtype=torch.float))

    def forward(self, x):
        return x * self.scale


class HeadConvsV2(nn.Module):
    """Convolution part of ``ImVoxelHeadV2`` (``_init_layers``, imvoxel_head_v2.py:45-49)."""

    def __init__(self, n_classes, n_channels, n_reg_outs, n_scales):
        super().__init__()
        self.centerness_conv = nn.Conv3d(n_channels, 1, 3, padding=1, bias=False)
        self.reg_conv = nn.Conv3d(n_channels, n_reg_outs, 3, padding=1, bias=False)
        self.cls_conv = nn.Conv3d(n_channels, n_classes, 3, padding=1)
        self.scales = nn.ModuleList([_Scale(1.) for _ in range(n_scales)])

    def forward(self, xs):
        return head_forward(self, xs)


class HeadConvsV1(nn.Module):
    """Convolution part of ``ImVoxelHead`` (``_init_layers``, imvoxel_head.py:46-63)."""

    def __init__(self, n_classes, n_channels, n_convs, n_reg_outs, n_scales=3):
        super().__init__()

        def stack():
            return nn.Sequential(*[nn.Sequential(nn.Conv3d(n_channels, n_channels, 3, padding=1, bias=False),
                                                 nn.BatchNorm3d(n_channels), nn.ReLU(inplace=True))
                                   for _ in range(n_convs)])

        self.reg_convs, self.cls_convs = stack(), stack()
        self.centerness_conv = nn.Conv3d(n_channels, 1, 3, padding=1, bias=False)
        self.reg_conv = nn.Conv3d(n_channels, n_reg_outs, 3, padding=1, bias=False)
        self.cls_conv = nn.Conv3d(n_channels, n_classes, 3, padding=1)
        self.scales = nn.ModuleList([_Scale(1.) for _ in range(n_scales)])

    def forward(self, xs):
        return head_forward(self, xs)
