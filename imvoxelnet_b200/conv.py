"""Fused convolution layers of the 3D necks on top of the C-ABI kernels.

A ``FusedConv`` wraps the parameter-holding ``nn.Conv3d`` /
``nn.ConvTranspose3d`` (+ ``nn.BatchNorm3d``) modules of a reference block
(mmdet3d/models/necks/imvoxelnet.py) and runs them as ONE kernel in eval mode:
BN folded into scale/shift, bias folded into shift, residual / ReLU / the
decoder's ``(x + y) / 2`` in the epilogue.  Activations are channels-last
``[B, X, Y, Z, C]`` tensors between layers.
"""
import torch
from torch import nn

from . import _lib

import contextlib

__all__ = ["FusedConv", "upsample2x_trilinear", "to_ndhwc", "from_ndhwc", "launch_count", "tf32", "fp32x3",
           "split_tf32"]

_launches = 0
#: bf16 layers with c_in, c_out multiples of 64 run on tcgen05; False forces the CUDA-core kernel
USE_TENSOR_CORES = True
_DT = {torch.float32: _lib.IVX_F32, torch.bfloat16: _lib.IVX_BF16}
_TF32 = False
#: fp32 layers with c_in % 32 == 0 and c_out % 64 == 0 run on tcgen05 as three tf32 products ("3xTF32",
#: ``ivx_conv3d_ndhwc_f32x3_tc``): fp32-grade results (1e-5).  False keeps them on the CUDA-core kernel.
_FP32X3 = True


@contextlib.contextmanager
def tf32(enabled=True):
    """Inside: fp32 layers with c_in % 32 == 0 and c_out % 64 == 0 run on the tensor cores with tf32
    math (``ivx_conv3d_ndhwc_tf32_tc``, about 1e-3 of the fp32 result), like
    ``torch.backends.cudnn.allow_tf32``.  The necks' ``precision = "tf32"`` switches it on."""
    global _TF32
    old, _TF32 = _TF32, bool(enabled)
    try:
        yield
    finally:
        _TF32 = old


@contextlib.contextmanager
def fp32x3(enabled=True):
    """Inside: whether eligible fp32 layers take the 3xTF32 tensor-core kernel (the default) or the
    CUDA-core kernel (``enabled=False``; the necks' ``precision = "fp32_simt"``)."""
    global _FP32X3
    old, _FP32X3 = _FP32X3, bool(enabled)
    try:
        yield
    finally:
        _FP32X3 = old


def _rna_tf32(t):
    """Round an fp32 tensor to tf32 (nearest, ties away from zero: PTX ``cvt.rna.tf32.f32``)."""
    return ((t.contiguous().view(torch.int32) + 0x1000) & -0x2000).view(torch.float32)


def split_tf32(x):
    """[..., C] fp32 -> [..., 2C]: (hi | lo) tf32 halves with hi + lo == x up to 2^-22 (``ivx_split_tf32_ndhwc``)."""
    global _launches
    assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()
    c = x.shape[-1]
    y = torch.empty(x.shape[:-1] + (2 * c,), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().ivx_split_tf32_ndhwc(x.data_ptr(), y.data_ptr(), x.numel() // c, c, _stream()),
                   "ivx_split_tf32_ndhwc")
    _launches += 1
    return y


def launch_count():
    return _launches


def _stream():
    return torch.cuda.current_stream().cuda_stream


def to_ndhwc(x):
    """Logical [B,C,X,Y,Z] tensor -> dense [B,X,Y,Z,C] (free for channels_last_3d input)."""
    if not x.is_cuda:
        raise _lib.IvxError("the 3D-neck kernels need CUDA tensors: there is no CPU implementation")
    return x.permute(0, 2, 3, 4, 1).contiguous()


def from_ndhwc(y):
    """Dense [B,X,Y,Z,C] -> logical [B,C,X,Y,Z] view (channels_last_3d strides)."""
    return y.permute(0, 4, 1, 2, 3)


def _triple(v):
    return tuple(v) if isinstance(v, (tuple, list)) else (v, v, v)


class FusedConv:
    """conv (+ BN) (+ ReLU) as one launch; parameters stay in the nn.Modules."""

    def __init__(self, conv, bn=None, relu=False):
        self.conv, self.bn, self.relu = conv, bn, relu
        self.transposed = isinstance(conv, nn.ConvTranspose3d)
        if not self.transposed:
            assert tuple(conv.dilation) == (1, 1, 1) and conv.groups == 1
        self._packed = None
        self._packed_tc = None
        self._packed_tf32 = None
        self._packed_x3 = None
        self._ws = {}

    # ---- parameter packing (cached until a parameter or buffer changes) ----
    def _key(self, device):
        tensors = [self.conv.weight, self.conv.bias]
        if self.bn is not None:
            tensors += [self.bn.weight, self.bn.bias, self.bn.running_mean, self.bn.running_var]
        return (str(device),) + tuple((t.data_ptr(), t._version) if t is not None else None
                                      for t in tensors)

    def packed(self, device):
        key = self._key(device)
        if self._packed is None or self._packed[0] != key:
            with torch.no_grad():
                w = self.conv.weight.detach().to(device=device, dtype=torch.float32)
                if self.transposed:       # [Cin, Cout, kx, ky, kz] -> [Cin][taps][Cout]
                    wp = w.permute(0, 2, 3, 4, 1).contiguous()
                else:                     # [Cout, Cin, kx, ky, kz] -> [taps][Cin][Cout]
                    wp = w.permute(2, 3, 4, 1, 0).contiguous()
                scale = shift = None
                bias = self.conv.bias
                if self.bn is not None:
                    bn = self.bn
                    scale = (bn.weight / torch.sqrt(bn.running_var + bn.eps)).to(device, torch.float32)
                    shift = (bn.bias - bn.running_mean * scale.to(bn.bias.device)).to(device, torch.float32)
                    if bias is not None:
                        shift = shift + bias.to(device, torch.float32) * scale
                    scale, shift = scale.contiguous(), shift.contiguous()
                elif bias is not None:
                    shift = bias.detach().to(device, torch.float32).contiguous()
            self._packed = (key, wp, scale, shift)
            self._packed_tc = self._packed_tf32 = self._packed_x3 = None
        return self._packed[1:]

    def packed_tc(self, device):
        """bf16 weights for the tensor-core kernel (+ fp32 scale/shift)."""
        _, scale, shift = self.packed(device)
        if self._packed_tc is None:
            with torch.no_grad():
                w = self.conv.weight.detach().to(device=device, dtype=torch.float32)
                if self.transposed:   # [Cin, Cout, kx, ky, kz] -> [taps][Cout][Cin]
                    self._packed_tc = w.permute(2, 3, 4, 1, 0).contiguous().to(torch.bfloat16)
                else:                 # [Cout, Cin, kx, ky, kz] -> [Cout][taps][Cin]
                    self._packed_tc = w.permute(0, 2, 3, 4, 1).contiguous().to(torch.bfloat16)
        return self._packed_tc, scale, shift

    def packed_tf32(self, device):
        """fp32 weights in the tensor-core layout (+ fp32 scale/shift) for the tf32 kernel."""
        _, scale, shift = self.packed(device)
        if self._packed_tf32 is None:
            with torch.no_grad():
                w = self.conv.weight.detach().to(device=device, dtype=torch.float32)
                # [taps][Cout][Cin] for transposed layers, [Cout][taps][Cin] otherwise
                perm = (2, 3, 4, 1, 0) if self.transposed else (0, 2, 3, 4, 1)
                self._packed_tf32 = w.permute(*perm).contiguous()
        return self._packed_tf32, scale, shift

    def packed_x3(self, device):
        """(hi | lo) tf32 halves of the weights for the 3xTF32 kernel: rows of the tensor-core layout
        ([c_out] x taps*c_in, transposed layers [taps*c_out] x c_in), each as [2][row]."""
        _, scale, shift = self.packed(device)
        if self._packed_x3 is None:
            with torch.no_grad():
                w = self.conv.weight.detach().to(device=device, dtype=torch.float32)
                if self.transposed:
                    rows = w.permute(2, 3, 4, 1, 0).reshape(-1, w.shape[0])
                else:
                    rows = w.permute(0, 2, 3, 4, 1).reshape(w.shape[0], -1)
                hi = _rna_tf32(rows)
                lo = _rna_tf32(rows - hi)
                self._packed_x3 = torch.stack([hi, lo], dim=1).contiguous()
        return self._packed_x3, scale, shift

    def _workspace(self, params, device):
        """Split-K scratch of this layer for this geometry: zero-filled once, the kernel
        leaves it zero-filled (include/ivx.h); None when the layer does not split."""
        key = (str(device), bytes(params))      # every field of the geometry decides the split
        ws = self._ws.get(key)
        if ws is None:
            nbytes = _lib.lib().ivx_conv3d_tc_workspace_bytes(params)
            ws = torch.zeros(nbytes, dtype=torch.uint8, device=device) if nbytes > 0 else False
            self._ws[key] = ws
        return ws if ws is not False else None

    def tc_eligible(self, x, out_dtype, residual=None):
        if self.transposed and residual is not None:
            return False
        return (x.dtype == torch.bfloat16 and out_dtype == torch.bfloat16
                and self.conv.in_channels % 64 == 0 and self.conv.out_channels % 64 == 0
                and USE_TENSOR_CORES)

    def tf32_eligible(self, x, out_dtype, residual=None):
        if self.transposed and residual is not None:
            return False
        return (_TF32 and x.dtype == torch.float32 and out_dtype == torch.float32
                and self.conv.in_channels % 32 == 0 and self.conv.out_channels % 64 == 0)

    def x3_eligible(self, x, out_dtype, residual=None):
        if self.transposed and residual is not None:
            return False
        return (_FP32X3 and not _TF32 and x.dtype == torch.float32 and out_dtype == torch.float32
                and self.conv.in_channels % 32 == 0 and self.conv.out_channels % 64 == 0)

    def out_dims(self, in_dims):
        k, s = _triple(self.conv.kernel_size), _triple(self.conv.stride)
        p = _triple(self.conv.padding)
        if self.transposed:
            return tuple(i * st for i, st in zip(in_dims, s))
        return tuple((i + 2 * pd - kk) // st + 1 for i, pd, kk, st in zip(in_dims, p, k, s))

    def make_params(self, b, in_dims, od, cin, cout, *, relu, after, post, in_dtype, out_dtype):
        """``ivx_conv3d_params`` of this layer for one input geometry."""
        p = _lib.ConvParams()
        p.batch, p.c_in, p.c_out = b, cin, cout
        k, s, pd = _triple(self.conv.kernel_size), _triple(self.conv.stride), _triple(self.conv.padding)
        for d in range(3):
            p.in_dims[d], p.out_dims[d] = in_dims[d], od[d]
            p.kernel[d], p.stride[d], p.pad[d] = k[d], s[d], pd[d]
        p.transposed, p.relu = int(self.transposed), int(relu)
        p.residual_after_act, p.post_scale = int(after), float(post)
        p.in_dtype, p.out_dtype = in_dtype, out_dtype
        return p

    def wants_autograd(self, x):
        """Training path: a gradient is wanted, or the norm is in ``.train()`` mode -- nn.BatchNorm3d
        then normalises with batch statistics and updates its running buffers whatever the grad
        mode (a validation loop that forgot ``.eval()``), so that case runs the training forward too."""
        if self.bn is not None and self.bn.training:
            return True
        return torch.is_grad_enabled() and (self.conv.training or x.requires_grad)

    def __call__(self, x, residual=None, residual_after_act=False, post_scale=1.0, out_dtype=None):
        """x: [B,X,Y,Z,Cin] dense CUDA tensor -> [B,X',Y',Z',Cout]."""
        global _launches
        if not x.is_cuda:
            raise _lib.IvxError("the 3D-neck kernels need CUDA tensors: there is no CPU implementation")
        if self.wants_autograd(x):
            from . import train
            return train.fused_layer_train(self, x, residual, residual_after_act, post_scale)
        assert x.is_contiguous() and x.dim() == 5
        b, in_dims, cin = x.shape[0], tuple(x.shape[1:4]), x.shape[4]
        cout = self.conv.out_channels
        assert cin == self.conv.in_channels, (cin, self.conv.in_channels)
        od = self.out_dims(in_dims)
        out_dtype = out_dtype or x.dtype
        y = torch.empty((b,) + od + (cout,), dtype=out_dtype, device=x.device)
        wp, scale, shift = self.packed(x.device)
        p = self.make_params(b, in_dims, od, cin, cout, relu=self.relu, after=residual_after_act,
                             post=post_scale, in_dtype=_DT[x.dtype], out_dtype=_DT[out_dtype])
        if residual is not None:
            assert residual.shape == y.shape and residual.dtype == y.dtype and residual.is_contiguous()
        res_ptr = residual.data_ptr() if residual is not None else None
        with torch.cuda.device(x.device):
            if self.tc_eligible(x, out_dtype, residual):
                wt, scale, shift = self.packed_tc(x.device)
                ws = self._workspace(p, x.device)
                _lib.check(_lib.lib().ivx_conv3d_ndhwc_bf16_tc(
                    p, x.data_ptr(), wt.data_ptr(), scale.data_ptr() if scale is not None else None,
                    shift.data_ptr() if shift is not None else None, res_ptr, y.data_ptr(),
                    ws.data_ptr() if ws is not None else None, ws.numel() if ws is not None else 0,
                    _stream()), "ivx_conv3d_ndhwc_bf16_tc")
            elif self.tf32_eligible(x, out_dtype, residual):
                wt, scale, shift = self.packed_tf32(x.device)
                ws = self._workspace(p, x.device)
                _lib.check(_lib.lib().ivx_conv3d_ndhwc_tf32_tc(
                    p, x.data_ptr(), wt.data_ptr(), scale.data_ptr() if scale is not None else None,
                    shift.data_ptr() if shift is not None else None, res_ptr, y.data_ptr(),
                    ws.data_ptr() if ws is not None else None, ws.numel() if ws is not None else 0,
                    _stream()), "ivx_conv3d_ndhwc_tf32_tc")
            elif self.x3_eligible(x, out_dtype, residual):
                wt, scale, shift = self.packed_x3(x.device)
                xs = split_tf32(x)
                p.in_dtype = _lib.IVX_F32X2
                ws = self._workspace(p, x.device)
                _lib.check(_lib.lib().ivx_conv3d_ndhwc_f32x3_tc(
                    p, xs.data_ptr(), wt.data_ptr(), scale.data_ptr() if scale is not None else None,
                    shift.data_ptr() if shift is not None else None, res_ptr, y.data_ptr(),
                    ws.data_ptr() if ws is not None else None, ws.numel() if ws is not None else 0,
                    _stream()), "ivx_conv3d_ndhwc_f32x3_tc")
            else:
                _lib.check(_lib.lib().ivx_conv3d_ndhwc(
                    p, x.data_ptr(), wp.data_ptr(), scale.data_ptr() if scale is not None else None,
                    shift.data_ptr() if shift is not None else None, res_ptr, y.data_ptr(), _stream()),
                    "ivx_conv3d_ndhwc")
        _launches += 1
        return y


def upsample2x_trilinear(x):
    """F.interpolate(scale_factor=2, mode='trilinear', align_corners=False) on [B,X,Y,Z,C]."""
    global _launches
    assert x.is_cuda and x.dim() == 5
    if torch.is_grad_enabled() and x.requires_grad:
        from . import train
        return train.upsample2x_train(x)
    assert x.is_contiguous()
    b, xs, ys, zs, c = x.shape
    y = torch.empty((b, 2 * xs, 2 * ys, 2 * zs, c), dtype=x.dtype, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().ivx_upsample2x_trilinear_ndhwc(
            x.data_ptr(), y.data_ptr(), b, c, _lib.i32([xs, ys, zs]), _DT[x.dtype], _stream()),
            "ivx_upsample2x_trilinear_ndhwc")
    _launches += 1
    return y
