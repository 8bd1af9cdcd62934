"""Training mode of the fused neck layers: batch-statistics BatchNorm and autograd.

The reference trains its necks with ordinary PyTorch autograd
(``forward_train``, mmdet3d/models/detectors/imvoxelnet.py:82-87; BatchNorm3d
in ``.train()`` mode normalises with the batch moments and updates the running
statistics).  Here every fused layer of ``conv.FusedConv`` becomes one
``torch.autograd.Function`` whose forward and backward are made of the fp32
CUDA kernels declared under "training mode" in include/ivx.h; only vectors of
length C (BN coefficients) are handled with torch ops.
"""
import ctypes

import torch
from torch import nn

from . import _lib

_F32 = _lib.IVX_F32
_launches = 0


def launch_count():
    return _launches


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return t.data_ptr() if t is not None else None


PROFILE = None      # set to a dict to collect {entry point: [calls, device ms]} (synchronises every call)


def _call(name, *args):
    global _launches
    if PROFILE is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    _lib.check(getattr(_lib.lib(), name)(*args, _stream()), name)
    _launches += 1
    if PROFILE is not None:
        e1.record()
        torch.cuda.synchronize()
        rec = PROFILE.setdefault(name, [0, 0.0])
        rec[0] += 1
        rec[1] += e0.elapsed_time(e1)


def _moments(a, b, rows, c):
    """(sum a, sum a*b) per channel, accumulated in double."""
    acc = torch.zeros(2 * c, dtype=torch.float64, device=a.device)
    _call("ivx_channel_moments_ndhwc", a.data_ptr(), b.data_ptr(), rows, c, acc.data_ptr())
    return acc[:c], acc[c:]


def _tf32_ok(c_in, c_out):
    from . import conv
    return conv._TF32 and c_in % 32 == 0 and c_out % 64 == 0


def _x3_ok(c_in, c_out):
    """fp32 training: convolutions of eligible layers (forward, stride-1 and transposed data gradients) as
    three tf32 products on the tensor cores (3xTF32, 1e-5 of fp32), like the inference path."""
    from . import conv
    return conv._FP32X3 and not conv._TF32 and c_in % 32 == 0 and c_out % 64 == 0


def _tc_ok(c_in, c_out):
    return _tf32_ok(c_in, c_out) or _x3_ok(c_in, c_out)


def _conv_tc(fc, params, x, w_tc, bias, y):
    """y = conv(x) + bias on the tensor cores: tf32 math under precision "tf32", else 3xTF32 (fp32 grade)."""
    from . import conv
    if conv._TF32:
        return _conv_tf32(fc, params, x, w_tc, bias, y)
    q = _copy_params(params, in_dtype=_lib.IVX_F32X2)
    rows = w_tc.reshape(-1, w_tc.shape[-1]) if params.transposed else w_tc.reshape(w_tc.shape[0], -1)
    hi = conv._rna_tf32(rows)
    w_split = torch.stack([hi, conv._rna_tf32(rows - hi)], dim=1).contiguous()
    xs = conv.split_tf32(x.contiguous())
    ws = fc._workspace(q, x.device)
    _call("ivx_conv3d_ndhwc_f32x3_tc", q, xs.data_ptr(), w_split.data_ptr(), None, _ptr(bias), None, y.data_ptr(),
          ws.data_ptr() if ws is not None else None, ws.numel() if ws is not None else 0)


def _conv_tf32(fc, params, x, w_tc, bias, y):
    """y = conv(x) + bias on the tensor cores with tf32 math (weights in the tensor-core layout)."""
    ws = fc._workspace(params, x.device)
    _call("ivx_conv3d_ndhwc_tf32_tc", params, x.data_ptr(), w_tc.data_ptr(), None, _ptr(bias), None, y.data_ptr(),
          ws.data_ptr() if ws is not None else None, ws.numel() if ws is not None else 0)


def _channel_major(t, axes, shifts=None, out_len=None, step=1, halo=0):
    """[B, X, Y, Z, C] -> [n, B, D2 + 2 halo, D1, runs, C, 32]: the tiled channel-major operand layout of
    ivx_conv3d_wgrad_tf32_tc (D_i = axis axes[i]; runs of 32 positions along D0).  Copy j holds
    ``t[.., o * step + shifts[j]]`` at ``o < out_len`` (zero outside the volume and in the halo planes);
    default: one plain copy."""
    assert t.is_contiguous() and t.dtype == torch.float32
    b, dims, c = t.shape[0], tuple(t.shape[1:4]), t.shape[4]
    shifts = shifts or [0]
    out_len = out_len or dims[axes[0]]
    d1, d2 = dims[axes[1]], dims[axes[2]]
    out = torch.empty((len(shifts), b, d2 + 2 * halo, d1, -(-out_len // 32), c, 32), dtype=torch.float32,
                      device=t.device)
    _call("ivx_channel_major_ndhwc", t.data_ptr(), b, _lib.i32(dims), c, _lib.i32(axes), len(shifts), shifts[0],
          step, out_len, halo, out.data_ptr())
    return out


def _wgrad_tf32(p, x, dz):
    """Weight gradient on the tensor cores in torch's layout, or None when the layer is not one it takes."""
    from . import conv
    x3 = not conv._TF32 and conv._FP32X3        # fp32 training: three tf32 products over split operands
    if not (conv._TF32 or x3) or p.c_in % 32 or p.c_out % 32:
        return None
    if p.transposed:
        # y[s*i + t] += x[i] * w[ci, co, t]: the plain convolution from the OUTPUT grid (channels c_out) to
        # the input grid (channels c_in) with kernel == stride, whose "input" is grad_z and whose "output
        # gradient" is x
        q = _copy_params(p, c_in=p.c_out, c_out=p.c_in, in_dims=tuple(p.out_dims), out_dims=tuple(p.in_dims),
                         transposed=0)
        src, grad = dz, x
    else:
        q, src, grad = p, x, dz
    axes = _lib.i32([0, 0, 0])
    if _lib.lib().ivx_conv3d_wgrad_tf32_axes(q, axes) != 0:
        return None
    axes = list(axes)
    a0 = axes[0]        # innermost axis: its tap shift and stride are baked into one copy of x per tap position
    if PROFILE is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    if x3:      # (hi | lo) halves of both operands: 2 C channels each from here on
        src, grad = conv.split_tf32(src.contiguous()), conv.split_tf32(grad.contiguous())
    x_t = _channel_major(src, axes, shifts=[j - q.pad[a0] for j in range(q.kernel[a0])],
                         out_len=q.out_dims[a0], step=q.stride[a0], halo=q.pad[axes[2]])
    dz_t = _channel_major(grad, axes)
    if PROFILE is not None:
        e1.record()
        torch.cuda.synchronize()
        rec = PROFILE.setdefault("(both channel-major copies of a layer)", [0, 0.0])
        rec[0] += 1
        rec[1] += e0.elapsed_time(e1)
    k = tuple(q.kernel)
    dw = torch.zeros((k[0], k[1], k[2], q.c_out, q.c_in), dtype=torch.float32, device=x.device)
    _call("ivx_conv3d_wgrad_f32x3_tc" if x3 else "ivx_conv3d_wgrad_tf32_tc", q, x_t.data_ptr(), dz_t.data_ptr(),
          dw.data_ptr())
    if p.transposed:
        return dw.permute(3, 4, 0, 1, 2)    # [k][k][k][Cin_T][Cout_T] -> torch's [Cin_T, Cout_T, kX, kY, kZ]
    return dw.permute(3, 4, 0, 1, 2)        # torch's [Cout, Cin, kX, kY, kZ]


def _dgrad_strided_tf32(fc, p, dz, wp, dx):
    """Data gradient of a strided convolution on the tensor cores, phase by phase.

    Input voxels i = s*j + r (phase r, per axis) only receive the taps t with (r + pad - t) % s == 0, from the
    outputs o = j + (r + pad - t) / s: a stride-1 convolution of ``dz`` with that subset of the kernel,
    written into the strided view dx[r::s].  ``wp`` is the layer's weight as [kX][kY][kZ][Cin][Cout]."""
    s3, k3, p3 = list(p.stride), list(p.kernel), list(p.pad)
    in_dims, out_dims = tuple(p.in_dims), tuple(p.out_dims)
    cin, cout = p.c_in, p.c_out
    # per axis and phase: the taps (in the order of increasing o - j) and the low-side padding
    plans = []
    for d in range(3):
        per_phase = []
        for r in range(s3[d]):
            taps = [t for t in range(k3[d]) if (r + p3[d] - t) % s3[d] == 0]
            if not taps:
                per_phase.append(None)
                continue
            e = {t: (r + p3[d] - t) // s3[d] for t in taps}
            e_min, e_max = min(e.values()), max(e.values())
            by_u = {e[t] - e_min: t for t in taps}
            per_phase.append(([by_u.get(u) for u in range(e_max - e_min + 1)], -e_min))
        plans.append(per_phase)
    if any(ph is not None and ph[1] < 0 for per in plans for ph in per):
        return False
    sx, sy, sz = in_dims[1] * in_dims[2] * cin, in_dims[2] * cin, cin      # element strides of dx [B,X,Y,Z,Cin]
    sb = in_dims[0] * sx
    full = all(ph is not None for per in plans for ph in per)
    if not full:
        dx.zero_()                  # phases no tap reaches keep a zero gradient
    from . import conv as _conv
    x3 = not _conv._TF32            # fp32 training: the same phases with three tf32 products (3xTF32)
    dz_in = _conv.split_tf32(dz.contiguous()) if x3 else dz
    n_taps = k3[0] * k3[1] * k3[2]
    # the layer's taps as rows, plus one zero row for the holes of a phase's sub-kernel
    taps_ext = torch.cat((wp.reshape(n_taps, cin, cout), wp.new_zeros((1, cin, cout))))
    cache = fc.__dict__.setdefault("_phase_index", {})
    for r0, ph0 in enumerate(plans[0]):
        for r1, ph1 in enumerate(plans[1]):
            for r2, ph2 in enumerate(plans[2]):
                if ph0 is None or ph1 is None or ph2 is None:
                    continue
                view_dims = tuple(-(-(in_dims[d] - r) // s3[d]) for d, r in enumerate((r0, r1, r2)))
                if min(view_dims) <= 0:
                    continue
                sel = [ph0[0], ph1[0], ph2[0]]
                key = (str(dz.device), tuple(k3), tuple(s3), tuple(p3), r0, r1, r2)
                idx = cache.get(key)
                if idx is None:     # which tap of the layer sits at (u0, u1, u2) of this phase's sub-kernel
                    rows = [(t0 * k3[1] + t1) * k3[2] + t2 if None not in (t0, t1, t2) else n_taps
                            for t0 in sel[0] for t1 in sel[1] for t2 in sel[2]]
                    idx = cache[key] = torch.tensor(rows, dtype=torch.long, device=dz.device)
                w = taps_ext.index_select(0, idx).permute(1, 0, 2).contiguous()    # [Cin][taps'][Cout]
                q = _copy_params(p, c_in=cout, c_out=cin, in_dims=out_dims, out_dims=view_dims,
                                 kernel=tuple(len(x) for x in sel), stride=(1, 1, 1), pad=(ph0[1], ph1[1], ph2[1]),
                                 relu=0, residual_after_act=0, post_scale=1.0)
                entry = "ivx_conv3d_ndhwc_tf32_tc_view"
                if x3:
                    q.in_dtype = _lib.IVX_F32X2
                    rows = w.reshape(cin, -1)
                    hi = _conv._rna_tf32(rows)
                    w = torch.stack([hi, _conv._rna_tf32(rows - hi)], dim=1).contiguous()
                    entry = "ivx_conv3d_ndhwc_f32x3_tc_view"
                ws = fc._workspace(q, dz.device)
                base = dx.data_ptr() + 4 * (r0 * sx + r1 * sy + r2 * sz)
                _call(entry, q, dz_in.data_ptr(), w.data_ptr(), None, None, base,
                      _lib.i64([s3[0] * sx, s3[1] * sy, s3[2] * sz, sb]),
                      ws.data_ptr() if ws is not None else None, ws.numel() if ws is not None else 0)
    return True


def _copy_params(p, **changes):
    q = _lib.ConvParams()
    ctypes.memmove(ctypes.byref(q), ctypes.byref(p), ctypes.sizeof(q))
    for k, v in changes.items():
        if isinstance(v, (tuple, list)):
            for d in range(3):
                getattr(q, k)[d] = v[d]
        else:
            setattr(q, k, v)
    return q


class _FusedLayer(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, gamma, beta, residual, fc, after, post):
        conv, bn = fc.conv, fc.bn
        dev = x.device
        b, in_dims, cin = x.shape[0], tuple(x.shape[1:4]), x.shape[4]
        cout = conv.out_channels
        od = fc.out_dims(in_dims)
        p = fc.make_params(b, in_dims, od, cin, cout, relu=0, after=0, post=1.0, in_dtype=_F32, out_dtype=_F32)
        w = weight.detach().float()
        wp = (w.permute(0, 2, 3, 4, 1) if fc.transposed else w.permute(2, 3, 4, 1, 0)).contiguous()
        z = torch.empty((b,) + od + (cout,), dtype=torch.float32, device=dev)
        bias_f = bias.detach().float().contiguous() if bias is not None else None
        if _tc_ok(cin, cout):        # the convolution itself on the tensor cores ("tf32": tf32 math; "fp32": 3xTF32)
            w_tc = (w.permute(2, 3, 4, 1, 0) if fc.transposed else w.permute(0, 2, 3, 4, 1)).contiguous()
            _conv_tc(fc, p, x, w_tc, bias_f, z)
        else:
            _call("ivx_conv3d_ndhwc", p, x.data_ptr(), wp.data_ptr(), None, _ptr(bias_f), None, z.data_ptr())
        rows = z.numel() // cout
        sc = sh = mean = invstd = None
        batch_stats = bn is not None and bn.training
        if bn is not None:
            if batch_stats:
                s1, s2 = _moments(z, z, rows, cout)
                mean64 = s1 / rows
                var64 = (s2 / rows - mean64 * mean64).clamp_(min=0)
                mean, invstd = mean64.float(), torch.rsqrt(var64 + bn.eps).float()
                if bn.track_running_stats and bn.running_mean is not None:
                    with torch.no_grad():
                        bn.num_batches_tracked += 1
                        m = bn.momentum if bn.momentum is not None else 1.0 / float(bn.num_batches_tracked)
                        bn.running_mean.mul_(1 - m).add_(mean, alpha=m)
                        bn.running_var.mul_(1 - m).add_((var64 * (rows / max(rows - 1, 1))).float(), alpha=m)
            else:
                mean = bn.running_mean.detach().float()
                invstd = torch.rsqrt(bn.running_var.detach().float() + bn.eps)
            sc = (gamma.detach().float() * invstd).contiguous()
            sh = (beta.detach().float() - mean * sc).contiguous()
        trivial = bn is None and not fc.relu and residual is None and post == 1.0
        if trivial:
            y = z
        else:
            y = torch.empty_like(z)
            _call("ivx_affine_act_ndhwc", z.data_ptr(), _ptr(sc), _ptr(sh), _ptr(residual), int(fc.relu),
                  int(after), float(post), rows, cout, y.data_ptr())
        from . import conv as _conv
        ctx.tf32 = _conv._TF32       # the backward pass runs outside the neck's tf32 context: remember it
        ctx.x3 = _conv._FP32X3
        ctx.fc, ctx.p, ctx.flags = fc, p, (after, post, trivial, batch_stats, rows)
        ctx.has_bias = bias is not None
        keep_res = residual if (residual is not None and fc.relu and not after) else None
        ctx.has_res = residual is not None
        ctx.save_for_backward(x, wp, z, sc, sh, mean, invstd, keep_res)
        return y

    @staticmethod
    def backward(ctx, dy):
        from . import conv as _conv
        with _conv.tf32(ctx.tf32), _conv.fp32x3(ctx.x3):
            return _FusedLayer._backward(ctx, dy)

    @staticmethod
    def _backward(ctx, dy):
        fc, p = ctx.fc, ctx.p
        after, post, trivial, batch_stats, rows = ctx.flags
        x, wp, z, sc, sh, mean, invstd, res = ctx.saved_tensors
        cout, cin = p.c_out, p.c_in
        dy = dy.contiguous().float()
        dres = None
        if trivial:
            g = dy
        else:
            g = torch.empty_like(z)
            g_after = torch.empty_like(z) if (ctx.has_res and after) else None
            _call("ivx_affine_act_backward_ndhwc", dy.data_ptr(), z.data_ptr(), _ptr(sc), _ptr(sh), _ptr(res),
                  int(fc.relu), int(after), float(post), rows, cout, g.data_ptr(), _ptr(g_after))
            if ctx.has_res:
                dres = g_after if after else g
        dgamma = dbeta = None
        if fc.bn is not None:
            s1, s2 = _moments(g, z, rows, cout)                    # sum g, sum g*z
            inv64, mean64 = invstd.double(), mean.double()
            s2h = inv64 * (s2 - mean64 * s1)                       # sum g * x_hat
            dbeta, dgamma = s1.float(), s2h.float()
            if batch_stats:
                sc64 = sc.double()
                A = sc
                B = (-sc64 * inv64 * s2h / rows).float()
                C = (-sc64 * s1 / rows + sc64 * inv64 * mean64 * s2h / rows).float()
            else:                                                   # frozen statistics: dz = sc * g
                A, B, C = sc, torch.zeros_like(sc), torch.zeros_like(sc)
            dz = torch.empty_like(z)
            _call("ivx_channel_lincomb_ndhwc", g.data_ptr(), z.data_ptr(), A.contiguous().data_ptr(),
                  B.contiguous().data_ptr(), C.contiguous().data_ptr(), rows, cout, dz.data_ptr())
        else:
            dz = g
        dbias = None
        if ctx.has_bias:
            dbias = _moments(dz, dz, rows, cout)[0].float()
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            unit_stride = all(p.stride[d] == 1 for d in range(3))
            if fc.transposed:    # data gradient of a kernel == stride transposed conv: a plain strided conv of dz
                q = _copy_params(p, c_in=cout, c_out=cin, in_dims=tuple(p.out_dims), out_dims=tuple(p.in_dims),
                                 transposed=0, relu=0, residual_after_act=0, post_scale=1.0)
                if _tc_ok(cout, cin):
                    _conv_tc(fc, q, dz, wp.contiguous(), None, dx)   # wp is [Cin][k][k][k][Cout] already
                else:
                    wt = wp.permute(1, 2, 3, 4, 0).contiguous()    # [Cin,k,k,k,Cout] -> [taps][Cout][Cin]
                    _call("ivx_conv3d_ndhwc", q, dz.data_ptr(), wt.data_ptr(), None, None, None, dx.data_ptr())
            elif unit_stride and _tc_ok(cout, cin):
                # stride 1: the data gradient is the convolution of dz with the flipped, transposed kernel
                q = _copy_params(p, c_in=cout, c_out=cin, in_dims=tuple(p.out_dims), out_dims=tuple(p.in_dims),
                                 pad=tuple(p.kernel[d] - 1 - p.pad[d] for d in range(3)), relu=0,
                                 residual_after_act=0, post_scale=1.0)
                w_flip = wp.flip(0, 1, 2).permute(3, 0, 1, 2, 4).contiguous()   # [k][k][k][Cin][Cout] -> [Cin][k'][Cout]
                _conv_tc(fc, q, dz, w_flip, None, dx)
            elif _tc_ok(cout, cin) and _dgrad_strided_tf32(fc, p, dz, wp, dx):
                pass                                               # strided layer: one tensor-core launch per phase
            else:
                wt = wp.permute(0, 1, 2, 4, 3).contiguous()        # [taps][Cin][Cout] -> [taps][Cout][Cin]
                _call("ivx_conv3d_dgrad_ndhwc", p, dz.data_ptr(), wt.data_ptr(), dx.data_ptr())
        dw = _wgrad_tf32(p, x, dz)
        if dw is None:
            dwp = torch.zeros_like(wp)
            _call("ivx_conv3d_wgrad_ndhwc", p, x.data_ptr(), dz.data_ptr(), dwp.data_ptr())
            dw = dwp.permute(0, 4, 1, 2, 3) if fc.transposed else dwp.permute(4, 3, 0, 1, 2)
        return dx, dw.contiguous(), dbias, dgamma, dbeta, dres, None, None, None


def fused_layer_train(fc, x, residual=None, residual_after_act=False, post_scale=1.0):
    """Differentiable forward of a ``FusedConv`` (fp32 channels-last tensors)."""
    if x.dtype != torch.float32:
        raise NotImplementedError("training keeps fp32 tensors: set precision='fp32' or 'tf32' on the neck")
    bn = fc.bn
    return _FusedLayer.apply(x.contiguous(), fc.conv.weight, fc.conv.bias,
                             bn.weight if bn is not None else None, bn.bias if bn is not None else None,
                             residual.contiguous() if residual is not None else None, fc,
                             bool(residual_after_act), float(post_scale))


class _Upsample2x(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        b, xs, ys, zs, c = x.shape
        y = torch.empty((b, 2 * xs, 2 * ys, 2 * zs, c), dtype=x.dtype, device=x.device)
        _call("ivx_upsample2x_trilinear_ndhwc", x.data_ptr(), y.data_ptr(), b, c, _lib.i32([xs, ys, zs]), _F32)
        ctx.shape = tuple(x.shape)
        return y

    @staticmethod
    def backward(ctx, dy):
        b, xs, ys, zs, c = ctx.shape
        dx = torch.zeros(ctx.shape, dtype=torch.float32, device=dy.device)
        _call("ivx_upsample2x_trilinear_backward_ndhwc", dy.contiguous().data_ptr(), dx.data_ptr(), b, c,
              _lib.i32([xs, ys, zs]))
        return dx


def upsample2x_train(x):
    return _Upsample2x.apply(x.contiguous())
