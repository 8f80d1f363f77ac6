"""Hand-written sm_100a compute ops (K5-K9): thin autograd wrappers over ``libewdml.so``.

Activations are bf16 in NHWC (torch ``channels_last``) order, weights are read from the bf16 shadow
copy that the fused optimizer kernel maintains next to the fp32 master parameters, and weight
gradients are accumulated by the kernels straight into the flat fp32 gradient buffer.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from ..parallel import overlap as _overlap
from ..runtime import lib as L

_lib = None


def _l():
    global _lib
    if _lib is None:
        _lib = L.load()
        vp, i32, f32 = C.c_void_p, C.c_int, C.c_float
        i64, u32 = C.c_int64, C.c_uint32
        sig = {
            "ewdml_gemm_bf16": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp, f32, i32, vp],
            "ewdml_im2col": [vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp],
            "ewdml_conv3x3_igemm": [i32, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, i32, vp],
            "ewdml_pad_nhwc": [vp, vp, i32, i32, i32, i32, vp],
            "ewdml_conv_small_fwd": [vp, i32, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp],
            "ewdml_conv_small_wgrad": [vp, i32, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp],
            "ewdml_col2im": [vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp],
            "ewdml_conv_direct_fwd": [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp],
            "ewdml_conv_direct_dgrad": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp],
            "ewdml_conv_direct_wgrad": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp],
            "ewdml_colstats": [vp, vp, i64, i32, i32, vp],
            "ewdml_bn_finalize": [vp, vp, vp, vp, vp, vp, i32, i64, f32, f32, vp],
            "ewdml_bn_finalize2": [vp, vp, vp, vp, vp, vp, i32, i64, f32, f32, vp, vp],
            "ewdml_bn_finalize3": [vp, vp, vp, vp, vp, vp, vp, i32, i64, f32, f32, vp, vp],
            "ewdml_bn_fwd_fused": [vp, vp, i32, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, f32, f32, vp],
            "ewdml_bn_bwd_fused": [vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
            "ewdml_f32_to_bf16_stats": [vp, vp, vp, vp, i64, i32, vp],
            "ewdml_bn_act_fwd": [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp],
            "ewdml_bn_act_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp],
            "ewdml_add_vec_f32": [vp, vp, i32, vp],
            "ewdml_cast_f32_bf16": [vp, vp, i64, vp],
            "ewdml_pad_rows_bf16": [vp, vp, i32, i32, i32, vp],
            "ewdml_relu_fwd": [vp, vp, i64, vp],
            "ewdml_relu_bwd": [vp, vp, vp, i64, vp],
            "ewdml_dropout": [vp, vp, i64, f32, u32, u32, vp, vp],
            "ewdml_avgpool_fwd": [vp, vp, i32, i32, i32, vp],
            "ewdml_avgpool_bwd": [vp, vp, i32, i32, i32, vp],
            "ewdml_xent": [vp, vp, vp, vp, i32, i32, i32, vp],
        }
        for name, args in sig.items():
            fn = getattr(_lib, name)
            fn.restype = i32
            fn.argtypes = args
    return _lib


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t: Optional[torch.Tensor]):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


OUT_BF16, OUT_F32, OUT_ACCUM = 0, 1, 2
USE_SMALL_CONV = False   # direct first-layer kernels (kept for reference; im2col + tcgen05 GEMM measured faster on B200)
USE_IGEMM = True      # implicit-GEMM path for 3x3 / stride 1 / pad 1 convolutions with C_in, C_out % 64 == 0
# cooperative one-launch BatchNorm forward / backward (un-pooled blocks).  Correct and tested, but NOT faster on B200: the grid
# barrier + cooperative launch cost what the two saved launches cost (ResNet-50 7.77 ms fused vs 7.60 ms with the separate
# streaming kernels, VGG-11 1.157 vs 1.141 ms; profiles/README.md), so it stays opt-in (EWDML_FUSED_BN=1).
USE_FUSED_BN = os.environ.get("EWDML_FUSED_BN", "0") == "1"


def _stream_ok(c: int) -> bool:
    """Shapes the streaming / fused BatchNorm kernels cover: C/8 a power of two <= 256 (csrc/ops/nn.cu stream_ok)."""
    c8n = c // 8
    return c % 8 == 0 and 1 <= c8n <= 256 and (c8n & (c8n - 1)) == 0

# device step state of the active trainer (uint32[4] tensor) -- lets dropout masks change on graph replays
STEP_STATE = {"tensor": None}


def gemm(a: torch.Tensor, b: torch.Tensor, a_mn: bool = False, b_mn: bool = False,
         out: Optional[torch.Tensor] = None, out_mode: int = OUT_BF16, bias: Optional[torch.Tensor] = None,
         alpha: float = 1.0, split_k: int = 0) -> torch.Tensor:
    """tcgen05 GEMM ``out (op)= alpha * op(a) @ op(b) (+ bias)`` on bf16 row-major matrices.

    ``a_mn=False``: a is [M, K];  ``a_mn=True``: a is [K, M] (used as its transpose).
    ``b_mn=False``: b is [N, K];  ``b_mn=True``: b is [K, N].
    Row pitches may exceed the logical width (views of wider matrices are fine).
    """
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.dim() == 2 and b.dim() == 2
    assert a.stride(1) == 1 and b.stride(1) == 1
    (K, M) = a.shape if a_mn else (a.shape[1], a.shape[0])
    (Kb, N) = b.shape if b_mn else (b.shape[1], b.shape[0])
    assert K == Kb, (a.shape, b.shape, a_mn, b_mn)
    if out is None:
        out = torch.empty(M, N, dtype=torch.bfloat16 if out_mode == OUT_BF16 else torch.float32, device=a.device)
        if out_mode == OUT_ACCUM:
            out.zero_()
    assert out.stride(-1) == 1
    ldc = out.stride(0) if out.dim() == 2 else N
    L.check(_l().ewdml_gemm_bf16(_p(a), _p(b), _p(out), M, N, K, a.stride(0), b.stride(0), ldc, int(a_mn), int(b_mn),
                                 out_mode, _p(bias), float(alpha), split_k, _stream()), "gemm_bf16")
    return out


def gemm_tn(a, b, **kw):
    return gemm(a, b, False, False, **kw)


# ======================================================================================
# zero-initialised fp32 scratch: one arena per device, cleared with ONE memset per step
# ======================================================================================
class _Arena:
    """Bump allocator over a pre-zeroed fp32 buffer.  ``begin_step`` clears the part used by the previous step, so
    the dozens of small zero-filled scratch vectors of a step (BatchNorm sums, split-K accumulators) cost one memset
    instead of one fill kernel each.  Without ``begin_step`` (unit tests) it falls back to ``torch.zeros``."""

    def __init__(self):
        self.buf, self.pos, self.hwm, self.active = None, 0, 0, False

    def begin_step(self, device, nfloats: int = 48 << 20) -> None:
        if self.buf is None or self.buf.device != device:
            self.buf = torch.zeros(nfloats, dtype=torch.float32, device=device)
            self.hwm = 0
        self.hwm = max(self.hwm, self.pos)
        if self.hwm:
            self.buf[:self.hwm].zero_()
        self.pos, self.active = 0, True

    def zeros(self, n: int, device) -> torch.Tensor:
        n_al = (n + 63) // 64 * 64
        if not self.active or self.buf is None or self.buf.device != device or self.pos + n_al > self.buf.numel():
            return torch.zeros(n, dtype=torch.float32, device=device)
        out = self.buf[self.pos:self.pos + n]
        self.pos += n_al
        return out


ARENA = _Arena()


def begin_step(device) -> None:
    ARENA.begin_step(device)


def end_step() -> None:
    """Stop handing out arena slices (anything after this point gets ordinary ``torch.zeros``)."""
    ARENA.hwm = max(ARENA.hwm, ARENA.pos)
    ARENA.active = False


# ======================================================================================
# layer functions
# ======================================================================================
BF16 = torch.bfloat16


def _r8(n: int) -> int:
    return (n + 7) // 8 * 8


def _empty2d(rows: int, cols: int, device, zero: bool = False) -> torch.Tensor:
    """[rows, cols] bf16 view of a buffer whose row pitch is a multiple of 8 elements (TMA operand rule)."""
    pitch = _r8(cols)
    buf = (torch.zeros if (zero or pitch != cols) else torch.empty)(rows, pitch, dtype=BF16, device=device)
    return buf[:, :cols] if pitch != cols else buf


def to_nhwc_bf16(x: torch.Tensor) -> torch.Tensor:
    """[N,C,H,W] (any format/dtype) -> contiguous [N,H,W,C] bf16."""
    xn = x.permute(0, 2, 3, 1)
    if x.dtype == BF16:
        return xn.contiguous()
    xn = xn.contiguous()
    if x.dtype == torch.float32:
        out = torch.empty(xn.shape, dtype=BF16, device=x.device)
        L.check(_l().ewdml_cast_f32_bf16(_p(xn), _p(out), xn.numel(), _stream()), "cast")
        return out
    return xn.to(BF16)


def _weight_bf16(w: torch.nn.Parameter, rows: int, k: int, pad: bool = True) -> torch.Tensor:
    """bf16 [rows, Kp] K-major weight matrix: the shadow copy kept by the fused optimizer when present."""
    sh = getattr(w, "_ewdml_bf16", None)
    if sh is None:
        src = w.detach()
        src = src.permute(0, 2, 3, 1) if src.dim() == 4 else src
        sh = src.reshape(rows, k).to(BF16).contiguous().view(-1)
    wb = sh.view(rows, k)
    if k % 8 == 0 or not pad:
        return wb
    kp = _r8(k)
    out = torch.empty(rows, kp, dtype=BF16, device=w.device)
    L.check(_l().ewdml_pad_rows_bf16(_p(wb), _p(out), rows, k, kp, _stream()), "pad_rows")
    return out


def _grad2d(w: torch.nn.Parameter, rows: int, k: int) -> torch.Tensor:
    """fp32 [rows, k] view of the parameter's gradient in storage order (a slice of the flat gradient buffer)."""
    if w.grad is None:
        w.grad = torch.zeros_like(w)
    g = w.grad
    g = g.permute(0, 2, 3, 1) if g.dim() == 4 else g
    if not g.is_contiguous():
        raise RuntimeError("native compute needs conv weights/gradients in channels-last storage (FlatModel(channels_last=True))")
    return g.reshape(rows, k)


def _colsum_into(dy2d: torch.Tensor, target: torch.Tensor) -> None:
    """target (fp32 [C]) += column sums of dy2d ([M, C] bf16 contiguous)."""
    M, Cn = dy2d.shape
    sums = ARENA.zeros(2 * Cn, dy2d.device)
    L.check(_l().ewdml_colstats(_p(dy2d), _p(sums), M, Cn, 0, _stream()), "colstats")
    L.check(_l().ewdml_add_vec_f32(_p(sums), _p(target), Cn, _stream()), "add_vec")


def _conv_supported(conv, x) -> bool:
    """Square kernel / stride / padding, no groups, no dilation.  C_out % 8 == 0 goes to the tcgen05 GEMM paths, anything
    else (LeNet's 20 / 50 channels) to the direct small-channel kernels (csrc/ops/conv_direct.cu)."""
    return (conv.groups == 1 and conv.dilation == (1, 1)
            and conv.kernel_size[0] == conv.kernel_size[1] and conv.stride[0] == conv.stride[1]
            and conv.padding[0] == conv.padding[1] and isinstance(conv.padding, tuple))


class _ConvBlockFn(torch.autograd.Function):
    """out = [maxpool2x2] act( bn( conv(x) + bias ) [+ residual] ) with every stage on the native kernels."""

    @staticmethod
    def forward(ctx, x, residual, _w, conv, bn, relu, pool):      # _w: the weight, so autograd sees a differentiable input
        lib = _l()
        st = _stream()
        dev = x.device
        N, Cin, H, W = x.shape
        kh, stride, pad = conv.kernel_size[0], conv.stride[0], conv.padding[0]
        Cout = conv.out_channels
        HO, WO = (H + 2 * pad - kh) // stride + 1, (W + 2 * pad - kh) // stride + 1
        M, K = N * HO * WO, kh * kh * Cin
        Kp = _r8(K)
        generic = Cout % 8 != 0                       # direct CUDA-core kernels: channel counts the GEMM tiles do not take
        small = (USE_SMALL_CONV and Cin <= 4 and kh * kh * Cin <= 36 and Cout in (64, 128) and not ctx.needs_input_grad[0]
                 and x.dtype in (torch.float32, BF16))
        if small:
            xn = x.permute(0, 2, 3, 1).contiguous()          # NHWC view of a channels-last tensor: no copy, no cast
        else:
            xn = to_nhwc_bf16(x)
        direct = kh == 1 and stride == 1 and pad == 0 and Cin % 8 == 0 and not generic
        igemm = USE_IGEMM and kh == 3 and stride == 1 and pad == 1 and Cin % 64 == 0 and Cout % 64 == 0
        wb = _weight_bf16(conv.weight, Cout, K, pad=not (small or generic))
        bias = conv.bias.detach() if conv.bias is not None else None
        train_bn = bn is not None and bn.training
        sums = ARENA.zeros(2 * Cout + 32, dev) if train_bn else None       # + one barrier word for the fused (cooperative) BN kernel
        # un-pooled training BatchNorm with C/8 a power of two: statistics, finalize and normalise run as ONE cooperative kernel
        fuse_bn = USE_FUSED_BN and train_bn and not pool and _stream_ok(Cout)
        if generic:
            if train_bn and Cout % 2:
                raise RuntimeError("native BatchNorm needs an even channel count")
            igemm = direct = small = False
            col = xn
            y = torch.empty(M, Cout, dtype=BF16, device=dev)
            L.check(lib.ewdml_conv_direct_fwd(_p(xn), _p(wb), _p(bias), _p(y), N, H, W, Cin, Cout, kh, kh, stride, pad, st),
                    "conv_direct_fwd")
            if train_bn:
                L.check(lib.ewdml_colstats(_p(y), _p(sums), M, Cout, 1, st), "colstats")
        elif small:
            igemm = direct = False
            col = xn
            y = torch.empty(M, Cout, dtype=BF16, device=dev)
            L.check(lib.ewdml_conv_small_fwd(_p(xn), int(xn.dtype == torch.float32), _p(wb), _p(bias), _p(y), N, H, W, Cin, Cout,
                                             kh, kh, stride, pad, st), "conv_small_fwd")
            if train_bn:
                L.check(lib.ewdml_colstats(_p(y), _p(sums), M, Cout, 1, st), "colstats")
        elif igemm:
            # implicit GEMM: zero-bordered copy of x, every filter tap is a row-shifted TMA load of the same matrix
            col = torch.empty(N, H + 2, W + 2, Cin, dtype=BF16, device=dev)
            L.check(lib.ewdml_pad_nhwc(_p(xn), _p(col), N, H, W, Cin, st), "pad_nhwc")
            m_rows, kb = N * (H + 2) * (W + 2), 9 * (Cin // 64)
        else:
            if direct:
                col = xn.view(M, Cin)
            else:
                col = torch.empty(M, Kp, dtype=BF16, device=dev)
                L.check(lib.ewdml_im2col(_p(xn), _p(col), N, H, W, Cin, kh, kh, stride, pad, st), "im2col")
            m_rows, kb = M, (Kp + 63) // 64
        if small or generic:
            m_rows, kb = M, 1
        small = small or generic                   # from here on: "the output y and its statistics already exist"
        tiles = ((m_rows + 127) // 128) * ((Cout + 127) // 128)
        split = (not small) and tiles <= 96 and kb >= 16
        # few output tiles + long reduction: split-K into an fp32 accumulator, then one pass that adds the bias, rounds
        # to bf16 and gathers the BatchNorm statistics
        acc = ARENA.zeros(M * Cout, dev).view(M, Cout) if split else None
        if not small:
            y = torch.empty(M, Cout, dtype=BF16, device=dev)
        if small:
            pass
        elif igemm:
            L.check(lib.ewdml_conv3x3_igemm(1, _p(col), _p(wb), _p(acc if split else y), N, H, W, Cin, Cout,
                                            OUT_ACCUM if split else OUT_BF16, None if split else _p(bias), 0, st), "igemm fprop")
        elif split:
            gemm(col, wb, out=acc, out_mode=OUT_ACCUM)
        else:
            gemm(col, wb, out=y, bias=bias)
        shifted = False
        have_sums = small or split
        if split:
            L.check(lib.ewdml_f32_to_bf16_stats(_p(acc), _p(bias), _p(y), _p(sums), M, Cout, st), "f32_to_bf16_stats")
        elif fuse_bn:
            pass                                                       # the fused kernel gathers the statistics itself
        elif train_bn and not small:
            # shifted sums (relative to row 0 of y): the variance does not cancel when |mean| >> std
            c8n = Cout // 8
            shifted = Cout % 8 == 0 and c8n <= 256 and (c8n & (c8n - 1)) == 0
            L.check(lib.ewdml_colstats(_p(y), _p(sums), M, Cout, 2 if shifted else 1, st), "colstats")
        stats = None
        resn = to_nhwc_bf16(residual) if residual is not None else None
        fused_done = False
        if fuse_bn:
            stats = torch.empty(4 * Cout, dtype=torch.float32, device=dev)
            out = torch.empty(N, HO, WO, Cout, dtype=BF16, device=dev)
            mom = bn.momentum if bn.momentum is not None else 0.1
            track = bn.track_running_stats
            L.check(lib.ewdml_bn_fwd_fused(_p(y), _p(sums), int(have_sums), _p(bn.weight.detach()), _p(bn.bias.detach()),
                                           _p(bn.running_mean) if track else None, _p(bn.running_var) if track else None,
                                           _p(bn.num_batches_tracked) if track and bn.num_batches_tracked is not None else None,
                                           _p(stats), _p(resn), _p(out), N, HO, WO, Cout, int(relu), float(mom), float(bn.eps), st),
                    "bn_fwd_fused")
            fused_done = True
        elif bn is not None:
            stats = torch.empty(4 * Cout, dtype=torch.float32, device=dev)
            if bn.training:
                mom = bn.momentum if bn.momentum is not None else 0.1
                track = bn.track_running_stats
                L.check(lib.ewdml_bn_finalize3(_p(sums), _p(y) if shifted else None, _p(bn.weight.detach()), _p(bn.bias.detach()),
                                               _p(bn.running_mean) if track else None, _p(bn.running_var) if track else None,
                                               _p(stats), Cout, M, float(mom), float(bn.eps),
                                               _p(bn.num_batches_tracked) if track and bn.num_batches_tracked is not None else None, st),
                        "bn_finalize")
            else:
                rstd = torch.rsqrt(bn.running_var + bn.eps)
                sc = bn.weight.detach() * rstd
                stats.copy_(torch.cat([bn.running_mean, rstd, sc, bn.bias.detach() - bn.running_mean * sc]))
        if not fused_done:
            HP, WP = (HO // 2, WO // 2) if pool else (HO, WO)
            out = torch.empty(N, HP, WP, Cout, dtype=BF16, device=dev)
            L.check(lib.ewdml_bn_act_fwd(_p(y), _p(stats), _p(resn), _p(out), N, HO, WO, Cout, int(relu), int(pool), st), "bn_act_fwd")
        ctx.save_for_backward(col, y, stats if stats is not None else torch.empty(0, device=dev),
                              resn if resn is not None else torch.empty(0, device=dev))
        ctx.meta = (conv, bn, relu, pool, N, H, W, Cin, HO, WO, Cout, kh, stride, pad, M, K, Kp, direct, igemm, small, generic)
        ctx.has_res = residual is not None
        return out.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dout):
        lib = _l()
        st = _stream()
        col, y, stats, resn = ctx.saved_tensors
        conv, bn, relu, pool, N, H, W, Cin, HO, WO, Cout, kh, stride, pad, M, K, Kp, direct, igemm, small, generic = ctx.meta
        dev = dout.device
        stats = stats if stats.numel() else None
        resn = resn if resn.numel() else None
        dn = to_nhwc_bf16(dout)
        need_dz = bn is not None or relu or pool
        dres = None
        if need_dz:
            dy = torch.empty(M, Cout, dtype=BF16, device=dev)
            dres = torch.empty(M, Cout, dtype=BF16, device=dev) if ctx.has_res else None
            sums = ARENA.zeros(2 * Cout + 32, dev)
            dg = db = None
            if bn is not None:
                if bn.weight.grad is None:
                    bn.weight.grad = torch.zeros_like(bn.weight)
                if bn.bias.grad is None:
                    bn.bias.grad = torch.zeros_like(bn.bias)
                dg, db = bn.weight.grad, bn.bias.grad
            if USE_FUSED_BN and stats is not None and not pool and _stream_ok(Cout):
                L.check(lib.ewdml_bn_bwd_fused(_p(y), _p(stats), _p(resn), _p(dn), _p(sums), _p(dy), _p(dres), _p(dg), _p(db),
                                               N, HO, WO, Cout, int(relu), st), "bn_bwd_fused")
            else:
                L.check(lib.ewdml_bn_act_bwd(_p(y), _p(stats), _p(resn), _p(dn), _p(sums), _p(dy), _p(dres), _p(dg), _p(db),
                                             N, HO, WO, Cout, int(relu), int(pool), st), "bn_act_bwd")
        else:
            dy = dn.view(M, Cout)
            dres = dy if ctx.has_res else None
        gw = _grad2d(conv.weight, Cout, K)
        if conv.bias is not None and bn is None:
            if conv.bias.grad is None:
                conv.bias.grad = torch.zeros_like(conv.bias)
            _colsum_into(dy, conv.bias.grad)
        dx = None
        if generic:
            L.check(lib.ewdml_conv_direct_wgrad(_p(col), _p(dy), _p(gw), N, H, W, Cin, Cout, kh, kh, stride, pad, st), "conv_direct_wgrad")
            if ctx.needs_input_grad[0]:
                wb = _weight_bf16(conv.weight, Cout, K, pad=False)
                dxn = torch.empty(N, H, W, Cin, dtype=BF16, device=dev)
                L.check(lib.ewdml_conv_direct_dgrad(_p(dy), _p(wb), _p(dxn), N, H, W, Cin, Cout, kh, kh, stride, pad, st), "conv_direct_dgrad")
                dx = dxn.permute(0, 3, 1, 2)
        elif small:
            L.check(lib.ewdml_conv_small_wgrad(_p(col), int(col.dtype == torch.float32), _p(dy), _p(gw), N, H, W, Cin, Cout, kh, kh,
                                               stride, pad, st), "conv_small_wgrad")
        elif igemm:
            dyp = torch.empty(N, H + 2, W + 2, Cout, dtype=BF16, device=dev)
            L.check(lib.ewdml_pad_nhwc(_p(dy), _p(dyp), N, H, W, Cout, st), "pad_nhwc dy")
            L.check(lib.ewdml_conv3x3_igemm(3, _p(dyp), _p(col), _p(gw), N, H, W, Cin, Cout, OUT_ACCUM, None, 0, st), "igemm wgrad")
            if ctx.needs_input_grad[0]:
                wb = _weight_bf16(conv.weight, Cout, K)
                tiles = ((N * (H + 2) * (W + 2) + 127) // 128) * ((Cin + 127) // 128)
                dxn = torch.empty(N, H, W, Cin, dtype=BF16, device=dev)
                if tiles <= 96:
                    acc = ARENA.zeros(M * Cin, dev)
                    L.check(lib.ewdml_conv3x3_igemm(2, _p(dyp), _p(wb), _p(acc), N, H, W, Cin, Cout, OUT_ACCUM, None, 0, st), "igemm dgrad")
                    L.check(lib.ewdml_f32_to_bf16_stats(_p(acc), None, _p(dxn), None, M, Cin, st), "f32_to_bf16")
                else:
                    L.check(lib.ewdml_conv3x3_igemm(2, _p(dyp), _p(wb), _p(dxn), N, H, W, Cin, Cout, OUT_BF16, None, 0, st), "igemm dgrad")
                dx = dxn.permute(0, 3, 1, 2)
        else:
            # weight gradient: dW[Cout, K] += dy^T col (both operands MN-major, split-K, fp32 atomics into the flat buffer)
            gemm(dy, col[:, :K] if Kp != K else col, True, True, out=gw, out_mode=OUT_ACCUM)
            if ctx.needs_input_grad[0]:
                wb = _weight_bf16(conv.weight, Cout, K)                  # [Cout, Kp]: used as the MN-major B operand [K=Cout][N=Kp]
                tiles = ((M + 127) // 128) * ((Kp + 127) // 128)
                if tiles <= 96 and Cout >= 1024:
                    acc = ARENA.zeros(M * Kp, dev).view(M, Kp)
                    gemm(dy, wb, False, True, out=acc, out_mode=OUT_ACCUM)
                    dcol = torch.empty(M, Kp, dtype=BF16, device=dev)
                    L.check(lib.ewdml_f32_to_bf16_stats(_p(acc), None, _p(dcol), None, M, Kp, st), "f32_to_bf16")
                else:
                    dcol = gemm(dy, wb, False, True)                     # [M, Kp] bf16
                if direct:
                    dxn = dcol.view(N, H, W, Cin)
                else:
                    dxn = torch.empty(N, H, W, Cin, dtype=BF16, device=dev)
                    L.check(lib.ewdml_col2im(_p(dcol), _p(dxn), N, H, W, Cin, kh, kh, stride, pad, st), "col2im")
                dx = dxn.permute(0, 3, 1, 2)
        dr = dres.view(N, HO, WO, Cout).permute(0, 3, 1, 2) if (ctx.has_res and ctx.needs_input_grad[1]) else None
        if _overlap._ACTIVE:       # gradients went straight into the flat buffer: tell the bucket scheduler
            for prm in (conv.weight, conv.bias, bn.weight if bn is not None else None, bn.bias if bn is not None else None):
                if prm is not None:
                    _overlap.notify_grad_ready(prm)
        return dx, dr, None, None, None, None, None


def conv_block(x, conv, bn=None, relu=False, residual=None, pool=False):
    """Fused conv (+bias) [+BatchNorm] [+residual] [+ReLU] [+2x2 max-pool]; falls back to torch for odd shapes."""
    if not _conv_supported(conv, x) or (pool and residual is not None):
        import torch.nn.functional as F
        y = F.conv2d(x.float() if x.dtype != conv.weight.dtype else x, conv.weight, conv.bias, conv.stride, conv.padding)
        if bn is not None:
            y = F.batch_norm(y, bn.running_mean, bn.running_var, bn.weight, bn.bias, bn.training, bn.momentum or 0.1, bn.eps)
        if residual is not None:
            y = y + residual
        if relu:
            y = F.relu(y)
        if pool:
            y = F.max_pool2d(y, 2, 2)
        return y
    return _ConvBlockFn.apply(x, residual, conv.weight, conv, bn, relu, pool)


class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, _w, lin, relu):
        lib = _l()
        B = x.shape[0]
        if x.dtype != BF16 or x.stride(1) != 1 or x.stride(0) % 8:
            xb = _empty2d(B, x.shape[1], x.device)
            xb.copy_(x)
            x = xb
        wb = _weight_bf16(lin.weight, lin.out_features, lin.in_features)[:, :lin.in_features]   # in_features % 8 != 0: padded pitch, logical K
        y = _empty2d(B, lin.out_features, x.device, zero=True)
        gemm(x, wb, out=y, bias=lin.bias.detach() if lin.bias is not None else None)
        if relu:
            buf = y if y.is_contiguous() else None
            if buf is not None:
                L.check(lib.ewdml_relu_fwd(_p(y), _p(y), y.numel(), _stream()), "relu")
            else:
                y.clamp_(min=0)
        ctx.save_for_backward(x, y if relu else torch.empty(0, device=x.device))
        ctx.meta = (lin, relu)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _l()
        x, yact = ctx.saved_tensors
        lin, relu = ctx.meta
        B = x.shape[0]
        if dy.dtype != BF16 or dy.stride(1) != 1 or dy.stride(0) % 8:
            d = _empty2d(B, dy.shape[1], dy.device, zero=True)
            d.copy_(dy)
            dy = d
        if relu:
            d2 = _empty2d(B, dy.shape[1], dy.device, zero=True)
            if dy.is_contiguous() and yact.is_contiguous():
                L.check(lib.ewdml_relu_bwd(_p(dy), _p(yact), _p(d2), dy.numel(), _stream()), "relu_bwd")
            else:
                d2.copy_(dy * (yact > 0))
            dy = d2
        gw = _grad2d(lin.weight, lin.out_features, lin.in_features)
        gemm(dy, x, True, True, out=gw, out_mode=OUT_ACCUM)                 # dW += dy^T x
        if lin.bias is not None:
            if lin.bias.grad is None:
                lin.bias.grad = torch.zeros_like(lin.bias)
            if dy.is_contiguous():
                _colsum_into(dy, lin.bias.grad)
            else:
                lin.bias.grad += dy.float().sum(0)
        dx = None
        if ctx.needs_input_grad[0]:
            wb = _weight_bf16(lin.weight, lin.out_features, lin.in_features)
            dx = _empty2d(B, lin.in_features, dy.device)
            gemm(dy, wb, False, True, out=dx)                                # dx = dy W
        if _overlap._ACTIVE:
            _overlap.notify_grad_ready(lin.weight)
            if lin.bias is not None:
                _overlap.notify_grad_ready(lin.bias)
        return dx, None, None, None


def linear_block(x, lin, relu=False):
    return _LinearFn.apply(x, lin.weight, lin, relu)


class _XentFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        B, Cn = logits.shape
        if logits.dtype != BF16 or logits.stride(1) != 1:
            lb = _empty2d(B, Cn, logits.device, zero=True)
            lb.copy_(logits)
            logits = lb
        acc = torch.zeros(2, dtype=torch.float32, device=logits.device)
        dl = torch.zeros(B, logits.stride(0), dtype=BF16, device=logits.device)
        L.check(_l().ewdml_xent(_p(logits), _p(target), _p(acc), _p(dl), B, Cn, logits.stride(0), _stream()), "xent")
        ctx.save_for_backward(dl[:, :Cn])
        ctx.correct = acc[1:2]
        return acc[0].clone()

    @staticmethod
    def backward(ctx, g):
        (dl,) = ctx.saved_tensors
        return dl * g.to(dl.dtype), None


def cross_entropy(logits, target):
    return _XentFn.apply(logits, target)


class _DropoutFn(torch.autograd.Function):
    _counter = [0]

    @staticmethod
    def forward(ctx, x, p):
        _DropoutFn._counter[0] += 1
        seed = int(torch.initial_seed() & 0xFFFFFFFF)
        off = _DropoutFn._counter[0] & 0xFFFFFFFF
        xc = x if x.is_contiguous() else x.contiguous()
        y = torch.empty_like(xc)
        L.check(_l().ewdml_dropout(_p(xc), _p(y), xc.numel(), float(p), seed, off, _p(STEP_STATE["tensor"]), _stream()), "dropout")
        ctx.meta = (p, seed, off)
        return y

    @staticmethod
    def backward(ctx, dy):
        p, seed, off = ctx.meta
        d = dy if dy.is_contiguous() else dy.contiguous()
        dx = torch.empty_like(d)
        L.check(_l().ewdml_dropout(_p(d), _p(dx), d.numel(), float(p), seed, off, _p(STEP_STATE["tensor"]), _stream()), "dropout_bwd")
        return dx, None


def dropout(x, p):
    if x.dtype != BF16:
        x = x.to(BF16)
    return _DropoutFn.apply(x, p)


class _AvgPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        N, Cn, H, W = x.shape
        xn = to_nhwc_bf16(x)
        y = torch.empty(N, Cn, dtype=BF16, device=x.device)
        L.check(_l().ewdml_avgpool_fwd(_p(xn), _p(y), N, H * W, Cn, _stream()), "avgpool")
        ctx.meta = (N, Cn, H, W)
        return y.view(N, Cn, 1, 1)

    @staticmethod
    def backward(ctx, dy):
        N, Cn, H, W = ctx.meta
        d = dy.reshape(N, Cn).to(BF16).contiguous()
        dx = torch.empty(N, H, W, Cn, dtype=BF16, device=dy.device)
        L.check(_l().ewdml_avgpool_bwd(_p(d), _p(dx), N, H * W, Cn, _stream()), "avgpool_bwd")
        return dx.permute(0, 3, 1, 2)


def avg_pool(x, kernel_size):
    k = kernel_size if isinstance(kernel_size, int) else kernel_size[0]
    if x.shape[2] != k or x.shape[3] != k:
        return torch.nn.functional.avg_pool2d(x, k)
    return _AvgPoolFn.apply(x)


# ---- entry points used by ewdml.ops.layers (unfused module-by-module use) ---------------------------
def conv2d(x, weight, bias, stride, padding):
    raise RuntimeError("native conv is reached through conv_block(); see ewdml.ops.layers.Conv2d.forward")


def linear(x, weight, bias):
    raise RuntimeError("native linear is reached through linear_block(); see ewdml.ops.layers.Linear.forward")


class _ReluFn(torch.autograd.Function):
    """Stand-alone ReLU on the native elementwise kernels (nn.ReLU sites outside a fused conv / linear block)."""

    @staticmethod
    def forward(ctx, x):
        xc = x.contiguous() if x.dtype == BF16 else x.to(BF16).contiguous()
        y = torch.empty_like(xc)
        L.check(_l().ewdml_relu_fwd(_p(xc), _p(y), xc.numel(), _stream()), "relu_fwd")
        ctx.save_for_backward(y)
        ctx.in_dtype = x.dtype
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        d = dy.to(BF16).contiguous()
        dx = torch.empty_like(d)
        L.check(_l().ewdml_relu_bwd(_p(d), _p(y), _p(dx), d.numel(), _stream()), "relu_bwd")
        return dx.to(ctx.in_dtype)


class _Pool2x2Fn(torch.autograd.Function):
    """Stand-alone 2x2 / stride-2 max-pool (optionally followed by ReLU): the pooling half of the fused BatchNorm/activation
    kernels with an identity affine (``stats = NULL``), forward and backward."""

    @staticmethod
    def forward(ctx, x, relu):
        N, Cn, H, W = x.shape
        xn = to_nhwc_bf16(x)
        out = torch.empty(N, H // 2, W // 2, Cn, dtype=BF16, device=x.device)
        L.check(_l().ewdml_bn_act_fwd(_p(xn), None, None, _p(out), N, H, W, Cn, int(relu), 1, _stream()), "pool_fwd")
        ctx.save_for_backward(xn)
        ctx.meta = (N, Cn, H, W, relu, x.dtype)
        return out.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dout):
        (xn,) = ctx.saved_tensors
        N, Cn, H, W, relu, in_dtype = ctx.meta
        dn = to_nhwc_bf16(dout)
        dx = torch.empty(N, H, W, Cn, dtype=BF16, device=dout.device)
        sums = ARENA.zeros(2 * Cn + 32, dout.device)
        L.check(_l().ewdml_bn_act_bwd(_p(xn), None, None, _p(dn), _p(sums), _p(dx), None, None, None, N, H, W, Cn, int(relu), 1, _stream()),
                "pool_bwd")
        return dx.permute(0, 3, 1, 2).to(in_dtype), None


def relu(x):
    """nn.ReLU as a stand-alone native op (bf16 out)."""
    return _ReluFn.apply(x)


def max_pool2x2(x, relu=False):
    """nn.MaxPool2d(2, 2) [+ ReLU] as a stand-alone native op; odd spatial sizes fall back to torch (floor semantics)."""
    if x.shape[2] % 2 or x.shape[3] % 2 or x.shape[1] % 2:
        y = torch.nn.functional.max_pool2d(x, 2, 2)
        return torch.relu(y) if relu else y
    return _Pool2x2Fn.apply(x, relu)
