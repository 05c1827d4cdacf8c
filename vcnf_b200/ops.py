"""Autograd bindings of the C-ABI kernels.  Every function here launches hand-written sm_100a
kernels on the current CUDA stream; backward passes are hand-written kernels too (no autograd
tracing through ATen for the bijector math)."""
import math

import torch
from torch.autograd.function import once_differentiable

from . import _lib
from ._lib import call, ptr, stream, dtype_code

# kernel launches issued through this module (read by bench.py for `gpu_launches`)
LAUNCHES = 0

COL_PASS, COL_COND, COL_SHARED = 0, 1, 2
TAIL_LINEAR, TAIL_CIRCULAR, TAIL_LIST_LINEAR, TAIL_LIST_CIRCULAR, TAIL_NONE = 0, 1, 2, 3, 4
SCALE_MAPS = {"exp": 0, "sigmoid": 1, "sigmoid_inv": 2, "none": 3, "maf": 4}


def _launch(name, *args, n=1):
    global LAUNCHES
    LAUNCHES += n
    call(name, *args)


def _promote(*tensors):
    """Type promotion of the floating-point operands like the reference's ATen ops do (an fp32 batch
    fed to an `nfm.double()` model computes in fp64, runadultvdeq.py:172,183)."""
    dt = None
    for t in tensors:
        if t is not None:
            dt = t.dtype if dt is None else torch.promote_types(dt, t.dtype)
    return tuple(t if t is None or t.dtype == dt else t.to(dt) for t in tensors)


def _padded_empty(rows, cols, dtype, device):
    """[rows, cols] tensor whose row pitch is a multiple of 16 bytes (a view of wider storage when
    `cols` is not): what the TMA engine needs to move row tiles (tensor-map loads / stores in the
    GEMM, bulk tile copies in the spline kernels).  The pad columns are never read."""
    q = 16 // torch.empty((), dtype=dtype).element_size()
    pad = (cols + q - 1) // q * q
    if pad == cols:
        return torch.empty((rows, cols), dtype=dtype, device=device)
    return torch.empty((rows, pad), dtype=dtype, device=device)[:, :cols]


def _rows(t):
    """Row-major 2-D operand with unit column stride (any row pitch); copies only if it is not."""
    return t if (t.dim() == 2 and t.stride(1) == 1 and t.stride(0) >= t.shape[1]) else t.contiguous()


def _zeros_if_none(g, like_shape, ref):
    return torch.zeros(like_shape, dtype=ref.dtype, device=ref.device) if g is None else g.contiguous()


class SplineSpec:
    """Per-layer constant metadata of a spline layer, resident on the device."""

    def __init__(self, D, K, col_map, cond_modes, cond_bounds, Pc, wh_scale,
                 shared_modes, shared_bounds, Ps, outside_zero):
        self.D, self.K = D, K
        self.col_map_host = torch.as_tensor(col_map, dtype=torch.int32)
        self.cond_modes_host = torch.as_tensor(cond_modes, dtype=torch.uint8)
        self.cond_bounds_host = torch.as_tensor(cond_bounds, dtype=torch.float64)
        self.shared_modes_host = torch.as_tensor(shared_modes, dtype=torch.uint8)
        self.shared_bounds_host = torch.as_tensor(shared_bounds, dtype=torch.float64)
        self.Tc, self.Ts = len(self.cond_modes_host), len(self.shared_modes_host)
        self.Pc, self.Ps = Pc, Ps
        self.wh_scale = float(wh_scale)
        self.outside_zero = int(outside_zero)
        self._cache = {}

    def on(self, device, dtype):
        key = (str(device), dtype)
        if key not in self._cache:
            self._cache[key] = (
                self.col_map_host.to(device),
                self.cond_modes_host.to(device), self.cond_bounds_host.to(device, dtype),
                self.shared_modes_host.to(device), self.shared_bounds_host.to(device, dtype))
        return self._cache[key]


class _RqsLayer(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cond, shared, spec, inverse, want_bins, pre=None):
        x = x.contiguous()
        B, D = x.shape
        if D != spec.D:
            raise ValueError("Expected features = {}, got {}.".format(spec.D, D))
        cmap, cm, cb, sm, sb = spec.on(x.device, x.dtype)
        cond = _rows(cond) if cond is not None else None
        shared = shared.contiguous() if shared is not None else None
        if pre is not None:            # (y, logdet) already computed by the fused net + spline kernel: only the graph node
            ctx.spec, ctx.inverse = spec, int(inverse)
            ctx.save_for_backward(x, cond, shared)
            return pre[0], pre[1]
        y = torch.empty_like(x)
        logdet = torch.empty(B, dtype=x.dtype, device=x.device)
        bins = torch.empty((B, D), dtype=torch.int8, device=x.device) if want_bins else None
        _launch("vcnf_rqs_layer_fwd", ptr(x), D, ptr(y), D, ptr(logdet), 0, B, D, spec.K, ptr(cmap),
                ptr(cond), cond.stride(0) if cond is not None else 0, spec.Tc, spec.Pc, ptr(cm), ptr(cb), spec.wh_scale,
                ptr(shared), spec.Ts, spec.Ps, ptr(sm), ptr(sb), int(inverse), spec.outside_zero,
                ptr(bins), None, dtype_code(x), stream())
        ctx.spec, ctx.inverse = spec, int(inverse)
        ctx.save_for_backward(x, cond, shared)
        if want_bins:
            ctx.mark_non_differentiable(bins)
            return y, logdet, bins
        return y, logdet

    @staticmethod
    @once_differentiable
    def backward(ctx, gy, gld, *unused):
        x, cond, shared = ctx.saved_tensors
        spec = ctx.spec
        B, D = x.shape
        cmap, cm, cb, sm, sb = spec.on(x.device, x.dtype)
        gy = _zeros_if_none(gy, x.shape, x)
        gld = _zeros_if_none(gld, (B,), x)
        gx = torch.empty_like(x)
        gcond = None
        if cond is not None:   # gradient in the layout the parameters came in (padded pitch stays padded)
            gcond = (_padded_empty(*cond.shape, cond.dtype, cond.device) if cond.stride(0) != cond.shape[1]
                     else torch.empty_like(cond))
        gshared = torch.empty_like(shared) if shared is not None else None
        ws = None
        if spec.Ts > 0:
            nbytes = _lib.query("vcnf_rqs_layer_bwd_workspace", spec.Ts, spec.Ps, dtype_code(x))
            ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
        _launch("vcnf_rqs_layer_bwd", ptr(x), D, ptr(gy), D, ptr(gld), ptr(gx), D, B, D, spec.K, ptr(cmap),
                ptr(cond), cond.stride(0) if cond is not None else 0, spec.Tc, spec.Pc, ptr(cm), ptr(cb), spec.wh_scale,
                ptr(gcond), gcond.stride(0) if gcond is not None else 0, ptr(shared), spec.Ts, spec.Ps, ptr(sm), ptr(sb),
                ptr(gshared), ptr(ws), ctx.inverse, spec.outside_zero, dtype_code(x), stream(),
                n=2 if spec.Ts > 0 else 1)
        return gx, gcond, gshared, None, None, None, None


def rqs_layer(x, cond, shared, spec, inverse, want_bins=False, pre=None):
    """Spline layer: returns (y, logdet[, bin_idx]).  pre = (y, logdet) computed elsewhere (ops.resnet_spline): only the
    autograd node is built."""
    x, cond, shared = _promote(x, cond, shared)
    return _RqsLayer.apply(x, cond, shared, spec, inverse, want_bins, pre)


class _GatherCols(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cols, pad, pre=None):
        x = x.contiguous()
        B, D = x.shape
        n = cols.numel()
        if pre is not None:            # already gathered by the fused net + spline kernel: only the graph node
            ctx.save_for_backward(cols)
            ctx.D = D
            return pre
        out = _padded_empty(B, n, x.dtype, x.device) if pad else torch.empty((B, n), dtype=x.dtype, device=x.device)
        _launch("vcnf_gather_cols", ptr(x), D, ptr(out), out.stride(0), B, n, ptr(cols), dtype_code(x), stream())
        ctx.save_for_backward(cols)
        ctx.D = D
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        (cols,) = ctx.saved_tensors
        g = _rows(g)
        B, n = g.shape
        gx = torch.empty((B, ctx.D), dtype=g.dtype, device=g.device)
        _launch("vcnf_scatter_cols", ptr(g), g.stride(0), ptr(gx), ctx.D, B, n, ctx.D, ptr(cols), 0, dtype_code(g), stream())
        return gx, None, None, None


def gather_cols(x, cols_i32, pad=False, pre=None):
    """x[:, cols].  pad=True returns a view whose row pitch is a 16-byte multiple (for consumers that
    take a leading dimension, i.e. ops.linear)."""
    return _GatherCols.apply(x, cols_i32, pad, pre)


class _AffineCoupling(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, param, c2_start, C2, scale_map, inverse, swap_ts):
        z, param = z.contiguous(), param.contiguous()
        B, C = z.shape[0], z.shape[1]
        S = math.prod(z.shape[2:])
        out = torch.empty_like(z)
        logdet = torch.empty(B, dtype=z.dtype, device=z.device)
        _launch("vcnf_affine_coupling_fwd", ptr(z), ptr(param), ptr(out), ptr(logdet), B, C, S, c2_start, C2,
                scale_map, int(inverse), int(swap_ts), dtype_code(z), stream())
        ctx.cfg = (B, C, S, c2_start, C2, scale_map, int(inverse), int(swap_ts))
        ctx.save_for_backward(z, param)
        return out, logdet

    @staticmethod
    @once_differentiable
    def backward(ctx, gout, gld):
        z, param = ctx.saved_tensors
        B, C, S, c2_start, C2, scale_map, inverse, swap_ts = ctx.cfg
        gout = _zeros_if_none(gout, z.shape, z)
        gld = _zeros_if_none(gld, (B,), z)
        gz, gparam = torch.empty_like(z), torch.empty_like(param)
        _launch("vcnf_affine_coupling_bwd", ptr(z), ptr(param), ptr(gout), ptr(gld), ptr(gz), ptr(gparam),
                B, C, S, c2_start, C2, scale_map, inverse, swap_ts, dtype_code(z), stream())
        return gz, gparam, None, None, None, None, None


def affine_coupling(z, param, c2_start, C2, scale_map, inverse, swap_ts=False):
    z, param = _promote(z, param)
    return _AffineCoupling.apply(z, param, c2_start, C2, scale_map, inverse, swap_ts)


class _MaskedAffine(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, mask, s, t, inverse):
        z = z.contiguous()
        B = z.shape[0]
        n = math.prod(z.shape[1:])
        s = s.contiguous() if s is not None else None
        t = t.contiguous() if t is not None else None
        out = torch.empty_like(z)
        logdet = torch.empty(B, dtype=z.dtype, device=z.device)
        _launch("vcnf_masked_affine_fwd", ptr(z), ptr(mask), ptr(s), ptr(t), ptr(out), ptr(logdet), B, n,
                int(inverse), dtype_code(z), stream())
        ctx.inverse = int(inverse)
        ctx.save_for_backward(z, mask, s, t)
        return out, logdet

    @staticmethod
    @once_differentiable
    def backward(ctx, gout, gld):
        z, mask, s, t = ctx.saved_tensors
        B, n = z.shape[0], math.prod(z.shape[1:])
        gout = _zeros_if_none(gout, z.shape, z)
        gld = _zeros_if_none(gld, (B,), z)
        gz = torch.empty_like(z)
        gs = torch.empty_like(z) if s is not None else None
        gt = torch.empty_like(z) if t is not None else None
        _launch("vcnf_masked_affine_bwd", ptr(z), ptr(mask), ptr(s), ptr(t), ptr(gout), ptr(gld), ptr(gz),
                ptr(gs), ptr(gt), B, n, ctx.inverse, dtype_code(z), stream())
        return gz, None, gs, gt, None


def masked_affine(z, mask, s, t, inverse):
    z, mask, s, t = _promote(z, mask, s, t)
    return _MaskedAffine.apply(z, mask, s, t, inverse)


class _ConstAffine(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, s, t, S, inverse):
        z = z.contiguous()
        s_, t_ = s.reshape(-1).contiguous(), t.reshape(-1).contiguous()
        Pn = s_.numel()
        B = z.numel() // (Pn * S)
        out = torch.empty_like(z)
        _launch("vcnf_const_affine_fwd", ptr(z), ptr(s_), ptr(t_), ptr(out), B, Pn, S, int(inverse),
                dtype_code(z), stream())
        ctx.cfg = (B, Pn, S, int(inverse), s.shape, t.shape)
        ctx.save_for_backward(z, s_, t_)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, gout):
        z, s_, t_ = ctx.saved_tensors
        B, Pn, S, inverse, s_shape, t_shape = ctx.cfg
        gout = gout.contiguous()
        gz, gs, gt = torch.empty_like(z), torch.empty_like(s_), torch.empty_like(t_)
        ws = torch.empty(_lib.query("vcnf_const_affine_bwd_workspace", Pn, dtype_code(z)), dtype=torch.uint8,
                         device=z.device)
        _launch("vcnf_const_affine_bwd", ptr(z), ptr(s_), ptr(t_), ptr(gout), ptr(gz), ptr(gs), ptr(gt), ptr(ws),
                B, Pn, S, inverse, dtype_code(z), stream(), n=2)
        return gz, gs.view(s_shape), gt.view(t_shape), None, None


def const_affine(z, s, t, S, inverse):
    z, s, t = _promote(z, s, t)
    return _ConstAffine.apply(z, s, t, S, inverse)


class _PermuteChannels(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, perm, inv_perm):
        z = z.contiguous()
        B, C = z.shape[0], z.shape[1]
        S = math.prod(z.shape[2:])
        out = torch.empty_like(z)
        _launch("vcnf_permute_channels", ptr(z), ptr(out), B, C, S, ptr(perm), dtype_code(z), stream())
        ctx.save_for_backward(perm, inv_perm)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        perm, inv_perm = ctx.saved_tensors
        return _PermuteChannels.apply(g, inv_perm, perm), None, None


def permute_channels(z, perm_i32, inv_perm_i32):
    """out[:, j] = z[:, perm[j]]"""
    return _PermuteChannels.apply(z, perm_i32, inv_perm_i32)


class _Squeeze(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, inverse):
        z = z.contiguous()
        B, C, H, W = z.shape
        if not inverse:
            c, h, w = C // 4, H, W
            out = torch.empty((B, c, 2 * h, 2 * w), dtype=z.dtype, device=z.device)
        else:
            c, h, w = C, H // 2, W // 2
            out = torch.empty((B, 4 * c, h, w), dtype=z.dtype, device=z.device)
        _launch("vcnf_squeeze", ptr(z), ptr(out), B, c, h, w, int(inverse), dtype_code(z), stream())
        ctx.inverse = int(inverse)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        return _Squeeze.apply(g, 1 - ctx.inverse), None


def squeeze(z, inverse):
    return _Squeeze.apply(z, inverse)


class _SmallLinear(torch.autograd.Function):
    """y[r, out_idx[o]] = sum_i W[o, i] x[r, in_idx[i]] + bias[o]  (or W (x - bias))."""

    @staticmethod
    def forward(ctx, x, W, bias, in_idx, out_idx, sub_bias_first):
        x, W = x.contiguous(), W.contiguous()
        D = W.shape[0]
        S = math.prod(x.shape[2:])
        rows = x.shape[0] * S
        y = torch.empty_like(x)
        _launch("vcnf_small_linear", ptr(x), ptr(y), ptr(W), ptr(bias), ptr(in_idx), ptr(out_idx), rows, D, S,
                int(sub_bias_first), dtype_code(x), stream())
        ctx.cfg = (D, S, rows, int(sub_bias_first))
        ctx.save_for_backward(x, W, bias, in_idx, out_idx)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, gy):
        x, W, bias, in_idx, out_idx = ctx.saved_tensors
        D, S, rows, sub = ctx.cfg
        gy = gy.contiguous()
        gx = gW = gb = None
        if S == 1 and not sub and x.dim() == 2:
            # density direction on tabular rows (LULinearPermute.inverse, the training path): input, weight and
            # bias gradients from ONE pass over (x, gy) -- vcnf_small_linear_bwd
            want_w = ctx.needs_input_grad[1]
            want_b = bias is not None and ctx.needs_input_grad[2]
            gx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
            gW = torch.empty((D, D), dtype=x.dtype, device=x.device) if want_w else None
            gb = torch.empty(D, dtype=x.dtype, device=x.device) if want_b else None
            ws = torch.empty(_lib.query("vcnf_small_linear_bwd_workspace", D, dtype_code(x)), dtype=torch.uint8,
                             device=x.device)
            _launch("vcnf_small_linear_bwd", ptr(x), ptr(gy), ptr(W), ptr(in_idx), ptr(out_idx), ptr(gx), ptr(gW),
                    ptr(gb), ptr(ws), rows, D, dtype_code(x), stream(), n=2)
            return gx, gW, gb, None, None, None
        if ctx.needs_input_grad[0] or (sub and bias is not None and ctx.needs_input_grad[2]):
            gx = torch.empty_like(x)
            Wt = W.t().contiguous()
            _launch("vcnf_small_linear", ptr(gy), ptr(gx), ptr(Wt), None, ptr(out_idx), ptr(in_idx), rows, D, S, 0,
                    dtype_code(x), stream())
        # D x D parameter gradients: a plain [D, rows] x [rows, D] library GEMM (not on the hot path's
        # byte budget; SURVEY.md 8d "LULinearPermute ... + D^2 param grads")
        if ctx.needs_input_grad[1] or (bias is not None and ctx.needs_input_grad[2]):
            if S != 1 and not sub and bias is None and rows > 0:
                # image tensors: both operands read in place by vcnf_small_linear_wgrad_img (no transposed copies)
                gW = torch.empty((D, D), dtype=x.dtype, device=x.device)
                ws = torch.empty(_lib.query("vcnf_small_linear_wgrad_img_workspace", D, dtype_code(x)), dtype=torch.uint8,
                                 device=x.device)
                _launch("vcnf_small_linear_wgrad_img", ptr(x), ptr(gy), ptr(in_idx), ptr(out_idx), ptr(gW), None, ptr(ws),
                        x.shape[0], D, S, dtype_code(x), stream(), n=2)
                return gx, gW, None, None, None, None
            if S == 1:
                xg = x if in_idx is None else x.index_select(1, in_idx.long())
                gg = gy if out_idx is None else gy.index_select(1, out_idx.long())
            else:
                xg = (x if in_idx is None else x.index_select(1, in_idx.long())).flatten(2).transpose(1, 2).reshape(-1, D)
                gg = (gy if out_idx is None else gy.index_select(1, out_idx.long())).flatten(2).transpose(1, 2).reshape(-1, D)
            if sub and bias is not None:
                xg = xg - bias
            if x.dtype == torch.float32 and not sub and rows > 0:
                # [D, rows] x [rows, D] with rows = batch: the split-K tcgen05 weight-gradient kernel
                # (cuBLAS picks a 64x64 SIMT tile here: 2.6 ms per cfg-2 step for sixteen 29 x 29 results)
                xg, gg = xg.contiguous(), gg.contiguous()
                gW = torch.empty((D, D), dtype=x.dtype, device=x.device)
                gb_ = torch.empty(D, dtype=x.dtype, device=x.device)
                ws = torch.empty(_lib.query("vcnf_linear_wgrad_workspace", rows, D, D), dtype=torch.uint8,
                                 device=x.device)
                _launch("vcnf_linear_wgrad_tf32x3", ptr(gg), D, ptr(xg), D, ptr(gW), ptr(gb_), ptr(ws), rows, D, D,
                        0, stream(), n=2)
                gb = gb_ if (bias is not None and ctx.needs_input_grad[2]) else None
                if not ctx.needs_input_grad[1]:
                    gW = None
                return gx, gW, gb, None, None, None
            if ctx.needs_input_grad[1]:
                gW = gg.t() @ xg
            if bias is not None and ctx.needs_input_grad[2]:
                gb = -(gg @ W).sum(0) if sub else gg.sum(0)
        return gx, gW, gb, None, None, None


SMALL_LINEAR_MAX = 64      # widest per-row map of vcnf_small_linear (W in shared memory, one row per thread group)


def _wide_linear(x, W, bias, in_idx, out_idx, sub_bias_first):
    """The same map for D > 64 (LULinearPermute on wide tabular data, Glow levels with > 64 channels): a dense
    contraction, so it runs on the conditioner GEMM (vcnf_linear_tf32x3, fp32) with the permutations as column
    gathers; fp64 operands use the library GEMM (no fp64 tensor-core path)."""
    D = W.shape[0]
    img = x.dim() == 4
    if img:
        B, _, H, Wd = x.shape
        rows = im2col(x, "nchw", B, D, H, Wd, 1) if x.dtype == torch.float32 else x.permute(0, 2, 3, 1).reshape(-1, D)
    else:
        rows = x
    if in_idx is not None:
        rows = gather_cols(rows, in_idx) if rows.dtype in (torch.float32, torch.float64) else rows[:, in_idx.long()]
    b_eff = bias
    if sub_bias_first and bias is not None:
        b_eff = -(W @ bias)                                        # W (x - b) = W x - W b
    if rows.dtype == torch.float32 and linear_supported(rows, W):
        out = linear(rows, W, b_eff)
    else:
        out = torch.nn.functional.linear(rows, W, b_eff)
    if out_idx is not None:                                        # out[:, out_idx[o]] = v[:, o]
        out = gather_cols(out, torch.argsort(out_idx.long()).to(torch.int32))
    if img:
        out = out.reshape(B, H, Wd, D).permute(0, 3, 1, 2).contiguous()
    return out


def small_linear(x, W, bias=None, in_idx=None, out_idx=None, sub_bias_first=False):
    x, W, bias = _promote(x, W, bias)
    if W.shape[0] > SMALL_LINEAR_MAX:
        return _wide_linear(x, W, bias, in_idx, out_idx, sub_bias_first)
    return _SmallLinear.apply(x, W, bias, in_idx, out_idx, sub_bias_first)


# ------------------------------------------------------------------ conditioner dense layers
class _LuCompose(torch.autograd.Function):
    """(lower_entries, upper_entries, unconstrained_upper_diag) -> (W = L U, log|det W|): one kernel each
    way instead of the index_put / softplus / matmul / log / sum chain (mixing.py:352-396)."""

    @staticmethod
    def forward(ctx, lower, upper, diag_raw, eps):
        lower, upper, diag_raw = lower.contiguous(), upper.contiguous(), diag_raw.contiguous()
        D = diag_raw.numel()
        W = torch.empty((D, D), dtype=diag_raw.dtype, device=diag_raw.device)
        logabsdet = torch.empty((), dtype=diag_raw.dtype, device=diag_raw.device)
        _launch("vcnf_lu_compose_fwd", ptr(lower), ptr(upper), ptr(diag_raw), float(eps), D, ptr(W), ptr(logabsdet),
                dtype_code(diag_raw), stream())
        ctx.eps = float(eps)
        ctx.save_for_backward(lower, upper, diag_raw)
        return W, logabsdet

    @staticmethod
    @once_differentiable
    def backward(ctx, gW, gl):
        lower, upper, diag_raw = ctx.saved_tensors
        D = diag_raw.numel()
        gW = gW.contiguous() if gW is not None else None
        gl = gl.contiguous() if gl is not None else None
        g_lower, g_upper, g_diag = torch.empty_like(lower), torch.empty_like(upper), torch.empty_like(diag_raw)
        _launch("vcnf_lu_compose_bwd", ptr(lower), ptr(upper), ptr(diag_raw), ctx.eps, D, ptr(gW), ptr(gl),
                ptr(g_lower), ptr(g_upper), ptr(g_diag), dtype_code(diag_raw), stream())
        return g_lower, g_upper, g_diag, None


class _LuComposeBatched(torch.autograd.Function):
    """_LuCompose for n layers at once (vcnf_lu_compose_fwd/bwd_batched, one CTA per layer): inputs
    (eps, n, lower_0..n-1, upper_0..n-1, diag_0..n-1), outputs (W_0..n-1, logabsdet_0..n-1).  The backward runs when the
    last of the 2n output gradients has arrived, i.e. once, after the first layer's backward."""

    @staticmethod
    def forward(ctx, eps, n, *params):
        params = [p.contiguous() for p in params]
        lower, upper, diag = params[:n], params[n:2 * n], params[2 * n:]
        D = diag[0].numel()
        dt, dev = diag[0].dtype, diag[0].device
        Ws = [torch.empty((D, D), dtype=dt, device=dev) for _ in range(n)]
        lds = [torch.empty((), dtype=dt, device=dev) for _ in range(n)]
        _launch("vcnf_lu_compose_fwd_batched", _ptr_array(lower), _ptr_array(upper), _ptr_array(diag), float(eps), D, n,
                _ptr_array(Ws), _ptr_array(lds), dtype_code(diag[0]), stream())
        ctx.cfg = (float(eps), n, D)
        ctx.save_for_backward(*params)
        return (*Ws, *lds)

    @staticmethod
    @once_differentiable
    def backward(ctx, *grads):
        eps, n, D = ctx.cfg
        params = ctx.saved_tensors
        lower, upper, diag = params[:n], params[n:2 * n], params[2 * n:]
        gW = [g.contiguous() if g is not None else None for g in grads[:n]]
        gl = [g.contiguous() if g is not None else None for g in grads[n:]]
        g_lower, g_upper = [torch.empty_like(p) for p in lower], [torch.empty_like(p) for p in upper]
        g_diag = [torch.empty_like(p) for p in diag]
        _launch("vcnf_lu_compose_bwd_batched", _ptr_array(lower), _ptr_array(upper), _ptr_array(diag), eps, D, n,
                _ptr_array(gW), _ptr_array(gl), _ptr_array(g_lower), _ptr_array(g_upper), _ptr_array(g_diag),
                dtype_code(diag[0]), stream())
        return (None, None, *g_lower, *g_upper, *g_diag)


LU_BATCH_MAX = 32


def lu_compose_batched(triples, eps):
    """[(lower_entries, upper_entries, unconstrained_upper_diag)] * n -> [(W, logabsdet)] * n; equal D, one dtype,
    n <= LU_BATCH_MAX."""
    n = len(triples)
    out = _LuComposeBatched.apply(float(eps), n, *[t[0] for t in triples], *[t[1] for t in triples],
                                  *[t[2] for t in triples])
    return list(zip(out[:n], out[n:]))


def lu_compose(lower_entries, upper_entries, unconstrained_upper_diag, eps):
    """W = L U and log|det W| of the LU-parametrised linear flow (D <= 64)."""
    lower_entries, upper_entries, unconstrained_upper_diag = _promote(lower_entries, upper_entries, unconstrained_upper_diag)
    return _LuCompose.apply(lower_entries, upper_entries, unconstrained_upper_diag, eps)


class _GlowW(torch.autograd.Function):
    """W = P (tril(L,-1) + I) (triu(U,1) + diag(sign_S exp(log_S))) and ld_scale * sum(log_S) (mixing.py:92-102,117-118)
    in one launch each way (vcnf_glow_w_fwd / _bwd)."""

    @staticmethod
    def forward(ctx, L, U, log_S, P, sign_S, ld_scale):
        L, U, log_S, P, sign_S = (t.contiguous() for t in (L, U, log_S, P, sign_S))
        C = L.shape[0]
        W = torch.empty((C, C), dtype=L.dtype, device=L.device)
        ld = torch.empty((), dtype=L.dtype, device=L.device)
        _launch("vcnf_glow_w_fwd", ptr(P), ptr(L), ptr(U), ptr(sign_S), ptr(log_S), None, C, float(ld_scale), ptr(W), None,
                ptr(ld), dtype_code(L), stream())
        ctx.ld_scale = float(ld_scale)
        ctx.save_for_backward(L, U, log_S, P, sign_S)
        return W, ld

    @staticmethod
    @once_differentiable
    def backward(ctx, gW, gld):
        L, U, log_S, P, sign_S = ctx.saved_tensors
        C = L.shape[0]
        gW = gW.contiguous() if gW is not None else None
        gld = gld.contiguous() if gld is not None else None
        gL, gU, gS = torch.empty_like(L), torch.empty_like(U), torch.empty_like(log_S)
        _launch("vcnf_glow_w_bwd", ptr(P), ptr(L), ptr(U), ptr(sign_S), ptr(log_S), None, None, None, None, C, ctx.ld_scale,
                ptr(gW), ptr(gld), ptr(gL), ptr(gU), ptr(gS), None, None, dtype_code(L), stream())
        return gL, gU, gS, None, None, None


GLOW_W_MAX = 64


def glow_w(L, U, log_S, P, sign_S, ld_scale):
    """(W [C, C], ld_scale * sum(log_S)) of Glow's LU-parametrised 1x1 convolution; C <= GLOW_W_MAX, one dtype."""
    return _GlowW.apply(L, U, log_S, P, sign_S, float(ld_scale))


class _GlowActNormConv(torch.autograd.Function):
    """ActNorm.inverse followed by Invertible1x1Conv.inverse (the density direction of a GlowBlock,
    flows/affine/glow.py:59-75): y = W ((z - t) e^-s) per pixel and the 0-dim log-det H W (sum log_S - sum s).
    The two per-pixel affine maps are folded into one weight by vcnf_glow_w_fwd, applied by vcnf_small_linear; the
    backward is the transposed map, vcnf_small_linear_wgrad_img and vcnf_glow_w_bwd (all parameter gradients of both
    layers): 2 + 5 launches instead of 5 + 10, one pass over z each way instead of two."""

    @staticmethod
    def forward(ctx, z, L, U, log_S, s, t, P, sign_S):
        z = z.contiguous()
        L, U, log_S, P, sign_S = (x.contiguous() for x in (L, U, log_S, P, sign_S))
        s1, t1 = s.reshape(-1).contiguous(), t.reshape(-1).contiguous()
        B, C = z.shape[0], z.shape[1]
        S = math.prod(z.shape[2:])
        W2 = torch.empty((C, C), dtype=L.dtype, device=L.device)
        W2t = torch.empty((C, C), dtype=L.dtype, device=L.device)
        ld = torch.empty((), dtype=L.dtype, device=L.device)
        _launch("vcnf_glow_w_fwd", ptr(P), ptr(L), ptr(U), ptr(sign_S), ptr(log_S), ptr(s1), C, float(S), ptr(W2), ptr(W2t),
                ptr(ld), dtype_code(L), stream())
        y = torch.empty_like(z)
        _launch("vcnf_small_linear", ptr(z), ptr(y), ptr(W2), ptr(t1), None, None, B * S, C, S, 1, dtype_code(z), stream())
        ctx.dims = (B, C, S)
        ctx.pshape = (s.shape, t.shape)
        ctx.save_for_backward(z, L, U, log_S, s1, t1, P, sign_S, W2, W2t)
        return y, ld

    @staticmethod
    @once_differentiable
    def backward(ctx, gy, gld):
        z, L, U, log_S, s1, t1, P, sign_S, W2, W2t = ctx.saved_tensors
        B, C, S = ctx.dims
        gy = gy.contiguous()
        gld = gld.contiguous() if gld is not None else None
        gz = None
        if ctx.needs_input_grad[0]:
            gz = torch.empty_like(z)
            _launch("vcnf_small_linear", ptr(gy), ptr(gz), ptr(W2t), None, None, None, B * S, C, S, 0, dtype_code(z),
                    stream())
        G = torch.empty((C, C), dtype=z.dtype, device=z.device)
        gb = torch.empty(C, dtype=z.dtype, device=z.device)
        ws = torch.empty(_lib.query("vcnf_small_linear_wgrad_img_workspace", C, dtype_code(z)), dtype=torch.uint8,
                         device=z.device)
        _launch("vcnf_small_linear_wgrad_img", ptr(z), ptr(gy), None, None, ptr(G), ptr(gb), ptr(ws), B, C, S,
                dtype_code(z), stream(), n=3)
        gL, gU, gS = torch.empty_like(L), torch.empty_like(U), torch.empty_like(log_S)
        gs, gt = torch.empty_like(s1), torch.empty_like(t1)
        _launch("vcnf_glow_w_bwd", ptr(P), ptr(L), ptr(U), ptr(sign_S), ptr(log_S), ptr(s1), ptr(t1), ptr(W2), ptr(gb), C,
                float(S), ptr(G), ptr(gld), ptr(gL), ptr(gU), ptr(gS), ptr(gs), ptr(gt), dtype_code(L), stream())
        return gz, gL, gU, gS, gs.view(ctx.pshape[0]), gt.view(ctx.pshape[1]), None, None


def glow_actnorm_conv(z, L, U, log_S, s, t, P, sign_S):
    """(W ((z - t) e^-s) per pixel, H W (sum log_S - sum s)): ActNorm.inverse + Invertible1x1Conv.inverse in one node."""
    return _GlowActNormConv.apply(z, L, U, log_S, s, t, P, sign_S)


def lu_inverse(lower_entries, upper_entries, unconstrained_upper_diag, eps):
    """(L U)^-1 and log|det L U| in one launch (sampling direction, no gradient; D <= 64)."""
    lo, up, dg = _promote(lower_entries.detach(), upper_entries.detach(), unconstrained_upper_diag.detach())
    lo, up, dg = lo.contiguous(), up.contiguous(), dg.contiguous()
    D = dg.shape[0]
    winv = torch.empty((D, D), dtype=dg.dtype, device=dg.device)
    logabsdet = torch.empty((), dtype=dg.dtype, device=dg.device)
    _launch("vcnf_lu_inverse", ptr(lo), ptr(up), ptr(dg), float(eps), D, ptr(winv), ptr(logabsdet), dtype_code(dg),
            stream())
    return winv, logabsdet


def lu_inverse_batched(triples, eps):
    """[(lower_entries, upper_entries, unconstrained_upper_diag)] * n -> [((L U)^-1, log|det L U|)] * n in one launch
    (no gradient; equal D, one dtype, n <= LU_BATCH_MAX)."""
    n = len(triples)
    flat = _promote(*[t.detach() for tr in triples for t in tr])
    lower = [flat[3 * i].contiguous() for i in range(n)]
    upper = [flat[3 * i + 1].contiguous() for i in range(n)]
    diag = [flat[3 * i + 2].contiguous() for i in range(n)]
    D = diag[0].numel()
    dt, dev = diag[0].dtype, diag[0].device
    winv = torch.empty((n, D, D), dtype=dt, device=dev)
    lds = torch.empty((n,), dtype=dt, device=dev)
    _launch("vcnf_lu_inverse_batched", _ptr_array(lower), _ptr_array(upper), _ptr_array(diag), float(eps), D, n,
            _ptr_array([winv[i] for i in range(n)]), _ptr_array([lds[i] for i in range(n)]), dtype_code(diag[0]), stream())
    return [(winv[i], lds[i]) for i in range(n)], lds


class _BroadcastScalar(torch.autograd.Function):
    """A 0-dim tensor repeated over the batch (`logabsdet * inputs.new_ones(batch_size)`, mixing.py:410,429):
    one copy kernel forward instead of fill + multiply; the adjoint (sum of the n incoming gradients) as two
    row reductions instead of the library's single-CTA reduction to a scalar (19 us at n = 65 536)."""

    @staticmethod
    def forward(ctx, s, n):
        ctx.shape = s.shape
        return s.detach().reshape(1).expand(n).contiguous()

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        n = g.shape[0]
        r = 256
        while r > 1 and n % r:
            r //= 2
        total = g.reshape(r, n // r).sum(1).sum() if (r > 1 and n >= 4096) else g.sum()
        return total.reshape(ctx.shape), None


def broadcast_scalar(s, n):
    return _BroadcastScalar.apply(s, int(n))


# ------------------------------------------------------------------ small MLP conditioner (registers)
MLP_MAX_WIDTH, MLP_MAX_LAYERS = 32, 4


def mlp_small_supported(x, weights, slope):
    """vcnf_mlp_fwd / _bwd: CUDA fp32 / fp64 rows, <= 4 linear layers, every width <= 32, LeakyReLU slope >= 0."""
    return (x.is_cuda and x.dim() == 2 and x.dtype in (torch.float32, torch.float64) and slope >= 0.0
            and 1 <= len(weights) <= MLP_MAX_LAYERS and all(w.dtype == x.dtype and w.dim() == 2 for w in weights)
            and x.shape[1] == weights[0].shape[1]
            and all(max(w.shape) <= MLP_MAX_WIDTH for w in weights)
            # shared memory of the backward (224 KB limit; csrc/mlp_kernels.cu mlp_bwd_elems): both weight layouts,
            # per-warp staging tiles, 4 per-warp copies of the parameter gradient, the hidden activations of a tile
            and (2080 * len(weights) + 8320 + 4096 * (len(weights) - 1)
                 + 4 * sum(w.numel() + w.shape[0] for w in weights)) * x.element_size() <= 224 * 1024)


def _ptr_array(tensors):
    import ctypes
    return (ctypes.c_void_p * len(tensors))(*[ptr(t) for t in tensors])


class _MlpSmall(torch.autograd.Function):
    """y = Linear_n(act(... act(Linear_1(x)))) with act = LeakyReLU(slope): one kernel forward, one kernel (plus
    the fixed-order reduce) backward.  args = (x, slope, n, W_1..W_n, b_1..b_n) with None for missing biases."""

    @staticmethod
    def forward(ctx, x, slope, n, *wb):
        import ctypes
        weights, biases = [w.contiguous() for w in wb[:n]], [b.contiguous() if b is not None else None for b in wb[n:]]
        x = _rows(x)
        B = x.shape[0]
        dims = [weights[0].shape[1]] + [w.shape[0] for w in weights]
        cdims = (ctypes.c_int32 * len(dims))(*dims)
        y = torch.empty((B, dims[-1]), dtype=x.dtype, device=x.device)
        _launch("vcnf_mlp_fwd", ptr(x), x.stride(0), ptr(y), y.stride(0), _ptr_array(weights), _ptr_array(biases),
                cdims, n, float(slope), B, dtype_code(x), stream())
        ctx.cfg = (float(slope), n, dims)
        ctx.has_bias = [b is not None for b in biases]
        ctx.save_for_backward(x, *weights, *[b for b in biases if b is not None])
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, gy):
        import ctypes
        slope, n, dims = ctx.cfg
        saved = ctx.saved_tensors
        x, weights = saved[0], list(saved[1:1 + n])
        it = iter(saved[1 + n:])
        biases = [next(it) if hb else None for hb in ctx.has_bias]
        gy = _rows(gy)
        B = x.shape[0]
        cdims = (ctypes.c_int32 * len(dims))(*dims)
        psize = sum(dims[l + 1] * dims[l] + dims[l + 1] for l in range(n))
        gx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        gparams = torch.empty(psize, dtype=x.dtype, device=x.device)
        ws = torch.empty(max(1, _lib.query("vcnf_mlp_bwd_workspace", cdims, n, dtype_code(x))), dtype=torch.uint8,
                         device=x.device)
        _launch("vcnf_mlp_bwd", ptr(x), x.stride(0), ptr(gy), gy.stride(0), _ptr_array(weights), _ptr_array(biases),
                cdims, n, slope, ptr(gx), gx.stride(0) if gx is not None else 0, ptr(gparams), ptr(ws), B,
                dtype_code(x), stream(), n=2)
        gws, gbs, off = [], [], 0
        for l in range(n):
            nw = dims[l + 1] * dims[l]
            gws.append(gparams[off:off + nw].view(dims[l + 1], dims[l]))
            gbs.append(gparams[off + nw:off + nw + dims[l + 1]] if ctx.has_bias[l] else None)
            off += nw + dims[l + 1]
        return (gx, None, None, *gws, *gbs)


def mlp_small(x, weights, biases, slope):
    return _MlpSmall.apply(x, float(slope), len(weights), *weights, *biases)


KSLICE = 128   # contraction width of one vcnf_linear_tf32x3 / vcnf_linear_wgrad_tf32x3 launch (W resident in TMEM)


def linear_supported(x, weight):
    """The tcgen05 3xTF32 kernels cover fp32 CUDA row-major operands; in_features > 128 (MADE, hidden 256)
    run as column slices of 128 that accumulate in place through the kernel's `residual` operand."""
    return (x.is_cuda and x.dtype == torch.float32 and weight.dtype == torch.float32 and x.dim() == 2
            and weight.dim() == 2 and x.shape[1] == weight.shape[1] and (weight.shape[1] <= KSLICE or weight.shape[1] % 4 == 0))


def resblock_supported(x, weight):
    """`residual_block` (one autograd node, four launches) needs square weights no wider than one slice."""
    return linear_supported(x, weight) and weight.shape[0] == weight.shape[1] <= KSLICE


class _Linear(torch.autograd.Function):
    """y = act(x) W^T + b (+ residual), act = ReLU when relu_in (pre-activation residual blocks,
    nets/resnet.py:49-55).  Forward and input-gradient run `vcnf_linear_tf32x3` (tcgen05, fp32-faithful);
    the ReLU, the bias, the residual add and the ReLU-backward mask are fused into the kernels."""

    @staticmethod
    def forward(ctx, x, weight, bias, residual, relu_in, pad_out, pre=None):
        # anything but the Parameter objects themselves may have been written by a kernel earlier on this stream
        # (MADE's W * mask, a re-laid convolution weight): the GEMM then fetches W after its dependency wait
        is_param = lambda t: t is None or isinstance(t, torch.nn.Parameter)
        w_dyn = 0 if (is_param(weight) and is_param(bias)) else 2
        x, weight = _rows(x), weight.contiguous()
        M, K = x.shape
        N = weight.shape[0]
        bias_ = bias.contiguous() if bias is not None else None
        res_ = _rows(residual) if residual is not None else None
        if pre is not None:            # output already computed by the fused net + spline kernel: only the graph node
            ctx.relu_in, ctx.w_dyn = bool(relu_in), w_dyn
            ctx.has_bias, ctx.has_res = bias is not None, residual is not None
            ctx.save_for_backward(x, weight)
            return pre
        y = _padded_empty(M, N, x.dtype, x.device) if pad_out else torch.empty((M, N), dtype=x.dtype, device=x.device)
        for k0 in range(0, K, KSLICE):
            # in_features > 128: slices of the contraction accumulate in place (`residual` = y)
            first = k0 == 0
            r_ptr, r_ld = (ptr(res_), res_.stride(0) if res_ is not None else 0) if first else (ptr(y), y.stride(0))
            _launch("vcnf_linear_tf32x3", ptr(x) + 4 * k0, x.stride(0), ptr(weight) + 4 * k0, K, 1,
                    ptr(bias_) if first else None, r_ptr, r_ld, None, 0,
                    ptr(y), y.stride(0), M, N, min(K, k0 + KSLICE) - k0, int(relu_in) | w_dyn, stream())
        ctx.relu_in = bool(relu_in)
        ctx.w_dyn = w_dyn
        ctx.has_bias, ctx.has_res = bias is not None, residual is not None
        ctx.save_for_backward(x, weight)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, gy):
        x, weight = ctx.saved_tensors
        gy = _rows(gy)
        M, K = x.shape
        N = weight.shape[0]
        ldg, ldx = gy.stride(0), x.stride(0)
        gx = gw = gb = gres = None
        if ctx.needs_input_grad[0]:
            gx = _padded_empty(M, K, x.dtype, x.device) if ldx != K else torch.empty_like(x)
            ldo = gx.stride(0)
            # gx[m, k] = [x > 0] * sum_n gy[m, n] W[n, k]  : W read as [n' = k, k' = n] -> sw_n = 1, sw_k = K
            if N <= 128:
                _launch("vcnf_linear_tf32x3", ptr(gy), ldg, ptr(weight), 1, K, None, None, 0,
                        ptr(x) if ctx.relu_in else None, ldx, ptr(gx), ldo, M, K, N, ctx.w_dyn, stream())
            else:
                # out_features > 128 (the 3K-1 wide head): reduce over column slices of <= 128
                # (the ReLU mask rides on the LAST slice: flag 4 = mask after the running sum)
                first = True
                for n0 in range(0, N, 128):
                    n1 = min(N, n0 + 128)
                    last = n1 == N
                    _launch("vcnf_linear_tf32x3", ptr(gy) + 4 * n0, ldg, ptr(weight) + 4 * n0 * K, 1, K, None,
                            None if first else ptr(gx), ldo, ptr(x) if (ctx.relu_in and last) else None, ldx,
                            ptr(gx), ldo, M, K, n1 - n0, ctx.w_dyn | (4 if (ctx.relu_in and last) else 0), stream())
                    first = False
        want_w = ctx.needs_input_grad[1]
        want_b = ctx.has_bias and ctx.needs_input_grad[2]
        if want_w or want_b:
            # gW = gy^T act(x), gb = column sums of gy: one split-K tcgen05 kernel + a fixed-order reduce
            gb = torch.empty(N, dtype=x.dtype, device=x.device) if want_b else None
            ws = torch.empty(_lib.query("vcnf_linear_wgrad_workspace", M, N, min(K, KSLICE)), dtype=torch.uint8,
                             device=x.device)
            if K <= KSLICE:
                gw = torch.empty_like(weight)
                _launch("vcnf_linear_wgrad_tf32x3", ptr(gy), ldg, ptr(x), ldx, ptr(gw), ptr(gb), ptr(ws), M, N, K,
                        int(ctx.relu_in), stream(), n=2)
            else:
                # in_features > 128: one launch per slice of input columns, gW slices joined afterwards
                parts = []
                for k0 in range(0, K, KSLICE):
                    kw = min(K, k0 + KSLICE) - k0
                    part = torch.empty((N, kw), dtype=x.dtype, device=x.device)
                    _launch("vcnf_linear_wgrad_tf32x3", ptr(gy), ldg, ptr(x) + 4 * k0, ldx, ptr(part),
                            ptr(gb) if k0 == 0 else None, ptr(ws), M, N, kw, int(ctx.relu_in), stream(), n=2)
                    parts.append(part)
                gw = torch.cat(parts, dim=1)
            if not want_w:
                gw = None
        if ctx.has_res and ctx.needs_input_grad[3]:
            gres = gy
        return gx, gw, gb, gres, None, None, None


def linear(x, weight, bias=None, residual=None, relu_in=False, pad_out=False, pre=None):
    """pad_out=True: the result is a view with a 16-byte-multiple row pitch (see _padded_empty)."""
    return _Linear.apply(x, weight, bias, residual, relu_in, pad_out, pre)


class _ResBlock(torch.autograd.Function):
    """Pre-activation residual block  out = x + W1 relu(W0 relu(x) + b0) + b1  (nets/resnet.py:38-55 with
    ReLU, no batch norm / dropout / context) as ONE autograd node: two forward GEMM launches, and in the
    backward the skip-connection gradient rides on the `residual` operand of the second input-gradient
    GEMM instead of a separate [B, features] add kernel (32 such adds per cfg-2 step)."""

    @staticmethod
    def forward(ctx, x, w0, b0, w1, b1, pre=None):
        x, w0, w1 = _rows(x), w0.contiguous(), w1.contiguous()
        M, F_ = x.shape
        if pre is not None:            # (t, out) already computed by the fused net + spline kernel
            ctx.save_for_backward(x, pre[0], w0, w1)
            return pre[1]
        t = torch.empty((M, F_), dtype=x.dtype, device=x.device)
        out = torch.empty((M, F_), dtype=x.dtype, device=x.device)
        _launch("vcnf_linear_tf32x3", ptr(x), x.stride(0), ptr(w0), F_, 1, ptr(b0), None, 0, None, 0,
                ptr(t), F_, M, F_, F_, 1, stream())
        _launch("vcnf_linear_tf32x3", ptr(t), F_, ptr(w1), F_, 1, ptr(b1), ptr(x), x.stride(0), None, 0,
                ptr(out), F_, M, F_, F_, 1, stream())
        ctx.save_for_backward(x, t, w0, w1)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        x, t, w0, w1 = ctx.saved_tensors
        g = _rows(g)
        M, F_ = x.shape
        ldg, ldx = g.stride(0), x.stride(0)
        need_x, need_w0, need_b0, need_w1, need_b1 = ctx.needs_input_grad[:5]
        gx = gw0 = gb0 = gw1 = gb1 = None
        # g_t = [t > 0] * (g W1)
        gt = torch.empty((M, F_), dtype=x.dtype, device=x.device)
        _launch("vcnf_linear_tf32x3", ptr(g), ldg, ptr(w1), 1, F_, None, None, 0, ptr(t), F_, ptr(gt), F_,
                M, F_, F_, 0, stream())
        if need_x:   # g_x = [x > 0] * (g_t W0) + g   (skip connection through `residual`)
            gx = torch.empty((M, F_), dtype=x.dtype, device=x.device)
            _launch("vcnf_linear_tf32x3", ptr(gt), F_, ptr(w0), 1, F_, None, ptr(g), ldg, ptr(x), ldx, ptr(gx), F_,
                    M, F_, F_, 0, stream())
        nbytes = _lib.query("vcnf_linear_wgrad_workspace", M, F_, F_)
        if need_w1 or need_b1:
            gw1 = torch.empty_like(w1)
            gb1 = torch.empty(F_, dtype=x.dtype, device=x.device)
            ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
            _launch("vcnf_linear_wgrad_tf32x3", ptr(g), ldg, ptr(t), F_, ptr(gw1), ptr(gb1), ptr(ws), M, F_, F_, 1,
                    stream(), n=2)
        if need_w0 or need_b0:
            gw0 = torch.empty_like(w0)
            gb0 = torch.empty(F_, dtype=x.dtype, device=x.device)
            ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
            _launch("vcnf_linear_wgrad_tf32x3", ptr(gt), F_, ptr(x), ldx, ptr(gw0), ptr(gb0), ptr(ws), M, F_, F_, 1,
                    stream(), n=2)
        return (gx, gw0 if need_w0 else None, gb0 if need_b0 else None, gw1 if need_w1 else None,
                gb1 if need_b1 else None, None)


def residual_block(x, w0, b0, w1, b1, pre=None):
    """x + Linear1(relu(Linear0(relu(x)))) with square weights [F, F], F <= 128, biases required."""
    return _ResBlock.apply(x, w0, b0, w1, b1, pre)


# ------------------------------------------------------------------ convolutional conditioner (Glow, K9)
def _ceil_to(n, q):
    return (n + q - 1) // q * q


def _src_strides(x, layout, H, W):
    """Element strides (sb, sc, sh, sw) of an image batch held as NCHW [B, C, H, W] or as NHWC rows [B*H*W, C]
    (any row pitch)."""
    if layout == "nchw":
        return x.stride(0), x.stride(1), x.stride(2), x.stride(3)
    ld = x.stride(0)
    return H * W * ld, 1, W * ld, ld


def im2col_raw(x, layout, B, C, H, W, ksize, Kp, flip=False, act=False, slope=0.0, bf16=False):
    """Patch matrix [B*H*W, Kp] of vcnf_im2col (no autograd)."""
    P = B * H * W
    col = (torch.empty((P, Kp), dtype=torch.bfloat16, device=x.device) if bf16
           else _padded_empty(P, Kp, torch.float32, x.device))
    sb, sc, sh, sw = _src_strides(x, layout, H, W)
    _launch("vcnf_im2col", ptr(x), sb, sc, sh, sw, ptr(col), col.stride(0), B, C, H, W, ksize, Kp, int(flip), int(act),
            float(slope), int(bf16), stream())
    return col


class _Im2Col(torch.autograd.Function):
    """fp32 patch matrix of a stride-1 'same' convolution (ksize 1 or 3) with LeakyReLU(slope) optionally applied
    while reading; backward = vcnf_col2im (gather form)."""

    @staticmethod
    def forward(ctx, x, layout, B, C, H, W, ksize, Kp, act, slope):
        if layout == "nchw":
            x = x.contiguous()
        else:
            x = _rows(x)
        ctx.cfg = (layout, B, C, H, W, ksize, bool(act), float(slope))
        ctx.save_for_backward(x)
        return im2col_raw(x, layout, B, C, H, W, ksize, Kp, act=act, slope=slope)

    @staticmethod
    @once_differentiable
    def backward(ctx, gcol):
        (x,) = ctx.saved_tensors
        layout, B, C, H, W, ksize, act, slope = ctx.cfg
        gcol = _rows(gcol)
        if layout == "nchw":
            gx = torch.empty_like(x)
        else:
            gx = _padded_empty(x.shape[0], C, x.dtype, x.device) if x.stride(0) != C else torch.empty_like(x)
        sb, sc, sh, sw = _src_strides(gx, layout, H, W)
        assert (sb, sc, sh, sw) == _src_strides(x, layout, H, W)
        _launch("vcnf_col2im", ptr(gcol), gcol.stride(0), ptr(x) if act else None, ptr(gx), sb, sc, sh, sw, B, C, H, W,
                ksize, 0, int(act), slope, stream())
        return gx, None, None, None, None, None, None, None, None, None


def im2col(x, layout, B, C, H, W, ksize, Kp=None, act=False, slope=0.0):
    return _Im2Col.apply(x, layout, B, C, H, W, ksize, Kp if Kp is not None else ksize * ksize * C, act, slope)


def colsum(x, layout, rows, C, ld_or_S):
    """fp32 [C] column sums (bias gradients): layout 0 = bf16 rows [rows, ld], 1 = fp32 NCHW [rows, C, S]."""
    out = torch.empty(C, dtype=torch.float32, device=x.device)
    ws = torch.empty(_lib.query("vcnf_colsum_workspace", C), dtype=torch.uint8, device=x.device)
    _launch("vcnf_colsum", ptr(x), layout, rows, C, ld_or_S, ptr(out), ptr(ws), stream(), n=2)
    return out


def conv_gemm_bf16(act, B, H, W, taps, Wp, bias=None, mask=None, slope=0.0, epi=0, out_nchw_channels=None):
    """vcnf_conv_gemm_bf16: act [B*H*W, Cp] bf16 (NHWC), Wp [taps, Np, Cp] bf16.  Returns NHWC bf16 [B*H*W, Np], or
    NCHW fp32 [B, Cout, H, W] when out_nchw_channels = Cout."""
    Cp, Np = act.shape[1], Wp.shape[1]
    assert act.is_contiguous() and Wp.is_contiguous() and Wp.shape[0] == taps and Wp.shape[2] == Cp
    if out_nchw_channels is None:
        out = torch.empty((B * H * W, Np), dtype=torch.bfloat16, device=act.device)
        mode, cout, ldo = 0, Np, Np
    else:
        cout = int(out_nchw_channels)
        out = torch.empty((B, cout, H, W), dtype=torch.float32, device=act.device)
        mode, ldo = 1, 0
    _launch("vcnf_conv_gemm_bf16", ptr(act), B, H, W, Cp, taps, ptr(Wp), Np, ptr(bias), ptr(mask),
            mask.stride(0) if mask is not None else 0, float(slope), epi, ptr(out), mode, cout, ldo, stream())
    return out


def conv_wgrad_bf16(X, Y):
    """G [Mx, Ny] fp32 = X^T Y over the pixel rows (X [P, Mx], Y [P, Ny] bf16)."""
    P, Mx = X.shape
    Ny = Y.shape[1]
    G = torch.empty((Mx, Ny), dtype=torch.float32, device=X.device)
    ws = torch.empty(max(16, _lib.query("vcnf_conv_wgrad_workspace", P, Mx, Ny)), dtype=torch.uint8, device=X.device)
    _launch("vcnf_conv_wgrad_bf16", ptr(X), X.stride(0), Mx, ptr(Y), Y.stride(0), Ny, P, ptr(G), ptr(ws), stream(), n=2)
    return G


class _ConvNet3Bf16(torch.autograd.Function):
    """conv3x3 -> LeakyReLU -> conv1x1 -> LeakyReLU -> conv3x3 (normflow/nets/cnn.py:29-47 as GlowBlock builds it,
    flows/affine/glow.py:24-34) with bf16 operands on the tcgen05 tensor cores, fp32 accumulation, fp32 in / out
    (NCHW).  Activations between the layers are NHWC bf16 and never leave that layout; the 3x3 convolution over the
    hidden channels is nine shifted TMA loads per tile (no patch matrix); the two few-channel ends (coupling input,
    gradient of the parameter output) are laid out as patch matrices by the pack kernel that converts them anyway."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, w3, b3, slope):
        B, C0, H, W = x.shape                                      # any NCHW strides (a channel slice is read in place)
        Hd, C3 = w1.shape[0], w3.shape[0]
        bf = torch.bfloat16
        K1p, Np3, K3p, Np1 = _ceil_to(9 * C0, 64), _ceil_to(C3, 16), _ceil_to(9 * C3, 64), _ceil_to(C0, 16)
        a1 = im2col_raw(x, "nchw", B, C0, H, W, 3, K1p, bf16=True)
        # every bf16 weight layout of the three convolutions (forward and input-gradient operands) in one launch
        with_bwd = any(ctx.needs_input_grad[:7])
        f32 = lambda t: t.detach().float().contiguous()                # noqa: E731  (no-op for fp32 parameters)
        pack = torch.empty(_lib.query("vcnf_convnet3_pack_elems", C0, Hd, C3, int(with_bwd)), dtype=bf, device=x.device)
        b3p = torch.empty(Np3, dtype=torch.float32, device=x.device)
        _launch("vcnf_convnet3_pack", ptr(f32(w1)), ptr(f32(w2)), ptr(f32(w3)), ptr(f32(b3)) if b3 is not None else None,
                C0, Hd, C3, int(with_bwd), ptr(pack), ptr(b3p), stream())
        sizes = [Hd * K1p, Hd * Hd, 9 * Np3 * Hd] + ([Hd * K3p, Hd * Hd, 9 * Np1 * Hd] if with_bwd else [])
        segs = torch.split(pack, sizes)
        w1p, w2p, w3p = segs[0].view(1, Hd, K1p), segs[1].view(1, Hd, Hd), segs[2].view(9, Np3, Hd)
        h1 = conv_gemm_bf16(a1, B, H, W, 1, w1p, f32(b1), slope=slope, epi=1)
        h2 = conv_gemm_bf16(h1, B, H, W, 1, w2p, f32(b2), slope=slope, epi=1)
        out = conv_gemm_bf16(h2, B, H, W, 9, w3p, b3p, out_nchw_channels=C3)
        ctx.packs = (segs[3].view(1, Hd, K3p), segs[4].view(1, Hd, Hd), segs[5].view(9, Np1, Hd)) if with_bwd else None
        ctx.slope = float(slope)
        ctx.dims = (B, C0, H, W, Hd, C3, K1p)
        ctx.save_for_backward(a1, h1, h2, w1, w2, w3)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        a1, h1, h2, w1, w2, w3 = ctx.saved_tensors
        B, C0, H, W, Hd, C3, K1p = ctx.dims
        slope, bf = ctx.slope, torch.bfloat16
        P = B * H * W
        g = g.contiguous().float()
        need = ctx.needs_input_grad
        pad = torch.nn.functional.pad
        K3p = _ceil_to(9 * C3, 64)
        wd3, wd2, wd1 = ctx.packs
        a3g = im2col_raw(g, "nchw", B, C3, H, W, 3, K3p, flip=True, bf16=True)
        gh2 = conv_gemm_bf16(a3g, B, H, W, 1, wd3, mask=h2, slope=slope, epi=2)
        gh1 = conv_gemm_bf16(gh2, B, H, W, 1, wd2, mask=h1, slope=slope, epi=2)
        gx = gw1 = gb1 = gw2 = gb2 = gw3 = gb3 = None
        if need[0]:
            gx = conv_gemm_bf16(gh1, B, H, W, 9, wd1, out_nchw_channels=C0)
        if need[5]:
            G3 = conv_wgrad_bf16(a3g, h2)
            gw3 = G3[:9 * C3].view(3, 3, C3, Hd).permute(2, 3, 0, 1).contiguous().to(w3.dtype)
        if need[6]:
            gb3 = colsum(g, 1, B, C3, H * W).to(w3.dtype)
        if need[3]:
            gw2 = conv_wgrad_bf16(gh2, h1).view(Hd, Hd, 1, 1).to(w2.dtype)
        if need[4]:
            gb2 = colsum(gh2, 0, P, Hd, Hd).to(w2.dtype)
        if need[1]:
            G1 = conv_wgrad_bf16(gh1, a1)
            gw1 = G1[:, :9 * C0].view(Hd, 3, 3, C0).permute(0, 3, 1, 2).contiguous().to(w1.dtype)
        if need[2]:
            gb1 = colsum(gh1, 0, P, Hd, Hd).to(w1.dtype)
        return gx, gw1, gb1, gw2, gb2, gw3, gb3, None


def convnet3_bf16(x, w1, b1, w2, b2, w3, b3, slope):
    return _ConvNet3Bf16.apply(x, w1, b1, w2, b2, w3, b3, float(slope))


# ------------------------------------------------------------------ variational dequantisation (K11)
class _VdeqFused(torch.autograd.Function):
    """All categorical columns through their dequantisers in one launch (vcnf_vdeq_fwd); the gradient of the
    per-layer (log-scale, shift) table comes from vcnf_vdeq_bwd."""

    @staticmethod
    def forward(ctx, xcat, noise, tab, quants, alpha):
        xcat = _rows(xcat)
        B, n = xcat.shape
        L = tab.shape[1]
        noise, tab = noise.contiguous(), tab.contiguous()
        zc = torch.empty((B, n), dtype=xcat.dtype, device=xcat.device)
        logq = torch.empty(B, dtype=torch.float32, device=xcat.device)
        _launch("vcnf_vdeq_fwd", ptr(xcat), xcat.stride(0), ptr(noise), ptr(tab), ptr(quants), float(alpha), B, n, L,
                ptr(zc), ptr(logq), None, dtype_code(xcat), stream())
        ctx.alpha = float(alpha)
        ctx.save_for_backward(xcat, noise, tab, quants)
        return zc, logq

    @staticmethod
    @once_differentiable
    def backward(ctx, g_zc, g_logq):
        xcat, noise, tab, quants = ctx.saved_tensors
        B, n = xcat.shape
        L = tab.shape[1]
        g_zc = g_zc.contiguous().to(xcat.dtype) if g_zc is not None else None
        g_logq = g_logq.contiguous().float() if g_logq is not None else None
        g_tab = torch.empty_like(tab)
        ws = torch.empty(max(16, _lib.query("vcnf_vdeq_bwd_workspace", n, L)), dtype=torch.uint8, device=xcat.device)
        _launch("vcnf_vdeq_bwd", ptr(xcat), xcat.stride(0), ptr(noise), ptr(tab), ptr(quants), ctx.alpha, B, n, L,
                ptr(g_zc), ptr(g_logq), ptr(g_tab), ptr(ws), dtype_code(xcat), stream(), n=2)
        return None, None, g_tab, None, None


def vdeq_fused(xcat, noise, tab, quants, alpha):
    """(dequantised logits [B, n], triangularly accumulated log-det [B] fp32); tab [n, L, 2] fp64."""
    return _VdeqFused.apply(xcat, noise, tab, quants, alpha)


def vdeq_decode(zc, quants, log_q):
    """Sampling direction for all categorical columns (no gradient); log_q is updated in place."""
    zc = _rows(zc)
    B, n = zc.shape
    out = torch.empty((B, n), dtype=zc.dtype, device=zc.device)
    assert log_q.is_contiguous() and log_q.dtype == zc.dtype
    _launch("vcnf_vdeq_decode", ptr(zc), zc.stride(0), ptr(quants), B, n, ptr(out), ptr(log_q), dtype_code(zc), stream())
    return out


# ------------------------------------------------------------------ base-distribution ends (f1)
class _Gauss(torch.autograd.Function):
    """Diagonal (class-conditional) Gaussian: mode 0 = log_prob(z), mode 1 = (z, log_p) from standard-normal draws.
    `loc`, `log_scale` are passed in the module's own layout; (js, cs, classes) describe it (see vcnf_gauss_fwd)."""

    @staticmethod
    def forward(ctx, zin, loc, log_scale, y, log_temp, js, cs, classes, mode):
        zin = zin.contiguous()
        B = zin.shape[0]
        n = zin.numel() // B if B else int(loc.numel() // classes)
        loc_, ls_ = loc.contiguous(), log_scale.contiguous()
        logp = torch.empty(B, dtype=zin.dtype, device=zin.device)
        zout = torch.empty_like(zin) if mode == 1 else None
        _launch("vcnf_gauss_fwd", ptr(zin), ptr(loc_), ptr(ls_), js, cs, ptr(y), float(log_temp), None, ptr(zout),
                ptr(logp), B, n, mode, dtype_code(zin), stream())
        ctx.cfg = (float(log_temp), js, cs, classes, mode, n, loc.shape)
        ctx.save_for_backward(zin, loc_, ls_, y)
        if mode == 1:
            return zout, logp
        return logp

    @staticmethod
    @once_differentiable
    def backward(ctx, *grads):
        zin, loc_, ls_, y = ctx.saved_tensors
        log_temp, js, cs, classes, mode, n, pshape = ctx.cfg
        B = zin.shape[0]
        if mode == 1:
            g_z, g_logp = grads
            g_z = g_z.contiguous() if g_z is not None else None
        else:
            (g_logp,), g_z = grads, None
        g_logp = g_logp.contiguous() if g_logp is not None else None
        want_z = mode == 0 and ctx.needs_input_grad[0]
        want_p = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
        gz = torch.empty_like(zin) if want_z else None
        gl = gs = ws = None
        if want_p:
            gl = torch.empty((classes, n), dtype=zin.dtype, device=zin.device)
            gs = torch.empty_like(gl)
            ws = torch.empty(_lib.query("vcnf_gauss_bwd_workspace", n, classes, dtype_code(zin)), dtype=torch.uint8,
                             device=zin.device)
        if want_z or want_p:
            _launch("vcnf_gauss_bwd", ptr(zin), ptr(loc_), ptr(ls_), js, cs, ptr(y), log_temp, ptr(g_logp), ptr(g_z),
                    ptr(gz), ptr(gl), ptr(gs), ptr(ws), B, n, classes, mode, dtype_code(zin), stream(),
                    n=3 if want_p else 1)
        if want_p:
            # dense [classes, n] -> the parameter's layout ([1, *shape] or [*shape, classes])
            if cs == 0:
                gl, gs = gl.reshape(pshape), gs.reshape(pshape)
            else:
                gl, gs = gl.t().reshape(pshape), gs.t().reshape(pshape)
        return gz, gl, gs, None, None, None, None, None, None


def gauss_logprob(z, loc, log_scale, y=None, log_temp=0.0, classes=1):
    z, loc, log_scale = _promote(z, loc, log_scale)
    js, cs = (1, 0) if classes == 1 else (classes, 1)
    return _Gauss.apply(z, loc, log_scale, y, log_temp, js, cs, classes, 0)


def gauss_sample(eps, loc, log_scale, y=None, log_temp=0.0, classes=1):
    eps, loc, log_scale = _promote(eps, loc, log_scale)
    js, cs = (1, 0) if classes == 1 else (classes, 1)
    return _Gauss.apply(eps, loc, log_scale, y, log_temp, js, cs, classes, 1)


class _GmmLogProb(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, mu, log_scale, weight_scores):
        z, mu_, ls_, ws_ = z.contiguous(), mu.contiguous(), log_scale.contiguous(), weight_scores.contiguous()
        B, d = z.shape
        K = ws_.numel()
        logp = torch.empty(B, dtype=z.dtype, device=z.device)
        _launch("vcnf_gmm_logprob", ptr(z), ptr(mu_), ptr(ls_), ptr(ws_), None, ptr(logp), B, K, d, dtype_code(z),
                stream())
        ctx.shapes = (mu.shape, log_scale.shape, weight_scores.shape, K, d)
        ctx.save_for_backward(z, mu_, ls_, ws_)
        return logp

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        z, mu_, ls_, ws_ = ctx.saved_tensors
        s_mu, s_ls, s_ws, K, d = ctx.shapes
        B = z.shape[0]
        g = g.contiguous()
        gz = torch.empty_like(z) if ctx.needs_input_grad[0] else None
        gp = torch.empty(2 * K * d + K, dtype=z.dtype, device=z.device)
        ws = torch.empty(_lib.query("vcnf_gmm_bwd_workspace", K, d, dtype_code(z)), dtype=torch.uint8, device=z.device)
        _launch("vcnf_gmm_logprob_bwd", ptr(z), ptr(mu_), ptr(ls_), ptr(ws_), ptr(g), ptr(gz), ptr(gp), ptr(ws), B, K, d,
                dtype_code(z), stream(), n=2)
        return gz, gp[:K * d].view(s_mu), gp[K * d:2 * K * d].view(s_ls), gp[2 * K * d:].view(s_ws)


def gmm_logprob(z, mu, log_scale, weight_scores):
    z, mu, log_scale, weight_scores = _promote(z, mu, log_scale, weight_scores)
    return _GmmLogProb.apply(z, mu, log_scale, weight_scores)


# ------------------------------------------------------------------ preprocessing transforms (f4), checkerboard (f2)
class _Logit(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, a0, beta, to_logit, want_ld):
        z = z.contiguous()
        B = z.shape[0]
        n = z.numel() // max(B, 1)
        out = torch.empty_like(z)
        ld = torch.empty(B, dtype=z.dtype, device=z.device) if want_ld else None
        _launch("vcnf_logit_fwd", ptr(z), ptr(out), ptr(ld), B, n, float(a0), float(beta), int(to_logit),
                dtype_code(z), stream())
        ctx.cfg = (float(a0), float(beta), int(to_logit), B, n)
        ctx.save_for_backward(z)
        if want_ld:
            return out, ld
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, gy, gld=None):
        (z,) = ctx.saved_tensors
        a0, beta, to_logit, B, n = ctx.cfg
        gy = gy.contiguous() if gy is not None else None
        gld = gld.contiguous() if gld is not None else None
        gz = torch.empty_like(z)
        _launch("vcnf_logit_bwd", ptr(z), ptr(gy), ptr(gld), ptr(gz), B, n, a0, beta, to_logit, dtype_code(z), stream())
        return gz, None, None, None, None


def logit_transform(z, a0, beta, to_logit, want_logdet=True):
    """to_logit: logit(a0 + beta z) (+ per-sample log-det); else (sigmoid(z) - a0) / beta."""
    return _Logit.apply(z, a0, beta, to_logit, want_logdet)


def axpb(x, a=1.0, b=0.0, noise=None, c=0.0, out=None):
    """out = a x + b + c noise in one launch (no autograd; out may be x for the in-place Shift)."""
    x = x if x.is_contiguous() else x.contiguous()
    out = torch.empty_like(x) if out is None else out
    assert out.is_contiguous() and out.shape == x.shape
    noise = noise.contiguous() if noise is not None else None
    _launch("vcnf_axpb", ptr(x), ptr(noise), ptr(out), x.numel(), float(a), float(b), float(c), dtype_code(x), stream())
    return out


_CB_PARITY = {}


def _mid_parity(mid_shape, device):
    key = (tuple(mid_shape), str(device))
    if key not in _CB_PARITY:
        par = torch.zeros(mid_shape if len(mid_shape) else (1,), dtype=torch.int64)
        for d, size in enumerate(mid_shape):
            shape = [1] * len(mid_shape)
            shape[d] = size
            par = par + torch.arange(size).view(shape)
        _CB_PARITY[key] = (par % 2).to(torch.uint8).reshape(-1).contiguous().to(device)
    return _CB_PARITY[key]


class _Checkerboard(torch.autograd.Function):
    """merge = False: z -> (z1, z2); merge = True: (z1, z2) -> z.  The adjoint is the opposite index map."""

    @staticmethod
    def forward(ctx, inv, merge, *ts):
        if not merge:
            z = ts[0].contiguous()
            W = z.shape[-1]
            half = z.shape[:-1] + (W // 2,)
            z1, z2 = torch.empty(half, dtype=z.dtype, device=z.device), torch.empty(half, dtype=z.dtype, device=z.device)
        else:
            z1, z2 = ts[0].contiguous(), ts[1].contiguous()
            W = 2 * z1.shape[-1]
            z = torch.empty(z1.shape[:-1] + (W,), dtype=z1.dtype, device=z1.device)
        B = z.shape[0]
        mid = tuple(z.shape[1:-1])
        M = math.prod(mid) if mid else 1
        par = _mid_parity(mid, z.device)
        _launch("vcnf_checkerboard", ptr(z), ptr(z1), ptr(z2), ptr(par), B, M, W, int(inv), int(merge), dtype_code(z),
                stream())
        ctx.cfg = (inv, merge)
        return z if merge else (z1, z2)

    @staticmethod
    @once_differentiable
    def backward(ctx, *gs):
        inv, merge = ctx.cfg
        if merge:
            g1, g2 = _Checkerboard.apply(inv, False, gs[0])
            return None, None, g1, g2
        g1, g2 = gs
        ref = g1 if g1 is not None else g2
        g1 = g1 if g1 is not None else torch.zeros_like(ref)
        g2 = g2 if g2 is not None else torch.zeros_like(ref)
        return None, None, _Checkerboard.apply(inv, True, g1, g2)


def checkerboard_split(z, inv):
    if z.shape[-1] % 2:
        raise ValueError("checkerboard split needs an even last dimension")
    return _Checkerboard.apply(bool(inv), False, z)


def checkerboard_merge(z1, z2, inv):
    z1, z2 = _promote(z1, z2)
    return _Checkerboard.apply(bool(inv), True, z1, z2)


# ------------------------------------------------------------------ autoregressive sampling (K7)
def made_inverse(y, packed, map_code, K=0, modes=None, bounds=None, outside_zero=False):
    """All D passes of an autoregressive layer's sampling direction in one launch (vcnf_made_inverse); `packed` is
    the dict built by nets.made.MADE.pack_for_sampling().  Returns (x, logabsdet).  No gradient."""
    y = y.contiguous()
    B, D = y.shape
    x = torch.empty_like(y)
    ld = torch.empty(B, dtype=y.dtype, device=y.device)
    _launch("vcnf_made_inverse", ptr(y), ptr(x), ptr(ld), B, D, packed["H"], packed["P"], packed["n_layers"],
            ptr(packed["band"]), ptr(packed["fin"]), ptr(packed["band_idx"]), ptr(packed["order"]),
            packed["band_units"], ptr(bounds), ptr(modes), int(K), int(map_code), int(outside_zero), None, stream())
    return x, ld


# ------------------------------------------------------------------ fused conditioner + spline
FUSED_NET = True      # False: spline coupling layers keep the layer-by-layer kernels (the tests' reference path)


class FusedNetSpec:
    """Host-side constants of vcnf_resnet_spline_fwd for one coupling layer: column lists, tail modes / bounds (ctypes
    arrays: the kernel takes them as launch arguments) and the packed-parameter buffer."""

    def __init__(self, id_cols, tr_cols, modes, bounds, K, Pc, wh_scale, outside_zero, hidden, num_blocks,
                 shared_modes=(), shared_bounds=(), Ps=0):
        import ctypes
        self.Ts, self.Ps = len(shared_modes), int(Ps)
        self.smodes = (ctypes.c_uint8 * max(self.Ts, 1))(*[int(m) for m in shared_modes])
        self.sbounds = (ctypes.c_float * max(self.Ts, 1))(*[float(b) for b in shared_bounds])
        self.n_in, self.Tc, self.K, self.Pc = len(id_cols), len(tr_cols), int(K), int(Pc)
        self.hidden, self.num_blocks = int(hidden), int(num_blocks)
        self.id_cols = (ctypes.c_int32 * self.n_in)(*[int(c) for c in id_cols])
        self.tr_cols = (ctypes.c_int32 * self.Tc)(*[int(c) for c in tr_cols])
        self.modes = (ctypes.c_uint8 * self.Tc)(*[int(m) for m in modes])
        self.bounds = (ctypes.c_float * self.Tc)(*[float(b) for b in bounds])
        self.wh_scale, self.outside_zero = float(wh_scale), int(outside_zero)
        self.nbytes = _lib.query("vcnf_resnet_spline_pack_bytes", self.n_in, self.hidden, self.num_blocks, self.Tc, self.Pc)
        self.packed = {}

    @property
    def supported(self):
        return self.nbytes > 0 and self.K == 8 and self.Ts in (0, self.n_in)

    def pack(self, weights, biases):
        import ctypes
        dev = weights[0].device
        buf = self.packed.get(str(dev))
        if buf is None:
            buf = self.packed[str(dev)] = torch.empty(self.nbytes, dtype=torch.uint8, device=dev)
        n = len(weights)
        W = (ctypes.c_void_p * n)(*[w.data_ptr() for w in weights])
        b = (ctypes.c_void_p * n)(*[(v.data_ptr() if v is not None else None) for v in biases])
        _launch("vcnf_resnet_spline_pack", W, b, self.n_in, self.hidden, self.num_blocks, self.Tc, self.Pc, ptr(buf),
                stream())
        return buf


def resnet_spline(x, fspec, weights, biases, inverse, shared=None, want_params=False, want_acts=False, want_xid=False):
    """The whole ResidualNet conditioner + the coupling spline in one launch (vcnf_resnet_spline_fwd), no gradient.
    weights / biases: contiguous fp32 parameters in forward order; shared: [Ts, Ps] parameters of the unconditional
    spline on the identity columns (or None).  Returns (y, logabsdet[, params][, acts])."""
    x = x.contiguous()
    if (shared is None) != (fspec.Ts == 0):
        raise ValueError("resnet_spline: shared parameters do not match the layer's specification")
    shared = shared.detach().contiguous() if shared is not None else None
    B, D = x.shape
    packed = fspec.pack(weights, biases)
    y = torch.empty_like(x)
    ld = torch.empty(B, dtype=x.dtype, device=x.device)
    params = None
    if want_params:       # "padded": row pitch a multiple of 16 bytes, like ResidualNet's own output for the spline layer
        params = (_padded_empty(B, fspec.Tc * fspec.Pc, x.dtype, x.device) if want_params == "padded"
                  else torch.empty((B, fspec.Tc * fspec.Pc), dtype=x.dtype, device=x.device))
    acts = (torch.empty((1 + 2 * fspec.num_blocks, B, fspec.hidden), dtype=x.dtype, device=x.device)
            if want_acts else None)
    xid = None
    if want_xid:          # x[:, id_cols]: what gather_cols would return ("padded": 16-byte-multiple row pitch)
        xid = (_padded_empty(B, fspec.n_in, x.dtype, x.device) if want_xid == "padded"
               else torch.empty((B, fspec.n_in), dtype=x.dtype, device=x.device))
    _launch("vcnf_resnet_spline_fwd", ptr(x), D, ptr(y), D, ptr(ld), 0, B, D, fspec.K, fspec.id_cols, fspec.n_in,
            fspec.tr_cols, fspec.Tc, fspec.Pc, fspec.hidden, fspec.num_blocks, ptr(packed), fspec.modes, fspec.bounds,
            fspec.wh_scale, int(inverse), fspec.outside_zero, None, ptr(shared), fspec.Ts, fspec.Ps, fspec.smodes,
            fspec.sbounds, ptr(params),
            params.stride(0) if params is not None else 0, ptr(acts), ptr(xid), xid.stride(0) if xid is not None else 0,
            stream())
    out = (y, ld)
    if want_params:
        out += (params,)
    if want_acts:
        out += (acts,)
    if want_xid:
        out += (xid,)
    return out
