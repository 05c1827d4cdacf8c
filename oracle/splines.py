"""Oracle: monotone rational-quadratic spline with linear / circular tails.

Restates `normflow/utils/splines.py:12-193` in predicate form: the reference compacts
the in-range elements with boolean indexing (`splines.py:73-77`) and runs the dense
routine on them; here every element is evaluated on a clamped copy of the input and the
out-of-range ones are overwritten afterwards, which yields the same values and the same
(zero) parameter gradients for out-of-range elements.
The order of floating-point operations inside the bin follows the reference line by
line so fp32 results agree to rounding.
"""
import math

import torch
import torch.nn.functional as F

MIN_BIN_WIDTH = 1e-3       # splines.py:6
MIN_BIN_HEIGHT = 1e-3      # splines.py:7
MIN_DERIVATIVE = 1e-3      # splines.py:8


def boundary_derivative_constant(min_derivative=MIN_DERIVATIVE):
    """`np.log(np.exp(1 - min_derivative) - 1)`, splines.py:38."""
    return math.log(math.exp(1.0 - min_derivative) - 1.0)


def bin_search(knots, x, eps=1e-6):
    """splines.py:12-17: index = #{k : x >= knot_k} - 1 with the last knot raised by eps.

    The reference adds eps in place; here a copy is raised (the caller's knots are the
    un-raised ones in both, because the reference gathers *after* the mutation only from
    `cumwidths`/`cumheights` entries < K... see note in `rqs`)."""
    raised = knots.clone()
    raised[..., -1] = raised[..., -1] + eps
    return (x[..., None] >= raised).sum(-1) - 1


def _knots(unnorm, lo, hi, min_size):
    """splines.py:109-119 (widths) == :123-133 (heights).  Returns (knots[K+1], sizes[K])."""
    k = unnorm.shape[-1]
    p = F.softmax(unnorm, dim=-1)
    p = min_size + (1 - min_size * k) * p
    cum = torch.cumsum(p, dim=-1)
    cum = F.pad(cum, pad=(1, 0), mode="constant", value=0.0)
    if torch.is_tensor(lo):
        cum = (hi[..., None] - lo[..., None]) * cum + lo[..., None]
        first = lo.unsqueeze(-1)
        last = hi.unsqueeze(-1)
    else:
        cum = (hi - lo) * cum + lo
        first = torch.full_like(cum[..., :1], lo)
        last = torch.full_like(cum[..., :1], hi)
    cum = torch.cat([first.expand_as(cum[..., :1]), cum[..., 1:-1], last.expand_as(cum[..., :1])], dim=-1)
    return cum, cum[..., 1:] - cum[..., :-1]


def rqs(x, uw, uh, ud, inverse=False, left=0.0, right=1.0, bottom=0.0, top=1.0,
        min_bin_width=MIN_BIN_WIDTH, min_bin_height=MIN_BIN_HEIGHT,
        min_derivative=MIN_DERIVATIVE, return_bin=False):
    """Bounded spline, `rational_quadratic_spline` splines.py:88-193.

    x: [...]; uw, uh: [..., K]; ud: [..., K+1].  Returns (y, logabsdet[, bin_idx]).
    Note (splines.py:13): the reference's search raises the last knot *in place*, so the
    raised value is also what `cumwidths.gather` would return for index K -- which is
    never gathered (bin <= K-1), hence no visible effect beyond the search itself.
    """
    k = uw.shape[-1]
    if min_bin_width * k > 1.0:
        raise ValueError("Minimal bin width too large for the number of bins")
    if min_bin_height * k > 1.0:
        raise ValueError("Minimal bin height too large for the number of bins")
    cw, w = _knots(uw, left, right, min_bin_width)
    d = min_derivative + F.softplus(ud)
    ch, h = _knots(uh, bottom, top, min_bin_height)

    idx = bin_search(ch if inverse else cw, x)[..., None]
    # The reference would raise an index error for idx outside [0, K-1]; only reachable
    # when x is outside [left, right], which callers exclude.
    idx_c = idx.clamp(0, k - 1)
    g = lambda t: t.gather(-1, idx_c)[..., 0]
    x_k, w_k, y_k, h_k = g(cw), g(w), g(ch), g(h)
    delta = h / w
    s_k = g(delta)
    d_k, d_k1 = g(d), d[..., 1:].gather(-1, idx_c)[..., 0]

    if inverse:                                               # splines.py:152-177
        a = (x - y_k) * (d_k + d_k1 - 2 * s_k) + h_k * (s_k - d_k)
        b = h_k * d_k - (x - y_k) * (d_k + d_k1 - 2 * s_k)
        c = -s_k * (x - y_k)
        disc = b.pow(2) - 4 * a * c
        root = (2 * c) / (-b - torch.sqrt(disc))
        y = root * w_k + x_k
        t1mt = root * (1 - root)
        den = s_k + (d_k + d_k1 - 2 * s_k) * t1mt
        num = s_k.pow(2) * (d_k1 * root.pow(2) + 2 * s_k * t1mt + d_k * (1 - root).pow(2))
        lad = -(torch.log(num) - 2 * torch.log(den))
    else:                                                     # splines.py:178-193
        theta = (x - x_k) / w_k
        t1mt = theta * (1 - theta)
        numer = h_k * (s_k * theta.pow(2) + d_k * t1mt)
        den = s_k + (d_k + d_k1 - 2 * s_k) * t1mt
        y = y_k + numer / den
        num = s_k.pow(2) * (d_k1 * theta.pow(2) + 2 * s_k * t1mt + d_k * (1 - theta).pow(2))
        lad = torch.log(num) - 2 * torch.log(den)
    if return_bin:
        return y, lad, idx[..., 0]
    return y, lad


def unconstrained_rqs(x, uw, uh, ud, inverse=False, tails="linear", tail_bound=1.0,
                      min_bin_width=MIN_BIN_WIDTH, min_bin_height=MIN_BIN_HEIGHT,
                      min_derivative=MIN_DERIVATIVE, return_bin=False):
    """`unconstrained_rational_quadratic_spline`, splines.py:20-85.

    tails: 'linear' (ud has K-1 entries), 'circular' (K entries) or a per-feature list of
    those two (K+1 entries; the feature axis is the last axis of x).  tail_bound: python
    scalar or tensor broadcastable to x.shape.
    """
    if tails is None:
        # neural_spline/coupling.py:321-323, autoregressive.py:106-108: the bounded spline on [0, 1] is called
        # directly (tail_bound is not passed); splines.py:105-106 refuses inputs outside the box
        if x.numel() and (torch.min(x) < 0.0 or torch.max(x) > 1.0):
            raise ValueError("Input outside domain")
        return rqs(x, uw, uh, ud, inverse=inverse, min_bin_width=min_bin_width, min_bin_height=min_bin_height,
                   min_derivative=min_derivative, return_bin=return_bin)
    const = boundary_derivative_constant(min_derivative)
    if torch.is_tensor(tail_bound):
        bound = torch.broadcast_to(tail_bound.to(x.dtype), x.shape)
        inside = (x >= -bound) & (x <= bound)
        lo, hi = -bound, bound
        x_in = torch.where(inside, x, torch.zeros_like(x))
    else:
        inside = (x >= -tail_bound) & (x <= tail_bound)
        lo, hi = -tail_bound, tail_bound
        x_in = torch.where(inside, x, torch.zeros_like(x))

    if tails == "linear":                                     # splines.py:36-43
        edge = torch.full_like(ud[..., :1], const)
        ud_full = torch.cat([edge, ud, edge], dim=-1)
        outside_value = x
    elif tails == "circular":                                 # splines.py:44-49
        ud_full = torch.cat([ud, ud[..., :1]], dim=-1)
        outside_value = x
    elif isinstance(tails, (list, tuple)):                    # splines.py:50-57
        lin = torch.tensor([t == "linear" for t in tails], dtype=torch.bool)
        circ = torch.tensor([t == "circular" for t in tails], dtype=torch.bool)
        shape = [1] * (ud.dim() - 2) + [len(tails)]
        lin, circ = lin.view(shape), circ.view(shape)
        first = torch.where(lin, torch.full_like(ud[..., 0], const), ud[..., 0])
        last = torch.where(lin, torch.full_like(ud[..., -1], const), ud[..., -1])
        last = torch.where(circ, first, last)
        ud_full = torch.cat([first[..., None], ud[..., 1:-1], last[..., None]], dim=-1)
        # Hazard A.5: the list branch never copies the out-of-range inputs through.
        outside_value = torch.zeros_like(x)
    else:
        raise RuntimeError("{} tails are not implemented.".format(tails))

    res = rqs(x_in, uw, uh, ud_full, inverse=inverse, left=lo, right=hi, bottom=lo, top=hi,
              min_bin_width=min_bin_width, min_bin_height=min_bin_height,
              min_derivative=min_derivative, return_bin=return_bin)
    y = torch.where(inside, res[0], outside_value)
    lad = torch.where(inside, res[1], torch.zeros_like(x))
    if return_bin:
        return y, lad, torch.where(inside, res[2], torch.full_like(res[2], -1))
    return y, lad
