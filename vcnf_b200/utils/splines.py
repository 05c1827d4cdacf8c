"""Functional spline entry points with the reference's names (normflow/utils/splines.py),
evaluated by the `vcnf_rqs_layer_*` kernels (each scalar input is one "row" with D = 1)."""
import torch

from .. import ops

DEFAULT_MIN_BIN_WIDTH = 1e-3
DEFAULT_MIN_BIN_HEIGHT = 1e-3
DEFAULT_MIN_DERIVATIVE = 1e-3


def check_minimums(num_bins, min_bin_width, min_bin_height, min_derivative):
    """The kernels are compiled for the reference's default floors (splines.py:6-8)."""
    if (abs(min_bin_width - DEFAULT_MIN_BIN_WIDTH) > 1e-12 or abs(min_bin_height - DEFAULT_MIN_BIN_HEIGHT) > 1e-12
            or abs(min_derivative - DEFAULT_MIN_DERIVATIVE) > 1e-12):
        raise NotImplementedError("non-default min_bin_width / min_bin_height / min_derivative")
    if min_bin_width * num_bins > 1.0:
        raise ValueError("Minimal bin width too large for the number of bins")
    if min_bin_height * num_bins > 1.0:
        raise ValueError("Minimal bin height too large for the number of bins")


def unconstrained_rational_quadratic_spline(inputs, unnormalized_widths, unnormalized_heights,
                                            unnormalized_derivatives, inverse=False, tails="linear",
                                            tail_bound=1., min_bin_width=DEFAULT_MIN_BIN_WIDTH,
                                            min_bin_height=DEFAULT_MIN_BIN_HEIGHT,
                                            min_derivative=DEFAULT_MIN_DERIVATIVE):
    """splines.py:20-85 for string tails and a scalar tail bound; returns (outputs, logabsdet)
    with the shape of `inputs`."""
    num_bins = unnormalized_widths.shape[-1]
    check_minimums(num_bins, min_bin_width, min_bin_height, min_derivative)
    if tails == "linear":
        mode = ops.TAIL_LINEAR
    elif tails == "circular":
        mode = ops.TAIL_CIRCULAR
    else:
        raise RuntimeError("{} tails are not implemented.".format(tails))
    if torch.is_tensor(tail_bound):
        raise NotImplementedError("tensor tail_bound: use the layer classes")
    shape = inputs.shape
    P = 2 * num_bins + unnormalized_derivatives.shape[-1]
    params = torch.cat([unnormalized_widths, unnormalized_heights, unnormalized_derivatives], dim=-1).reshape(-1, P)
    spec = ops.SplineSpec(1, num_bins, [(ops.COL_COND << 16)], [mode], [float(tail_bound)], P, 1.0, [], [], P, False)
    y, lad = ops.rqs_layer(inputs.reshape(-1, 1), params, None, spec, inverse)
    return y.reshape(shape), lad.reshape(shape)
