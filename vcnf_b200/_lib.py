"""ctypes binding of libvcnf_b200.so (the C ABI declared in include/vcnf_b200.h).

There is deliberately NO fallback: if the library is missing or a tensor is not on a CUDA
device the call raises.  `python -m vcnf_b200.build` (or `__graft_entry__.build()`) compiles it.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvcnf_b200.so")

_i64, _i32, _dbl, _vp = ctypes.c_int64, ctypes.c_int, ctypes.c_double, ctypes.c_void_p

# name -> argument ctypes (pointers are passed as integers / None through c_void_p)
_SIGNATURES = {
    "vcnf_version": [],
    "vcnf_sm_count": [],
    "vcnf_rqs_layer_fwd": [_vp, _i64, _vp, _i64, _vp, _i32, _i64, _i32, _i32, _vp,
                           _vp, _i64, _i32, _i32, _vp, _vp, _dbl,
                           _vp, _i32, _i32, _vp, _vp, _i32, _i32, _vp, _vp, _i32, _vp],
    "vcnf_rqs_layer_bwd": [_vp, _i64, _vp, _i64, _vp, _vp, _i64, _i64, _i32, _i32, _vp,
                           _vp, _i64, _i32, _i32, _vp, _vp, _dbl, _vp, _i64,
                           _vp, _i32, _i32, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _vp],
    "vcnf_gather_cols": [_vp, _i64, _vp, _i64, _i64, _i32, _vp, _i32, _vp],
    "vcnf_scatter_cols": [_vp, _i64, _vp, _i64, _i64, _i32, _i32, _vp, _i32, _i32, _vp],
    "vcnf_affine_coupling_fwd": [_vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp],
    "vcnf_affine_coupling_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _i32, _i32,
                                 _i32, _vp],
    "vcnf_masked_affine_fwd": [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _vp],
    "vcnf_masked_affine_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _vp],
    "vcnf_const_affine_fwd": [_vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp],
    "vcnf_const_affine_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp],
    "vcnf_const_affine_bwd_static": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp],
    "vcnf_permute_channels": [_vp, _vp, _i64, _i32, _i32, _vp, _i32, _vp],
    "vcnf_squeeze": [_vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _vp],
    "vcnf_linear_tf32x3": [_vp, _i64, _vp, _i64, _i64, _vp, _vp, _i64, _vp, _i64, _vp, _i64, _i64, _i32, _i32, _i32, _vp],
    "vcnf_linear_wgrad_tf32x3": [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _vp],
    "vcnf_small_linear": [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp],
    "vcnf_small_linear_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _vp],
    "vcnf_lu_compose_fwd": [_vp, _vp, _vp, _dbl, _i32, _vp, _vp, _i32, _vp],
    "vcnf_mlp_fwd": [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _i32, _dbl, _i64, _i32, _vp],
    "vcnf_mlp_bwd": [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _i32, _dbl, _vp, _i64, _vp, _vp, _i64, _i32, _vp],
    "vcnf_lu_inverse": [_vp, _vp, _vp, _dbl, _i32, _vp, _vp, _i32, _vp],
    "vcnf_lu_compose_bwd": [_vp, _vp, _vp, _dbl, _i32, _vp, _vp, _vp, _vp, _vp, _i32, _vp],
    "vcnf_lu_compose_fwd_batched": [_vp, _vp, _vp, _dbl, _i32, _i32, _vp, _vp, _i32, _vp],
    "vcnf_lu_inverse_batched": [_vp, _vp, _vp, _dbl, _i32, _i32, _vp, _vp, _i32, _vp],
    "vcnf_lu_compose_bwd_batched": [_vp, _vp, _vp, _dbl, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _i32, _vp],
    "vcnf_im2col": [_vp, _i64, _i64, _i64, _i64, _vp, _i64, _i64, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _dbl, _i32, _vp],
    "vcnf_col2im": [_vp, _i64, _vp, _vp, _i64, _i64, _i64, _i64, _i64, _i32, _i32, _i32, _i32, _i32, _i32, _dbl, _vp],
    "vcnf_colsum": [_vp, _i32, _i64, _i32, _i64, _vp, _vp, _vp],
    "vcnf_conv_gemm_bf16": [_vp, _i64, _i32, _i32, _i32, _i32, _vp, _i32, _vp, _vp, _i64, _dbl, _i32, _vp, _i32, _i32,
                            _i64, _vp],
    "vcnf_conv_wgrad_bf16": [_vp, _i64, _i32, _vp, _i64, _i32, _i64, _vp, _vp, _vp],
    "vcnf_vdeq_fwd": [_vp, _i64, _vp, _vp, _vp, _dbl, _i64, _i32, _i32, _vp, _vp, _vp, _i32, _vp],
    "vcnf_vdeq_bwd": [_vp, _i64, _vp, _vp, _vp, _dbl, _i64, _i32, _i32, _vp, _vp, _vp, _vp, _i32, _vp],
    "vcnf_vdeq_decode": [_vp, _i64, _vp, _i64, _i32, _vp, _vp, _i32, _vp],
    "vcnf_gauss_fwd": [_vp, _vp, _vp, _i64, _i64, _vp, _dbl, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _vp],
    "vcnf_gauss_bwd": [_vp, _vp, _vp, _i64, _i64, _vp, _dbl, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32,
                       _vp],
    "vcnf_gmm_logprob": [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _vp],
    "vcnf_gmm_logprob_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _vp],
    "vcnf_logit_fwd": [_vp, _vp, _vp, _i64, _i32, _dbl, _dbl, _i32, _i32, _vp],
    "vcnf_logit_bwd": [_vp, _vp, _vp, _vp, _i64, _i32, _dbl, _dbl, _i32, _i32, _vp],
    "vcnf_axpb": [_vp, _vp, _vp, _i64, _dbl, _dbl, _dbl, _i32, _vp],
    "vcnf_checkerboard": [_vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _vp],
    "vcnf_adam_flat_guarded": [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _dbl, _dbl, _dbl, _dbl, _dbl, _vp, _i32, _i32,
                               _vp],
    "vcnf_adamax_flat_guarded": [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _dbl, _dbl, _dbl, _dbl, _dbl, _vp, _i32, _i32,
                               _vp],
    "vcnf_small_linear_wgrad_img": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _vp],
    "vcnf_convnet3_pack": [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp],
    "vcnf_glow_w_fwd": [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _dbl, _vp, _vp, _vp, _i32, _vp],
    "vcnf_glow_w_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _dbl, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32,
                        _vp],
    "vcnf_made_inverse": [_vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _i32,
                          _i32, _vp, _vp],
    "vcnf_adam_flat": [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _dbl, _dbl, _dbl, _dbl, _dbl, _i32, _vp],
    "vcnf_resnet_spline_pack": [_vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp, _vp],
    "vcnf_resnet_spline_fwd": [_vp, _i64, _vp, _i64, _vp, _i32, _i64, _i32, _i32, _vp, _i32, _vp, _i32, _i32, _i32, _i32,
                               _vp, _vp, _vp, _dbl, _i32, _i32, _vp, _vp, _i32, _i32, _vp, _vp, _vp, _i64, _vp, _vp, _i64,
                               _vp],
}
_I64_RETURN = {
    "vcnf_rqs_layer_bwd_workspace": [_i32, _i32, _i32],
    "vcnf_resnet_spline_pack_bytes": [_i32, _i32, _i32, _i32, _i32],
    "vcnf_const_affine_bwd_workspace": [_i32, _i32],
    "vcnf_small_linear_wgrad_img_workspace": [_i32, _i32],
    "vcnf_convnet3_pack_elems": [_i32, _i32, _i32, _i32],
    "vcnf_linear_wgrad_workspace": [_i64, _i32, _i32],
    "vcnf_small_linear_bwd_workspace": [_i32, _i32],
    "vcnf_mlp_bwd_workspace": [_vp, _i32, _i32],
    "vcnf_colsum_workspace": [_i32],
    "vcnf_gauss_bwd_workspace": [_i32, _i32, _i32],
    "vcnf_gmm_bwd_workspace": [_i32, _i32, _i32],
    "vcnf_vdeq_bwd_workspace": [_i32, _i32],
    "vcnf_conv_wgrad_workspace": [_i64, _i32, _i32],
}
EXPORTED = sorted(list(_SIGNATURES) + list(_I64_RETURN))

_lib = None


def lib():
    """The loaded library (loads on first use)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "vcnf_b200: %s is missing -- build it with `python -m vcnf_b200.build`; "
                "this package has no CPU or PyTorch fallback" % LIB_PATH)
        handle = ctypes.CDLL(LIB_PATH)
        for name, args in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.argtypes, fn.restype = args, ctypes.c_int
        for name, args in _I64_RETURN.items():
            fn = getattr(handle, name)
            fn.argtypes, fn.restype = args, ctypes.c_int64
        _lib = handle
    return _lib


def dtype_code(t):
    if t.dtype == torch.float32:
        return 0
    if t.dtype == torch.float64:
        return 1
    raise TypeError("vcnf_b200 kernels support float32 and float64, got %s" % t.dtype)


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("vcnf_b200: expected a CUDA tensor (there is no CPU path), got device %s" % t.device)
    return t.data_ptr()


def stream():
    return torch.cuda.current_stream().cuda_stream


def call(name, *args):
    rc = getattr(lib(), name)(*args)
    if rc != 0:
        raise RuntimeError("vcnf_b200: %s failed with code %d%s" % (
            name, rc, " (invalid argument)" if rc == -1 else " (unsupported size)" if rc == -2 else " (CUDA error)"))


def query(name, *args):
    return int(getattr(lib(), name)(*args))
