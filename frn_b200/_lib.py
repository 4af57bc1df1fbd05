"""ctypes binding of libfrn_b200.so (the C ABI declared in include/frn_b200.h).

The library is built in-tree by ``__graft_entry__.build()`` / ``make -C frn_b200/csrc``.
There is NO fallback: if the shared object is missing, importing succeeds (so that
CPU-only tooling can still import the package) but the first call raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# FRN_B200_LIB: instrumented builds of the same library (tools/prof_*.py); the default is the in-tree product build
LIB_PATH = os.environ.get("FRN_B200_LIB") or os.path.join(_HERE, "libfrn_b200.so")

c_float_p = C.c_void_p  # raw device pointers travel as integers


class CabParams(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in
                ("g_w", "g_b", "theta_w", "theta_b", "phi_w", "phi_b", "W_w", "W_b")]


class RefusionParams(C.Structure):
    _fields_ = [("conv0_rgb_w", C.c_void_p), ("conv0_rgb_b", C.c_void_p),
                ("conv0_evt_w", C.c_void_p), ("conv0_evt_b", C.c_void_p),
                ("rgb_ca", CabParams), ("event_ca", CabParams)]


class GemmDesc(C.Structure):
    _fields_ = ([("A", C.c_void_p), ("B", C.c_void_p), ("C", C.c_void_p)] +
                [(n, C.c_int) for n in ("M", "N", "K", "batch")] +
                [(n, C.c_long) for n in ("sAm", "sAk", "sAb", "sBk", "sBn", "sBb", "sCm", "sCn", "sCb")] +
                [("bias_m", C.c_void_p), ("bias_n", C.c_void_p), ("accumulate", C.c_int), ("reduce_batch", C.c_int)])


# name -> (restype, argtypes); mirrors include/frn_b200.h one to one
_SIGNATURES = {
    "frn_abi_version": (C.c_int, []),
    "frn_error_string": (C.c_char_p, [C.c_int]),
    "frn_kernel_inventory": (C.c_char_p, []),
    "frn_launch_count": (C.c_ulonglong, []),
    "frn_mean_std": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_float,
                               C.c_void_p, C.c_void_p]),
    "frn_adain_forward": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_void_p]),
    "frn_adain_backward": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_size_t]),
    "frn_adain_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "frn_adain_static_max_rows": (C.c_long, []),
    "frn_adain_backward_static": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "frn_cab_saved_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "frn_cab_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "frn_cab_forward": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                  C.POINTER(CabParams), C.c_void_p, C.c_long, C.c_void_p, C.c_size_t,
                                  C.c_void_p, C.c_size_t]),
    "frn_cab_backward": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                   C.POINTER(CabParams), C.c_void_p, C.c_void_p, C.c_long, C.c_void_p,
                                   C.c_void_p, C.c_int, C.POINTER(CabParams), C.c_void_p, C.c_size_t]),
    "frn_refusion_saved_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "frn_refusion_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "frn_refusion_forward": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                       C.POINTER(RefusionParams), C.c_void_p, C.c_void_p, C.c_size_t,
                                       C.c_void_p, C.c_size_t]),
    "frn_refusion_backward": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                        C.POINTER(RefusionParams), C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.POINTER(RefusionParams), C.c_void_p, C.c_size_t]),
    "frn_gemm_workspace_bytes": (C.c_size_t, [C.POINTER(GemmDesc)]),
    "frn_gemm_f32": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(GemmDesc), C.c_void_p, C.c_size_t]),
    "frn_attn_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "frn_attn_forward": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 6 + [C.c_size_t]),
    "frn_attn_backward": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 10 + [C.c_size_t]),
    "frn_refusion_infer_workspace_bytes": (C.c_size_t, [C.c_int] * 4),
    "frn_refusion_infer": (C.c_int, [C.c_void_p] + [C.c_int] * 4 + [C.c_void_p, C.c_void_p,
                                     C.POINTER(RefusionParams), C.c_void_p, C.c_void_p, C.c_size_t]),
    "frn_refusion_lateral_workspace_bytes": (C.c_size_t, [C.c_int] * 4),
    "frn_refusion_lateral_forward": (C.c_int, [C.c_void_p] + [C.c_int] * 4 + [C.c_void_p, C.c_void_p,
                                               C.POINTER(RefusionParams)] + [C.c_void_p] * 5 + [C.c_size_t, C.c_void_p,
                                               C.c_size_t]),
    "frn_refusion_lateral_backward": (C.c_int, [C.c_void_p] + [C.c_int] * 4 + [C.c_void_p, C.c_void_p,
                                                C.POINTER(RefusionParams)] + [C.c_void_p] * 6 +
                                      [C.POINTER(RefusionParams)] + [C.c_void_p] * 3 + [C.c_size_t]),
    "frn_conv1x1_workspace_bytes": (C.c_size_t, [C.c_int] * 4),
    "frn_conv1x1_forward": (C.c_int, [C.c_void_p] + [C.c_int] * 4 + [C.c_void_p] * 5 + [C.c_size_t]),
    "frn_conv1x1_backward": (C.c_int, [C.c_void_p] + [C.c_int] * 4 + [C.c_void_p] * 7 + [C.c_size_t]),
    "frn_anchors_count": (C.c_long, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int]),
    "frn_anchors_generate": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                       C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_double), C.c_int,
                                       C.POINTER(C.c_double), C.c_void_p]),
    "frn_focal_loss_saved_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "frn_focal_loss_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "frn_focal_loss_forward": (C.c_int, [C.c_void_p] + [C.c_int] * 4 + [C.c_void_p] * 4 + [C.c_float, C.c_void_p,
                                         C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "frn_focal_loss_backward": (C.c_int, [C.c_void_p] + [C.c_int] * 4 + [C.c_void_p] * 4 + [C.c_float] +
                                [C.c_void_p] * 4),
    "frn_voxelize_workspace_bytes": (C.c_size_t, []),
    "frn_voxelize_events": (C.c_int, [C.c_void_p, C.c_long, C.c_int] + [C.c_void_p] * 4 + [C.c_int] * 4 +
                            [C.c_void_p, C.c_void_p, C.c_size_t]),
    "frn_voxel_normalize": (C.c_int, [C.c_void_p, C.c_long, C.c_int, C.c_void_p]),
    "frn_detect_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "frn_decode_boxes": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "frn_detect_postprocess": (C.c_int, [C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 4 + [C.c_int, C.c_int,
                                         C.c_float, C.c_float, C.c_int] + [C.c_void_p] * 6 + [C.c_size_t]),
    "frn_set_gemm_engine": (None, [C.c_int]),
    "frn_get_gemm_engine": (C.c_int, []),
    "frn_set_attn_engine": (None, [C.c_int]),
    "frn_get_attn_engine": (C.c_int, []),
    "frn_timing_enable": (None, [C.c_int]),
    "frn_timing_read": (C.c_int, [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_long)]),
    "frn_timing_kernels": (C.c_long, [C.c_char_p, C.c_long]),
}

FRN_F32, FRN_BF16 = 0, 1
EXPORTS = tuple(_SIGNATURES)
_lib = None


class FrnError(RuntimeError):
    pass


def load():
    """Load the shared object (once) and set the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FrnError(
            f"frn_b200: {LIB_PATH} is missing -- build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` or `make -C frn_b200/csrc`. There is no PyTorch/CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here == header/library drift
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(code: int, what: str):
    if code != 0:
        msg = load().frn_error_string(code).decode()
        raise FrnError(f"frn_b200.{what} failed: {msg} (code {code})")
