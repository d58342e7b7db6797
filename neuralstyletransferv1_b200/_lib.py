"""ctypes binding of ``libb200nst.so`` (C ABI declared in ``include/b200nst.h``).

There is deliberately no fallback: if the shared library has not been built
(``python -c "import __graft_entry__ as g; g.build()"``) importing the compute
path raises, and a forward on a machine without an sm_100 GPU raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NST_LIB_PATH") or os.path.join(_HERE, "libb200nst.so")   # override: A/B builds (tools/ab.sh)

NST_OK = 0
ARCH = {"johnson": 0, "nst": 1, "reconet": 2}
PRECISION = {"f16": 0, "f16x3": 1, "bf16": 2, "fp32_direct": 3}
COMPOSITE_MODE = {"keep": 0, "replace": 1}
MASK_TYPE = {"float32": 0, "float16": 1, "uint8": 2}

# symbol -> (restype, argtypes); must list every function include/b200nst.h declares
_c_float3 = C.POINTER(C.c_float)
PROTOTYPES = {
    "nst_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "nst_destroy": (None, [C.c_void_p]),
    "nst_load_weights": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_void_p),
                                   C.POINTER(C.c_int64), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "nst_output_shape": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "nst_workspace_bytes": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t)]),
    "nst_forward_f32_nchw": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                       C.c_void_p, C.c_size_t, C.c_void_p]),
    "nst_forward_u8_nhwc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                      _c_float3, _c_float3, _c_float3, _c_float3, _c_float3, C.c_int, C.c_double,
                                      C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]),
    "nst_forward_u8_nhwc_fit": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                          _c_float3, _c_float3, _c_float3, _c_float3, _c_float3, C.c_int, C.c_double,
                                          C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]),
    "nst_composite_regions_batch_u8": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                                 C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_float), C.c_void_p, C.c_int, C.c_int,
                                                 C.c_int, C.c_void_p]),
    "nst_composite_from_crops_f32": (C.c_int, [C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                               C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.POINTER(C.c_float), C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "nst_resize_post_pack_u8": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_int, C.c_void_p,
                                          C.c_int, C.c_int, C.c_void_p]),
    "nst_voronoi_masks_f32": (C.c_int, [C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                        C.c_void_p]),
    "nst_flow_ema_f32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_void_p]),
    "nst_pattern_masks_f32": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_int),
                                        C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "nst_rotate_masks_f32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_void_p]),
    "nst_warp_masks_organic_f32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                                             C.POINTER(C.c_double), C.c_void_p, C.c_void_p]),
    "nst_feather_masks_f32": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "nst_mask_bboxes": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_void_p]),
    "nst_mask_scratch_floats": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "nst_lab_ema_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_void_p]),
    "nst_motion_blend_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double,
                                      C.c_double, C.c_double, C.c_void_p]),
    "nst_blend_pack_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "nst_motion_blend_u8_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double,
                                              C.c_double, C.c_double, C.c_void_p]),
    "nst_composite_regions_u8": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                           C.POINTER(C.c_int), C.POINTER(C.c_float), C.c_void_p, C.c_int, C.c_int,
                                           C.c_void_p]),
    "nst_composite_regions_f32": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                            C.POINTER(C.c_int), C.POINTER(C.c_float), C.c_void_p, C.c_int, C.c_int,
                                            C.c_void_p]),
    "nst_num_layers": (C.c_int, [C.c_void_p]),
    "nst_layer_info": (C.c_int, [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_double),
                                 C.POINTER(C.c_double), C.POINTER(C.c_int)]),
    "nst_profile_layers_u8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                        C.c_void_p, C.c_size_t, C.c_void_p, C.POINTER(C.c_float)]),
    "nst_describe_plan": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int]),
    "nst_launch_count": (C.c_longlong, [C.c_void_p]),
    "nst_last_error": (C.c_char_p, []),
    "nst_version": (C.c_char_p, []),
}

_lib = None


def load() -> C.CDLL:
    """Load the shared library once; raise if it is missing (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: the B200 stylisation path has no fallback. Build it with "
            "`python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc)."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)           # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class NstError(RuntimeError):
    """Raised for every non-zero status of the C ABI (pipeline.py:1487-1490 expects exceptions)."""


def check(rc: int, what: str) -> None:
    if rc != NST_OK:
        msg = load().nst_last_error()
        raise NstError(f"{what} failed (status {rc}): {msg.decode(errors='replace') if msg else ''}")


def describe_plan(arch: str, precision: str, n: int, h: int, w: int) -> str:
    """Planner dump for a shape (host only, no GPU)."""
    import ctypes
    buf = ctypes.create_string_buffer(16384)
    check(load().nst_describe_plan(ARCH[arch], PRECISION[precision], n, h, w, buf, len(buf)), "nst_describe_plan")
    return buf.value.decode()
