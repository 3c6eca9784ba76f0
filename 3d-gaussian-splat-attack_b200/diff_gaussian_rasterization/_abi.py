"""ctypes loader of libgsr_b200.so, the C ABI declared in include/gsr_b200.h.

Replaces the pybind extension module `diff_gaussian_rasterization._C`
(DGR/ext.cpp:15-19; DGR = submodules/gaussian-splatting/submodules/diff-gaussian-rasterization).
There is no CPU fallback: if the library is missing or a call fails, this raises.
"""
import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
LIB_PATH = os.environ.get("GSR_B200_LIB") or os.path.join(_ROOT, "lib", "libgsr_b200.so")  # override: kernel experiments
CSRC_DIR = os.path.join(_ROOT, "csrc")

GSR_OK, GSR_ERR_INVALID, GSR_ERR_CUDA, GSR_ERR_CAPACITY, GSR_ERR_PREFILTERED = 0, -1, -2, -3, -4
ABI_VERSION = 4
MAX_FUSED_VIEWS = 16  # GSR_MAX_FUSED_VIEWS


class GsrInputs(C.Structure):
    """struct gsr_inputs (include/gsr_b200.h)."""
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("W", C.c_int32), ("H", C.c_int32), ("K", C.c_int32),
        ("prefiltered", C.c_int32), ("debug", C.c_int32),
        ("scale_modifier", C.c_float), ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("raw_params", C.c_int32),
        ("bg", C.c_void_p), ("means3D", C.c_void_p), ("shs", C.c_void_p), ("colors_precomp", C.c_void_p),
        ("opacities", C.c_void_p), ("scales", C.c_void_p), ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p),
        ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p), ("campos", C.c_void_p), ("obj_feats", C.c_void_p),
        ("shs_rest", C.c_void_p), ("dL_dshs_rest", C.c_void_p),
    ]


class GsrViewRef(C.Structure):
    """struct gsr_view_ref (include/gsr_b200.h): one forwarded view as the multi-view backward sees it."""
    _fields_ = [
        ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p), ("campos", C.c_void_p),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("W", C.c_int32), ("H", C.c_int32),
        ("geom", C.c_void_p), ("radii", C.c_void_p), ("accum", C.c_void_p), ("dL_dmean2D", C.c_void_p),
    ]


class GsrViewFwd(C.Structure):
    """struct gsr_view_fwd (include/gsr_b200.h): one view of the multi-view forward preprocess."""
    _fields_ = [
        ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p), ("campos", C.c_void_p),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("W", C.c_int32), ("H", C.c_int32),
        ("geom", C.c_void_p), ("radii", C.c_void_p),
    ]


EXPORTS = {
    # name: (restype, argtypes)
    "gsr_abi_version": (C.c_int, []),
    "gsr_last_error": (C.c_char_p, []),
    "gsr_higher_msb": (C.c_uint32, [C.c_uint32]),
    "gsr_geom_bytes": (C.c_size_t, [C.c_int32]),
    "gsr_image_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "gsr_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int64, C.c_int32, C.c_int32]),
    "gsr_sort_temp_bytes": (C.c_size_t, [C.c_int64]),
    "gsr_forward_geometry": (C.c_int, [C.POINTER(GsrInputs), C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                       C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsr_forward_render": (C.c_int, [C.POINTER(GsrInputs), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsr_forward_render_speculative": (C.c_int, [C.POINTER(GsrInputs), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p,
                                                 C.c_void_p]),
    "gsr_forward_binning": (C.c_int, [C.POINTER(GsrInputs), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_size_t, C.c_void_p]),
    "gsr_forward_blend": (C.c_int, [C.POINTER(GsrInputs), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p]),
    "gsr_backward": (C.c_int, [C.POINTER(GsrInputs), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                               C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p] + [C.c_void_p] * 9 + [C.c_void_p]),
    "gsr_backward_accumulate": (C.c_int, [C.POINTER(GsrInputs), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_size_t, C.c_void_p] + [C.c_void_p] * 8 + [C.c_void_p]),
    "gsr_pgd_linf_update": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float, C.c_void_p]),
    "gsr_pgd_l2_update": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_float, C.c_float,
                                    C.c_void_p, C.c_void_p]),
    "gsr_image_to_u8_hwc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "gsr_detector_scale": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "gsr_backward_accumulate_phase": (C.c_int, [C.POINTER(GsrInputs), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                                C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p] + [C.c_void_p] * 8 +
                                      [C.c_int32, C.c_void_p]),
    "gsr_forward_preprocess_views": (C.c_int, [C.POINTER(GsrInputs), C.c_int32, C.POINTER(GsrViewFwd), C.c_void_p]),
    "gsr_forward_order": (C.c_int, [C.POINTER(GsrInputs), C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p,
                                    C.c_void_p]),
    "gsr_view_accum_bytes": (C.c_size_t, [C.c_int32]),
    "gsr_backward_blend_view": (C.c_int, [C.POINTER(GsrInputs), C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p]),
    "gsr_backward_preprocess_views": (C.c_int, [C.POINTER(GsrInputs), C.c_int32, C.POINTER(GsrViewRef), C.c_int32] +
                                      [C.c_void_p] * 7 + [C.c_void_p]),
    "gsr_mark_visible": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gsr_sort_pairs_u64": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                                     C.c_void_p, C.c_size_t, C.c_void_p]),
    "gsr_sort_u32_temp_bytes": (C.c_size_t, [C.c_int64]),
    "gsr_sort_pairs_u32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32,
                                     C.c_void_p, C.c_size_t, C.c_void_p]),
    "gsr_sort_pairs_u32_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                                            C.c_void_p]),
    "gsr_profile_enable": (C.c_int, [C.c_int]),
    "gsr_profile_read": (C.c_int, [C.c_void_p, C.c_void_p]),
    "gsr_launch_count": (C.c_int64, [C.c_int]),
    "gsr_debug_counters": (C.c_int, [C.c_void_p, C.c_int]),
    "gsr_debug_sort_profile": (C.c_int, [C.c_void_p, C.c_int]),
    "gsr_debug_fetch": (C.c_int64, [C.c_char_p, C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
}

STAGES = ("preprocess_fwd", "duplicate_keys", "sort", "tile_ranges", "blend_fwd", "blend_bwd", "preprocess_bwd")
_lib = None


def build(verbose=False):
    """Compile csrc/*.cu for sm_100a into lib/libgsr_b200.so (nvcc cross-compiles without a GPU)."""
    import subprocess
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", CSRC_DIR, "-j8"], stdout=out)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} not found: the CUDA library is not built. Run `python __graft_entry__.py build` "
                f"(or `make -C {CSRC_DIR}`). There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in EXPORTS.items():
            fn = getattr(L, name)  # AttributeError here = header/library mismatch
            fn.restype = res
            fn.argtypes = args
        if L.gsr_abi_version() != ABI_VERSION:
            raise ImportError(f"libgsr_b200.so ABI {L.gsr_abi_version()} != expected {ABI_VERSION}")
        _lib = L
    return _lib


def last_error():
    return lib().gsr_last_error().decode(errors="replace")


def check(status, what):
    """Map a gsr_status to the exception type the reference raises for the same condition."""
    if status == GSR_OK:
        return
    msg = last_error()
    if status == GSR_ERR_INVALID:
        raise Exception(msg)  # reference: plain Exception on argument misuse (__init__.py:191-195)
    raise RuntimeError(f"{what}: {msg} (status {status})")
