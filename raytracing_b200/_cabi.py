"""ctypes binding of librtb.so (include/rtb.h).  There is no fallback: if the CUDA library is
missing or no GPU is present the product path raises."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTB_LIB_PATH") or os.path.join(_HERE, "csrc", "librtb.so")  # override: kernel tuning only

RTB_PTR_HOST, RTB_PTR_DEVICE = 0, 1
RTB_F_INTERFACE, RTB_F_EXACT_BASIS, RTB_F_NO_TAIL_ZERO, RTB_F_STATIC_GRID, RTB_F_RAY_QUEUE, RTB_F_SORT_FAN = 1, 2, 4, 8, 16, 32
RTB_OPT_QUEUE, RTB_OPT_TRAJ_STORE, RTB_OPT_SHOOT_GROUPS, RTB_OPT_SHOOT_SPECULATE = 1, 2, 3, 4
RTB_SWEEP_SNELL, RTB_SWEEP_CLOSURE, RTB_SWEEP_PX_CV = 1, 2, 3
RTB_ABI_VERSION = 2
RTB_OK, RTB_ERR_ARG, RTB_ERR_CUDA, RTB_ERR_NOFIELD, RTB_ERR_NOMEM = 0, -1, -2, -3, -4

# every symbol include/rtb.h declares
EXPORTS = ("rtb_create", "rtb_destroy", "rtb_set_stream", "rtb_last_error", "rtb_abi_version",
           "rtb_set_field", "rtb_field_eval", "rtb_trace", "rtb_fp64_peak", "rtb_shoot_coarse", "rtb_shoot_refine", "rtb_wavefront",
           "rtb_fit_field", "rtb_gradient", "rtb_get_field",
           "rtb_gather_alloc", "rtb_gather_open", "rtb_gather_close", "rtb_set_gather", "rtb_set_gather_cyclic",
           "rtb_sweep", "rtb_sweep_noalloc", "rtb_set_option", "rtb_build_id", "rtb_kernel_id")


class RtbSpline(C.Structure):
    _fields_ = [("tu", C.c_void_p), ("tv", C.c_void_p), ("c", C.c_void_p),
                ("nu", C.c_int32), ("nv", C.c_int32), ("ku", C.c_int32), ("kv", C.c_int32)]


class RtbParams(C.Structure):
    _fields_ = [("method", C.c_int32), ("flags", C.c_uint32), ("ptr_kind", C.c_int32), ("out_stride", C.c_int32),
                ("max_size", C.c_int64), ("gamma", C.c_double), ("step", C.c_double),
                ("limx_i", C.c_double), ("limx_s", C.c_double), ("limy_i", C.c_double), ("limy_s", C.c_double),
                ("steps", C.c_void_p), ("max_sizes", C.c_void_p)]


class RtbStats(C.Structure):
    _fields_ = [("kernel_ms", C.c_double), ("total_ms", C.c_double), ("ray_steps", C.c_int64),
                ("launches", C.c_int64), ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64),
                ("nonfinite_rays", C.c_int64)]


class RtbSweepSpec(C.Structure):
    _fields_ = [("steps", C.c_void_p), ("max_sizes", C.c_void_p), ("C", C.c_int64), ("metric", C.c_int32),
                ("skip_lo", C.c_int64), ("skip_hi", C.c_int64)]


class RtbShootProblem(C.Structure):
    _fields_ = [("src_x", C.c_double), ("src_y", C.c_double), ("steps", C.c_void_p), ("max_sizes", C.c_void_p),
                ("C", C.c_int64), ("target_x", C.c_void_p), ("target_y", C.c_void_p), ("Tn", C.c_int64)]


class RtbError(RuntimeError):
    def __init__(self, code, text):
        super().__init__("librtb error %d: %s" % (code, text))
        self.code = code


_lib = None


def load():
    """Load librtb.so and declare the prototypes.  Raises if the extension was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("raytracing_b200: %s is missing -- build it with `make -C raytracing_b200/csrc` "
                          "(or __graft_entry__.build()); there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, i32, u32, i64 = C.c_void_p, C.c_int32, C.c_uint32, C.c_int64
    lib.rtb_create.restype = C.c_int
    lib.rtb_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.rtb_destroy.restype = None
    lib.rtb_destroy.argtypes = [vp]
    lib.rtb_set_stream.restype = C.c_int
    lib.rtb_set_stream.argtypes = [vp, vp]
    lib.rtb_last_error.restype = C.c_char_p
    lib.rtb_last_error.argtypes = [vp]
    lib.rtb_abi_version.restype = C.c_int
    lib.rtb_abi_version.argtypes = []
    lib.rtb_set_field.restype = C.c_int
    lib.rtb_set_field.argtypes = [vp, C.POINTER(RtbSpline), C.POINTER(RtbSpline), C.POINTER(RtbSpline)]
    lib.rtb_field_eval.restype = C.c_int
    lib.rtb_field_eval.argtypes = [vp, i32, u32, vp, vp, i64, vp, vp, vp]
    lib.rtb_trace.restype = C.c_int
    lib.rtb_trace.argtypes = [vp, C.POINTER(RtbParams), vp, vp, vp, i64, vp, vp, vp, vp, vp, C.POINTER(RtbStats)]
    lib.rtb_fp64_peak.restype = C.c_int
    lib.rtb_fp64_peak.argtypes = [vp, C.POINTER(C.c_double)]
    lib.rtb_shoot_coarse.restype = C.c_int
    lib.rtb_shoot_coarse.argtypes = [vp, C.POINTER(RtbParams), C.POINTER(RtbShootProblem), vp, i64, i64, vp, vp, vp,
                                     C.POINTER(RtbStats)]
    lib.rtb_shoot_refine.restype = C.c_int
    lib.rtb_shoot_refine.argtypes = [vp, C.POINTER(RtbParams), C.POINTER(RtbShootProblem), vp, vp, vp, vp,
                                     C.POINTER(i64), C.POINTER(RtbStats)]
    lib.rtb_wavefront.restype = C.c_int
    lib.rtb_wavefront.argtypes = [vp, i32, vp, vp, i64, i64, vp, i32, vp]
    lib.rtb_fit_field.restype = C.c_int
    lib.rtb_fit_field.argtypes = [vp, i32, vp, i64, vp, i64, vp, C.c_double, C.POINTER(RtbStats)]
    lib.rtb_gradient.restype = C.c_int
    lib.rtb_gradient.argtypes = [vp, i32, vp, i64, i64, C.c_double, vp, vp]
    lib.rtb_get_field.restype = C.c_int
    lib.rtb_get_field.argtypes = [vp, i32, C.POINTER(i32), C.POINTER(i32), vp, vp, vp]
    lib.rtb_gather_alloc.restype = C.c_int
    lib.rtb_gather_alloc.argtypes = [vp, i64, C.POINTER(vp), vp]
    lib.rtb_gather_open.restype = C.c_int
    lib.rtb_gather_open.argtypes = [vp, vp, C.POINTER(vp)]
    lib.rtb_gather_close.restype = C.c_int
    lib.rtb_gather_close.argtypes = [vp, vp, i32]
    lib.rtb_set_gather.restype = C.c_int
    lib.rtb_set_gather.argtypes = [vp, i32, C.POINTER(vp), i64, i64]
    lib.rtb_set_gather_cyclic.restype = C.c_int
    lib.rtb_set_gather_cyclic.argtypes = [vp, i32, C.POINTER(vp), i64, i64, i32, i32]
    lib.rtb_sweep.restype = C.c_int
    lib.rtb_sweep.argtypes = [vp, C.POINTER(RtbParams), vp, vp, vp, i64, C.POINTER(RtbSweepSpec), vp, vp, vp,
                              C.POINTER(RtbStats)]
    lib.rtb_sweep_noalloc.restype = C.c_int
    lib.rtb_sweep_noalloc.argtypes = lib.rtb_sweep.argtypes
    lib.rtb_set_option.restype = C.c_int
    lib.rtb_set_option.argtypes = [vp, i32, i32]
    lib.rtb_build_id.restype = C.c_char_p
    lib.rtb_build_id.argtypes = []
    lib.rtb_kernel_id.restype = C.c_char_p
    lib.rtb_kernel_id.argtypes = []
    if lib.rtb_abi_version() != RTB_ABI_VERSION:
        raise ImportError("raytracing_b200: %s has ABI version %d, this package binds version %d -- rebuild it"
                          % (LIB_PATH, lib.rtb_abi_version(), RTB_ABI_VERSION))
    _lib = lib
    return lib
