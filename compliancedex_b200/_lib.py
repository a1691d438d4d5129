"""ctypes binding of libcdx.so (C ABI declared in include/cdx.h).  There is NO fallback: if the library is
missing or fails to load, importing the compute classes raises."""
import ctypes
import os
import weakref

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcdx.so")

c_void_p, c_int, c_i64, c_double, c_u64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_double, ctypes.c_uint64


class PregraspParams(ctypes.Structure):
    """struct cdx_pregrasp_params (include/cdx.h)."""
    _fields_ = [
        ("num_copies", ctypes.c_int32),
        ("coeffs", (ctypes.c_float * 4) * 4),
        ("weights", ctypes.c_float * 4),
        ("ref_q", ctypes.c_float * 16),
        ("box_lo", (ctypes.c_float * 3) * 4),
        ("box_hi", (ctypes.c_float * 3) * 4),
        ("com", ctypes.c_float * 3),
        ("dummy_k", ctypes.c_float),
        ("dummy_M", ctypes.c_float),
        ("gravity", ctypes.c_int32),
        ("uncertainty", c_double),
        ("optimize_target", ctypes.c_int32),
        ("optimize_palm", ctypes.c_int32),
        ("cos_mu", ctypes.c_float),
        ("noise_seed", c_u64),
        ("env_offset", c_i64),
        ("lr", c_double * 5),
        ("beta1", c_double),
        ("beta2", c_double),
        ("eps", c_double),
        ("compliance_min", c_double),
        ("track_after", ctypes.c_int32),
        ("variant", ctypes.c_int32),
        ("rms_alpha", ctypes.c_double),
    ]


# name -> (restype, argtypes); every symbol include/cdx.h declares
SIGNATURES = {
    "cdx_last_error": (ctypes.c_char_p, []),
    "cdx_version": (c_int, []),
    "cdx_create": (c_int, [c_int, ctypes.POINTER(c_void_p)]),
    "cdx_destroy": (c_int, [c_void_p]),
    "cdx_gpis_fit": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_double, c_i64, c_double, c_int, c_double, c_void_p]),
    "cdx_gpis_fit_profile": (c_int, [c_void_p, c_int]),
    "cdx_gpis_fit_profile_read": (c_int, [c_void_p, c_void_p, c_void_p]),
    "cdx_gpis_set_state": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_double, c_double, c_i64, c_int, c_double, c_void_p]),
    "cdx_gpis_get": (c_int, [c_void_p, c_int, c_void_p, c_void_p]),
    "cdx_gpis_info": (c_int, [c_void_p, ctypes.POINTER(c_i64), ctypes.POINTER(c_double), ctypes.POINTER(c_double),
                              ctypes.POINTER(c_double), ctypes.POINTER(c_int), c_void_p]),
    "cdx_gpis_query": (c_int, [c_void_p, c_void_p, c_i64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "cdx_gpis_set_var_mode": (c_int, [c_void_p, c_int]),
    "cdx_gpis_set_ext_mode": (c_int, [c_void_p, c_int]),
    "cdx_gpis_set_i8_variant": (c_int, [c_void_p, c_int]),
    "cdx_gpis_query_host": (c_int, [c_void_p, c_void_p, c_i64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "cdx_gpis_query_weights": (c_int, [c_void_p, c_void_p, c_void_p, c_i64, c_void_p, c_void_p, c_void_p]),
    "cdx_grid_surface": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "cdx_grid_surface_static": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "cdx_sdf_forward": (c_int, [c_void_p, c_void_p, c_i64, c_void_p, c_i64, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "cdx_sdf_set_mode": (c_int, [c_void_p, c_int]),
    "cdx_sdf_backward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_i64, c_int, c_void_p, c_void_p]),
    "cdx_hand_load": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "cdx_hand_fk": (c_int, [c_void_p, c_void_p, c_void_p, c_i64, c_void_p, c_void_p]),
    "cdx_hand_fk_recursive": (c_int, [c_void_p, c_void_p, c_void_p, c_i64, c_void_p, c_void_p]),
    "cdx_collision_cost": (c_int, [c_void_p, c_void_p, c_i64, c_int, c_void_p, c_void_p, c_int, c_double, c_void_p, c_void_p, c_void_p]),
    "cdx_pregrasp_init": (c_int, [c_void_p, ctypes.POINTER(PregraspParams), c_i64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "cdx_pregrasp_run": (c_int, [c_void_p, c_int, c_void_p]),
    "cdx_pregrasp_set_tuning": (c_int, [c_void_p, c_int, c_int, c_int]),
    "cdx_pregrasp_nan_count": (c_int, [c_void_p, ctypes.POINTER(c_i64), ctypes.POINTER(c_int), c_void_p]),
    "cdx_pregrasp_generation": (c_i64, [c_void_p, ctypes.POINTER(c_i64)]),
    "cdx_pregrasp_eval": (c_int, [c_void_p, c_i64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "cdx_pregrasp_results": (c_int, [c_void_p, c_i64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "cdx_pregrasp_state": (c_int, [c_void_p, c_i64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "cdx_pregrasp_latched_margin": (c_int, [c_void_p, c_i64, c_void_p, c_void_p, c_void_p]),
    "cdx_pregrasp_argmin": (c_int, [c_void_p, c_void_p, c_void_p]),
    "cdx_pregrasp_profile": (c_int, [c_void_p, c_int]),
    "cdx_pregrasp_profile_read": (c_int, [c_void_p, c_void_p, ctypes.POINTER(c_int), c_void_p]),
    "cdx_launch_count": (c_i64, [c_void_p]),
}

CDX_OK, CDX_ERR_ARG, CDX_ERR_CUDA, CDX_ERR_NOT_SPD, CDX_ERR_STATE, CDX_ERR_ALLOC = range(6)
KERNEL_IDS = {"tps": 0, "rbf": 1, "joint": 2}
MAT_E11, MAT_CHOL, MAT_CHOL_INV, MAT_E11_INV, VEC_ALPHA, VEC_Y1, MAT_X1 = range(7)

_lib = None


def load():
    """Load libcdx.so and bind every declared symbol; raises if the extension is missing (no CPU fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(cdx has no CPU fallback)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class CdxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"cdx error {code}: {msg}")
        self.code = code


def check(rc):
    if rc != CDX_OK:
        msg = load().cdx_last_error().decode("utf-8", "replace")
        if rc == CDX_ERR_NOT_SPD:
            import torch
            raise torch.linalg.LinAlgError(f"cdx: {msg}")
        if rc == CDX_ERR_ARG:
            raise ValueError(f"cdx: {msg}")
        raise CdxError(rc, msg)


_contexts = {}


def context(device_index):
    """One cdx_ctx per device, created lazily."""
    lib = load()
    if device_index not in _contexts:
        h = c_void_p()
        check(lib.cdx_create(int(device_index), ctypes.byref(h)))
        _contexts[device_index] = h
    return _contexts[device_index]


class Context(ctypes.c_void_p):
    """A cdx_ctx handle that is released with its last Python reference (cdx_destroy frees the factorised state -
    4 Np^2 doubles, 0.5 GB at N = 4096 - the INT8 slices and the pregrasp buffers).  Passed wherever the C ABI takes
    a cdx_ctx*."""

    def close(self):
        fin = getattr(self, "_finalizer", None)
        if fin is not None:
            fin()


def new_context(device_index):
    lib = load()
    h = Context()
    check(lib.cdx_create(int(device_index), ctypes.byref(h)))
    h._finalizer = weakref.finalize(h, lib.cdx_destroy, c_void_p(h.value))
    h._finalizer.atexit = False          # at interpreter exit the driver reclaims everything; do not race its teardown
    return h
