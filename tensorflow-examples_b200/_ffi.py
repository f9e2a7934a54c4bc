"""ctypes binding of include/b2mlp.h.  The library is csrc/libb2mlp.so, built by build.py."""
import ctypes
import os

from . import build

_LIB = None

# every symbol include/b2mlp.h declares (tests check the built library exports exactly these)
EXPORTS = [
    "b2_version", "b2_last_error", "b2_create", "b2_destroy", "b2_set_option", "b2_param_count",
    "b2_params_set", "b2_params_get", "b2_get_unique_id", "b2_comm_init", "b2_train_steps", "b2_eval",
    "b2_session_run_train", "b2_global_step", "b2_launch_count", "b2_kernel_times", "b2_kernel_class_name",
    "b2_linear_step_noalloc", "b2_peer_export", "b2_peer_import", "b2_matmul_bf16", "b2_random_normal",
]

UNIQUE_ID_BYTES = 128
OPT_FORCE_LAYERED = 1
OPT_KERNEL_TIMING = 2
KERNEL_CLASSES = 13
OPT_NO_PEER_EXCHANGE = 3
OPT_NO_BUCKET_OVERLAP = 4
OPT_NO_SPLIT3 = 6
OPT_NO_NARROW_HEAD = 7
OPT_NO_PERSISTENT = 8
OPT_MULTISTEP_FUSED = 9
OPT_SHARDED_APPLY = 10
OPT_ASYNC_PERIOD = 5
PEER_HANDLE_BYTES = 64


class B2Error(RuntimeError):
    """A non-zero return code from the C ABI; carries b2_last_error()."""

    def __init__(self, code, message):
        super().__init__("b2mlp error %d: %s" % (code, message))
        self.code = code


def lib_path():
    return build.LIB


def load_library():
    """dlopen csrc/libb2mlp.so.  Raises if it has not been built -- there is no fallback path."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise RuntimeError(
            "CUDA library %s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  There is no CPU fallback." % path)
    lib = ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
    c_void_p, c_int, c_long, c_float = ctypes.c_void_p, ctypes.c_int, ctypes.c_long, ctypes.c_float
    lib.b2_version.restype = ctypes.c_char_p
    lib.b2_last_error.restype = ctypes.c_char_p
    lib.b2_create.argtypes = [ctypes.POINTER(c_void_p), c_int, c_int, ctypes.POINTER(c_int), c_int, c_int, c_int, c_int]
    lib.b2_create.restype = c_int
    lib.b2_destroy.argtypes = [c_void_p]
    lib.b2_destroy.restype = None
    lib.b2_set_option.argtypes = [c_void_p, c_int, c_int]
    lib.b2_set_option.restype = c_int
    lib.b2_param_count.argtypes = [c_void_p]
    lib.b2_param_count.restype = c_long
    lib.b2_params_set.argtypes = [c_void_p, c_int, c_void_p, c_void_p, c_void_p]
    lib.b2_params_set.restype = c_int
    lib.b2_params_get.argtypes = [c_void_p, c_int, c_void_p, c_void_p, c_void_p]
    lib.b2_params_get.restype = c_int
    lib.b2_get_unique_id.argtypes = [c_void_p]
    lib.b2_get_unique_id.restype = c_int
    lib.b2_comm_init.argtypes = [c_void_p, c_int, c_int, c_void_p]
    lib.b2_comm_init.restype = c_int
    lib.b2_train_steps.argtypes = [c_void_p, c_void_p, c_void_p, c_long, c_int, c_long, c_int, c_long, c_float,
                                   c_void_p, c_void_p, c_void_p]
    lib.b2_train_steps.restype = c_int
    lib.b2_eval.argtypes = [c_void_p, c_void_p, c_void_p, c_long, c_void_p, c_void_p, c_void_p]
    lib.b2_eval.restype = c_int
    lib.b2_session_run_train.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_float,
                                         ctypes.POINTER(c_float), ctypes.POINTER(c_int)]
    lib.b2_session_run_train.restype = c_int
    lib.b2_global_step.argtypes = [c_void_p]
    lib.b2_global_step.restype = c_float
    lib.b2_launch_count.argtypes = [c_void_p]
    lib.b2_launch_count.restype = c_long
    lib.b2_kernel_times.argtypes = [c_void_p, ctypes.POINTER(c_float), ctypes.POINTER(c_long)]
    lib.b2_kernel_times.restype = c_int
    lib.b2_kernel_class_name.argtypes = [c_int]
    lib.b2_kernel_class_name.restype = ctypes.c_char_p
    lib.b2_linear_step_noalloc.argtypes = [c_void_p, c_void_p, c_int, c_long, c_float, c_void_p, c_void_p,
                                           c_void_p, c_void_p, c_void_p]
    lib.b2_linear_step_noalloc.restype = c_int
    lib.b2_peer_export.argtypes = [c_void_p, c_void_p]
    lib.b2_peer_export.restype = c_int
    lib.b2_peer_import.argtypes = [c_void_p, c_void_p]
    lib.b2_peer_import.restype = c_int
    lib.b2_matmul_bf16.argtypes = [c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                   c_void_p, c_void_p]
    lib.b2_matmul_bf16.restype = c_int
    lib.b2_random_normal.argtypes = [c_void_p, ctypes.c_long, ctypes.c_ulonglong, ctypes.c_ulonglong, ctypes.c_float,
                                     ctypes.c_float, c_void_p]
    lib.b2_random_normal.restype = c_int
    _LIB = lib
    return lib


def check(lib, rc):
    if rc != 0:
        raise B2Error(rc, lib.b2_last_error().decode("utf-8", "replace"))


def nccl_library_hint():
    """Path of the NCCL that ships with torch, exported as B2_NCCL_LIB for the C side's dlopen."""
    try:
        import nvidia.nccl as _n  # type: ignore
        cand = os.path.join(os.path.dirname(_n.__file__), "lib", "libnccl.so.2")
        if os.path.exists(cand):
            return cand
    except Exception:
        pass
    return None
