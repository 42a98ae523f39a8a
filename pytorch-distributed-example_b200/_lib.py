"""ctypes binding of csrc/libgradavg.so (the C ABI declared in include/gradavg.h)."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

GA_OK, GA_EINVAL, GA_ECUDA, GA_ECAPACITY, GA_ESTATE, GA_ETIMEOUT = 0, -1, -2, -3, -4, -5
VARIANT_AUTO, VARIANT_ONESHOT, VARIANT_TWOSHOT = 0, 1, 2
VARIANT_ONESHOT_PUSH, VARIANT_TWOSHOT_PUSH, VARIANT_TWOSHOT_PIPE = 3, 4, 5
VARIANT_HOSTED_DIRECT = 6
VARIANT_VIEW, VARIANT_NVLS = 7, 8
OP_F32_AVG, OP_I32_SUM, OP_F32_SUM = 0, 1, 2
OPT_ONESHOT_MAX_BYTES, OPT_TIMEOUT_MS, OPT_MAX_BLOCKS, OPT_PIPE_GROUP = 1, 2, 3, 4
OPT_HOST_CHUNK_BYTES = 5
OPT_TRACE = 6
OPT_VIEW_CHUNK_TILES = 7
OPT_VIEW_ONESHOT_MAX_BYTES = 8
OPT_VIEW_MIN_BYTES = 9
VIEW_MULTICAST = 1
VIEW_HANDLE_BYTES = 128
IPC_HANDLE_BYTES = 64
MAX_WORLD = 8
TILE_BYTES = 2048

# name -> (restype, argtypes); the single source for the loader and for tests/test_abi.py
_c = ctypes
SIGNATURES = {
    "ga_last_error": (_c.c_char_p, []),
    "ga_create": (_c.c_int, [_c.c_int, _c.c_int, _c.c_int, _c.c_size_t, _c.POINTER(_c.c_void_p)]),
    "ga_create_local": (_c.c_int, [_c.c_int, _c.c_int, _c.c_size_t, _c.POINTER(_c.c_void_p)]),
    "ga_export_handle": (_c.c_int, [_c.c_void_p, _c.c_char_p]),
    "ga_open_peers": (_c.c_int, [_c.c_void_p, _c.c_char_p]),
    "ga_average": (_c.c_int, [_c.c_void_p, _c.POINTER(_c.c_void_p), _c.POINTER(_c.c_int64), _c.c_int,
                              _c.c_void_p, _c.c_int]),
    "ga_allreduce_tensors": (_c.c_int, [_c.c_void_p, _c.POINTER(_c.c_void_p), _c.POINTER(_c.c_int64), _c.c_int,
                                        _c.c_int, _c.c_void_p, _c.c_int]),
    "ga_allreduce_tensors_ex": (_c.c_int, [_c.c_void_p, _c.POINTER(_c.c_void_p), _c.POINTER(_c.c_int64), _c.c_int,
                                           _c.c_int, _c.c_void_p, _c.c_int, _c.c_int]),
    "ga_broadcast_tensors": (_c.c_int, [_c.c_void_p, _c.POINTER(_c.c_void_p), _c.POINTER(_c.c_int64), _c.c_int,
                                        _c.c_int, _c.c_void_p]),
    "ga_view_create": (_c.c_int, [_c.c_void_p, _c.c_size_t, _c.c_int]),
    "ga_view_export": (_c.c_int, [_c.c_void_p, _c.c_char_p]),
    "ga_view_open": (_c.c_int, [_c.c_void_p, _c.c_char_p]),
    "ga_view_bind": (_c.c_int, [_c.c_void_p]),
    "ga_view_ptr": (_c.c_void_p, [_c.c_void_p, _c.c_int]),
    "ga_view_destroy": (_c.c_int, [_c.c_void_p]),
    "ga_view_bytes": (_c.c_size_t, [_c.c_void_p]),
    "ga_view_has_multicast": (_c.c_int, [_c.c_void_p]),
    "ga_multicast_supported": (_c.c_int, [_c.c_int]),
    "ga_average_view": (_c.c_int, [_c.c_void_p, _c.c_size_t, _c.c_int, _c.c_void_p, _c.c_int]),
    "ga_allreduce_bucket": (_c.c_int, [_c.c_void_p, _c.c_size_t, _c.c_int, _c.c_void_p, _c.c_int]),
    "ga_average_host": (_c.c_int, [_c.c_void_p, _c.POINTER(_c.c_void_p), _c.POINTER(_c.c_int64), _c.c_int,
                                   _c.c_void_p, _c.c_int]),
    "ga_bucket_acquire": (_c.c_int, [_c.c_void_p, _c.c_void_p]),
    "ga_send_ptr": (_c.c_void_p, [_c.c_void_p, _c.c_int]),
    "ga_recv_ptr": (_c.c_void_p, [_c.c_void_p, _c.c_int]),
    "ga_bucket_offset": (_c.c_int64, [_c.POINTER(_c.c_int64), _c.c_int, _c.c_int]),
    "ga_world": (_c.c_int, [_c.c_void_p]),
    "ga_rank": (_c.c_int, [_c.c_void_p]),
    "ga_n_local": (_c.c_int, [_c.c_void_p]),
    "ga_capacity": (_c.c_size_t, [_c.c_void_p]),
    "ga_pick_variant": (_c.c_int, [_c.c_void_p, _c.c_size_t]),
    "ga_auto_variant": (_c.c_int, [_c.c_int, _c.c_int, _c.c_size_t, _c.c_size_t]),
    "ga_launch_count": (_c.c_uint64, [_c.c_void_p]),
    "ga_last_grid": (_c.c_int, [_c.c_void_p, _c.POINTER(_c.c_int), _c.POINTER(_c.c_int)]),
    "ga_set_option": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_int64]),
    "ga_read_trace": (_c.c_int, [_c.c_void_p, _c.POINTER(_c.c_uint64), _c.c_size_t, _c.POINTER(_c.c_int),
                                 _c.POINTER(_c.c_int)]),
    "ga_status": (_c.c_int, [_c.c_void_p]),
    "ga_synchronize": (_c.c_int, [_c.c_void_p, _c.c_void_p]),
    "ga_average_static": (_c.c_int, [_c.c_int, _c.c_int, _c.POINTER(_c.c_void_p), _c.POINTER(_c.c_int64), _c.c_int,
                                     _c.c_void_p, _c.c_int]),
    "ga_static_status": (_c.c_int, [_c.c_int, _c.c_void_p]),
    "ga_close_peers": (_c.c_int, [_c.c_void_p]),
    "ga_destroy": (_c.c_int, [_c.c_void_p]),
}


class GradAvgError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("libgradavg error %d: %s" % (code, message))
        self.code = code


def lib_path():
    return os.path.join(_HERE, "csrc", "libgradavg.so")


def load_library():
    """dlopen csrc/libgradavg.so.  Raises (never falls back) if it has not been built."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise GradAvgError(GA_ESTATE, "%s is missing: build it with `python -c 'import __graft_entry__ as g; "
                           "g.build()'` (nvcc, sm_100a). There is no CPU or PyTorch fallback." % path)
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)        # AttributeError here = header and library out of step
        fn.restype = res
        fn.argtypes = args
    _LIB = lib
    return lib


def check(rc):
    if rc != 0:
        raise GradAvgError(rc, load_library().ga_last_error().decode("utf-8", "replace"))
    return rc
