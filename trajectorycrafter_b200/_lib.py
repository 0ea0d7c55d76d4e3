"""ctypes binding of libtcwarp.so (include/tcwarp.h).  No fallback: a missing library is an error."""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_longlong, c_size_t, c_void_p

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TCW_LIB", os.path.join(HERE, "libtcwarp.so"))   # TCW_LIB: A/B builds of the same library

# flags (include/tcwarp.h)
TCW_PER_UNIT_MAX = 1
TCW_CLEAN = 4
TCW_IS_IMAGE = 8
TCW_ATOMIC_F32 = 16
TCW_HOST_MASK_U8 = 32
TCW_MASK_U8 = 64
TCW_PACK = 128
TCW_IPC_HANDLE_BYTES = 64
TCW_ZB_STATIC_MAX_PIXELS = 1 << 20
# status bits a call ORs into its range_flag word
TCW_STATUS_RANGE = 1
TCW_STATUS_UNORDERED = 2
TCW_STATUS_SRC_INDEX = 4
TCW_STATUS_BOUNDS = 8
VERSION = 200

# every symbol include/tcwarp.h declares: name -> (restype, argtypes)
_FP, _IP, _VP = c_void_p, c_void_p, c_void_p      # device/host pointers are passed as raw addresses
SIGNATURES = {
    "tcw_create": (c_int, [c_int, POINTER(c_void_p)]),
    "tcw_destroy": (None, [c_void_p]),
    "tcw_version": (c_int, []),
    "tcw_last_error": (c_char_p, []),
    "tcw_launch_count": (c_longlong, [c_void_p]),
    "tcw_profile_enable": (c_int, [c_void_p, c_int]),
    "tcw_profile_read": (c_int, [c_void_p, POINTER(c_double), POINTER(c_longlong), POINTER(c_double)]),
    "tcw_workspace_bytes": (c_int, [c_void_p, c_int, c_int, c_int, c_int, POINTER(c_size_t)]),
    "tcw_forward_warp": (c_int, [c_void_p, _FP, _FP, _FP, _IP, _FP, c_int, c_int, c_int, c_int, c_int,
                                 _VP, _VP, _FP, _FP, _IP, _VP, c_size_t, c_void_p]),
    "tcw_forward_warp_scatter": (c_int, [c_void_p, _FP, _FP, _FP, _IP, _FP, _IP, c_int, c_int, c_int, c_int, c_int, c_int,
                                         _VP, _VP, _FP, _FP, _IP, _VP, c_size_t, c_void_p]),
    "tcw_forward_warp_host": (c_int, [c_void_p, _FP, _FP, _FP, _IP, _FP, c_int, c_int, c_int, c_int, c_int,
                                      _VP, _VP, _FP, _FP, POINTER(c_int)]),
    "tcw_forward_warp_pack": (c_int, [c_void_p, _FP, _FP, _FP, _IP, _FP, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                      _FP, _FP, _IP, _VP, c_size_t, c_void_p]),
    "tcw_project": (c_int, [c_void_p, _FP, _IP, _FP, c_int, c_int, c_int, c_int, _FP, c_void_p]),
    "tcw_splat": (c_int, [c_void_p, _FP, _FP, _FP, _FP, _FP, c_int, c_int, c_int, c_int, c_int,
                          _FP, _FP, _IP, _VP, c_size_t, c_void_p]),
    "tcw_clean_points": (c_int, [c_void_p, _FP, _FP, c_int, c_int, c_int, _VP, _VP, c_void_p]),
    "tcw_resize_pack": (c_int, [c_void_p, _FP, _FP, c_int, c_int, c_int, c_int, c_int, _FP, _FP, c_void_p]),
    "tcw_pc_lift": (c_int, [c_void_p, _FP, _FP, _FP, _FP, c_int, c_int, c_int, c_int, _FP, _FP, c_void_p]),
    "tcw_pc_render_zbuffer": (c_int, [c_void_p, _FP, _FP, POINTER(c_float), c_longlong, c_int, c_int, c_int, _FP, _FP, _FP, _VP, c_size_t, c_void_p]),
    "tcw_pc_render_zbuffer_first": (c_int, [c_void_p, _FP, _FP, POINTER(c_float), c_longlong, c_int, c_int, c_int, _FP, _FP, _FP, _VP, c_size_t, c_void_p]),
    "tcw_pc_render_zbuffer_static": (c_int, [c_void_p, _FP, _FP, POINTER(c_float), c_longlong, c_int, c_int, c_int, _FP, _FP, _FP, c_void_p]),
    "tcw_pc_render_disc": (c_int, [c_void_p, _FP, _FP, _FP, POINTER(c_float), c_longlong, c_int, c_int, c_float, POINTER(c_int), POINTER(c_float), c_int, _FP, _FP, _VP, c_size_t, c_void_p]),
    "tcw_pc_render_bilinear": (c_int, [c_void_p, _FP, _FP, _FP, POINTER(c_float), c_longlong, c_int, c_int, c_float, _FP, _FP, _VP, c_size_t, c_void_p]),
    "tcw_depth_post": (c_int, [c_void_p, _FP, c_int, c_int, c_int, c_int, c_float, c_float, c_float, _FP, _FP, c_void_p]),
    "tcw_ipc_alloc": (c_int, [c_void_p, c_size_t, POINTER(c_void_p), c_void_p]),
    "tcw_ipc_free": (c_int, [c_void_p, c_void_p]),
    "tcw_ipc_open": (c_int, [c_void_p, c_void_p, POINTER(c_void_p)]),
    "tcw_ipc_close": (c_int, [c_void_p, c_void_p]),
}

_lib = None


class TcwError(RuntimeError):
    pass


def load() -> ctypes.CDLL:
    """Load libtcwarp.so; raise (never fall back) if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise TcwError(
            f"{LIB_PATH} is missing: the CUDA library has not been built. Run "
            "`python -c 'import __graft_entry__ as g; g.build()'` (or trajectorycrafter_b200/build.py). "
            "There is no CPU or torch fallback for this path.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.tcw_version() != VERSION:
        raise TcwError(f"libtcwarp.so version {lib.tcw_version()} does not match the Python host ({VERSION}): rebuild it")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().tcw_last_error()
        raise TcwError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")
