"""ctypes loader for libvsk.so (built in-tree by vntrseek_b200/csrc/Makefile).

There is no fallback: if the library is missing or the machine has no CUDA
device the import / vsk_init fails loudly.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libvsk.so")

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not built: run `make -C vntrseek_b200/csrc` "
                "(or __graft_entry__.build()); libvsk has no CPU fallback")
        _lib = C.CDLL(LIB_PATH)
        _declare(_lib)
    return _lib


def _p(dtype):
    return np.ctypeslib.ndpointer(dtype, flags="C_CONTIGUOUS")


def _declare(L):
    vp = C.c_void_p
    L.vsk_init.argtypes = [C.POINTER(vp), C.c_int, _p(np.uint8)]
    L.vsk_destroy.argtypes = [vp]
    L.vsk_destroy.restype = None
    L.vsk_last_error.argtypes = [vp]
    L.vsk_last_error.restype = C.c_char_p
    L.vsk_launch_count.argtypes = [vp]
    L.vsk_launch_count.restype = C.c_uint64
    L.vsk_sync.argtypes = [vp]
    L.vsk_load_distance_file.argtypes = [C.c_char_p, _p(np.uint8)]
    L.vsk_profset_upload.argtypes = [vp, C.c_int32, _p(np.int64), _p(np.uint16), C.POINTER(vp)]
    L.vsk_profset_free.argtypes = [vp, vp]
    L.vsk_profset_free.restype = None
    L.vsk_pair_dp.argtypes = [vp, vp, vp, C.c_int64, _p(np.int32), _p(np.int32), _p(np.uint8),
                              C.c_int, _p(np.int32)]
    L.vsk_dp_batch_prepare.argtypes = [vp, vp, vp, C.c_int64, _p(np.int32), _p(np.int32),
                                       _p(np.uint8), C.c_int, C.POINTER(vp)]
    L.vsk_dp_batch_run.argtypes = [vp, vp]
    L.vsk_dp_batch_fetch.argtypes = [vp, vp, _p(np.int32)]
    L.vsk_dp_batch_cells.argtypes = [vp]
    L.vsk_dp_batch_cells.restype = C.c_int64
    L.vsk_dp_batch_free.argtypes = [vp, vp]
    L.vsk_dp_batch_free.restype = None
    L.vsk_lcs_pairs.argtypes = [vp, C.c_char_p, _p(np.int64), C.c_char_p, _p(np.int64), C.c_int64,
                                _p(np.int32), _p(np.int32), _p(np.int32)]
    L.vsk_ed_pairs.argtypes = [vp, C.c_char_p, _p(np.int64), _p(np.int32), C.c_char_p, _p(np.int64),
                               _p(np.int32), C.c_int64, C.c_int, _p(np.int32)]
