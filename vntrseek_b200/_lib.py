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
    L.vsk_pair_dp_cyclic.argtypes = [vp, vp, vp, C.c_int64, _p(np.int32), _p(np.int32), _p(np.int32)]
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


class DpDesc(C.Structure):
    _fields_ = [("left_off", C.c_int64), ("top_off", C.c_int64), ("l1", C.c_int32), ("l2", C.c_int32),
                ("br", C.c_int32), ("out_idx", C.c_int32), ("svector", C.c_double)]


class ProcluParams(C.Structure):
    _fields_ = [("cutoff", C.c_int32), ("maxerrors", C.c_int32), ("option_r", C.c_int32),
                ("count_cells", C.c_int32), ("key_budget", C.c_int64), ("links_on_device", C.c_int32),
                ("reserved_", C.c_int32), ("progress", C.c_void_p), ("progress_user", C.c_void_p)]


class ProcluStats(C.Structure):
    _fields_ = [("candidates_considered", C.c_uint64), ("lcs_alignments", C.c_uint64),
                ("profile_alignments", C.c_uint64), ("profile_alignments_success", C.c_uint64),
                ("confirmed", C.c_uint64), ("dp_cells", C.c_int64), ("candidate_keys", C.c_int64),
                ("chunks", C.c_int64), ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64),
                ("dp_alg_bytes", C.c_int64), ("dp_cells_sweep", C.c_int64), ("dp_alg_bytes_sweep", C.c_int64)]


class JoinStats(C.Structure):
    _fields_ = [("clusters_read", C.c_int32), ("clusters_after_join", C.c_int32), ("maxdepth", C.c_int32),
                ("pad_", C.c_int32), ("refs_read", C.c_int64), ("refs_written", C.c_int64),
                ("reads_read", C.c_int64), ("reads_written", C.c_int64)]


class RankStats(C.Structure):
    _fields_ = [("files", C.c_int64), ("read_lines", C.c_int64), ("map_rows", C.c_int64),
                ("rank_rows", C.c_int64), ("reads_ranked", C.c_int64)]


LINK_DTYPE = np.dtype([("refkey", np.int32), ("readkey", np.int32), ("dir", np.int32),
                       ("lerr", np.int32), ("rerr", np.int32), ("dsum", np.int32), ("sumw", np.int32)])


def _declare2(L):
    vp = C.c_void_p
    L.vsk_trset_load_leb36.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
    L.vsk_trset_load_leb36_raw.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
    L.vsk_trset_load_leb36_cached.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.POINTER(vp),
                                              C.POINTER(C.c_int)]
    L.vsk_trset_from_arrays.argtypes = [C.c_int32, _p(np.int32), _p(np.int32), _p(np.int64), _p(np.uint16),
                                        _p(np.int64), _p(np.uint8), C.c_int, C.c_int, C.c_int, C.c_int,
                                        C.POINTER(vp)]
    L.vsk_trset_apply_rotindex.argtypes = [vp, C.c_char_p]
    L.vsk_trsets_check_unique.argtypes = [vp, vp, C.POINTER(C.c_int32)]
    L.vsk_trset_pin.argtypes = [vp]
    L.vsk_trset_size.argtypes = [vp]
    L.vsk_trset_size.restype = C.c_int32
    L.vsk_trset_error.argtypes = [vp]
    L.vsk_trset_error.restype = C.c_char_p
    L.vsk_trset_free.argtypes = [vp]
    L.vsk_trset_free.restype = None
    L.vsk_trset_checksum.argtypes = [vp]
    L.vsk_trset_checksum.restype = C.c_uint64
    L.vsk_refs_create.argtypes = [vp, vp, C.POINTER(vp)]
    L.vsk_refs_create_cached.argtypes = [vp, vp, C.c_char_p, C.POINTER(vp)]
    L.vsk_index_cache_build.argtypes = [vp, C.c_char_p, C.POINTER(C.c_int)]
    L.vsk_index_cache_verify.argtypes = [vp, C.c_char_p]
    L.vsk_refs_free.argtypes = [vp, vp]
    L.vsk_refs_free.restype = None
    L.vsk_refs_seed_hits.argtypes = [vp]
    L.vsk_refs_seed_hits.restype = C.c_int64
    L.vsk_reads_upload.argtypes = [vp, vp, C.POINTER(vp)]
    L.vsk_reads_free.argtypes = [vp, vp]
    L.vsk_reads_free.restype = None
    L.vsk_proclu_run.argtypes = [vp, vp, vp, C.POINTER(ProcluParams), C.POINTER(vp)]
    L.vsk_proclu_run_resident.argtypes = [vp, vp, vp, C.POINTER(ProcluParams), C.POINTER(vp)]
    L.vsk_proclu_result_links.argtypes = [vp, C.POINTER(vp)]
    L.vsk_proclu_result_links.restype = C.c_int64
    L.vsk_proclu_result_links_dev.argtypes = [vp, C.POINTER(vp)]
    L.vsk_proclu_result_links_dev.restype = C.c_int64
    L.vsk_proclu_result_stats.argtypes = [vp, C.POINTER(ProcluStats)]
    L.vsk_proclu_result_stats.restype = None
    L.vsk_proclu_result_free.argtypes = [vp]
    L.vsk_proclu_result_free.restype = None
    L.vsk_clusters_write.argtypes = [vp, C.c_int64, C.c_char_p, C.c_char_p, C.POINTER(C.c_uint64)]
    L.vsk_join_clusters_files.argtypes = [C.POINTER(C.c_char_p), C.c_int32, C.c_char_p, C.POINTER(JoinStats),
                                          C.c_char_p, C.c_size_t]
    L.vsk_rankflank_dir.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.POINTER(RankStats)]
    L.vsk_join_clusters_dir.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(JoinStats), C.c_int, C.c_char_p,
                                        C.c_size_t]
    L.vsk_timer_start.argtypes = [vp]
    L.vsk_timer_stop.argtypes = [vp, C.POINTER(C.c_float)]
    L.vsk_profile_enable.argtypes = [vp, C.c_int]
    L.vsk_profile_read.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    L.vsk_profile_read_sweep.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_int64), C.POINTER(C.c_double)]
    L.vsk_profile_stages.argtypes = [vp, C.c_char_p, C.c_size_t]
    L.vsk_int_peak.argtypes = [vp, C.POINTER(C.c_double)]
    L.vsk_pair_dp_static.argtypes = [vp, vp, C.c_int64, vp, C.c_int64, C.POINTER(DpDesc), C.c_int64, vp]
    L.vsk_band_params.argtypes = [C.c_int32, C.c_int32, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int32)]
    L.vsk_edges_run.argtypes = [vp, vp, vp, C.c_char_p, C.c_char_p, C.POINTER(C.c_int64)]


_declare_base = _declare


def _declare(L):  # noqa: F811
    _declare_base(L)
    _declare2(L)
