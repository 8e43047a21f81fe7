"""vntrseek_b200 -- B200 (sm_100a) implementation of VNTRseek's alignment core.

The product is the C-ABI library `libvsk.so` (include/vsk.h) plus the drop-in
`psearch` / `flankalign` front-ends in vntrseek_b200/csrc/.  This Python package
is a thin ctypes mirror of that ABI used by the tests and bench.py.
"""
from .api import (Context, ProfileSet, RefIndex, ResidentReads, TRSetHandle, VskError, pack_profiles,  # noqa: F401
                  pack_strings, index_cache_build, index_cache_verify, join_clusters, proclu, rankflank,
                  write_clusters)
