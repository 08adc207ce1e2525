"""jn_cuclark_b200 -- B200-native cuCLARK / cuCLARK-l classification hot path.

The product is the CUDA shared library ``libcuclark_b200.so`` (C ABI declared in
``include/cuclark_b200.h``) and the ``bin/cuCLARK`` / ``bin/cuCLARK-l``
binaries built on it.  This package is a thin ctypes binding used by the
tests and ``bench.py``; it has NO CPU fallback: importing :mod:`.lib` raises if
the CUDA library has not been built (``python -c 'import __graft_entry__ as g;
g.build()'`` or ``make -C jn_cuclark_b200``).
"""
from . import host  # noqa: F401
from .lib import (  # noqa: F401
    CCB_EXTENDED, CCB_FORCE_DENSE, CCB_NO_ANCHOR, CCB_STATS, CCB_STATS_EX, CCB_WARP_KERNEL, CcbError, CuClarkDB, CuClarkGroup, Workload, lib_path, load_library,
)

__all__ = ["CuClarkDB", "CuClarkGroup", "Workload", "CcbError", "load_library", "lib_path",
           "CCB_EXTENDED", "CCB_STATS", "CCB_STATS_EX", "CCB_FORCE_DENSE", "CCB_NO_ANCHOR", "CCB_WARP_KERNEL"]
