"""duckhts_b200 — B200-native read_bam / bgunzip hot path of DuckHTS.

The product is native: `lib/libduckhts_cuda.so` (C-ABI in include/duckhts_cuda.h, sm_100a kernels) and
`lib/duckhts_b200.duckdb_extension.so` (DuckDB C-API table functions).  This Python package is a thin
ctypes mirror of the C-ABI used by tests and bench.py.  There is no CPU fallback: importing works
anywhere, but every compute entry point raises if the CUDA library or a GPU is missing.
"""
from ._lib import DuckhtsError, lib, lib_path  # noqa: F401
from .bgzf import BgzfDevice, Context, bgzf_deflate, bgzf_deflate_fd, bgzf_inflate, bgzf_walk  # noqa: F401
from .bam import BamFile, Index, bam_scan, batch_to_columns, STD_TAGS, VOFF_MAX  # noqa: F401,E402
from .bcf import BcfFile  # noqa: F401,E402
from .text import TextFile  # noqa: F401,E402
