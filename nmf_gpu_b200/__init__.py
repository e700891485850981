"""nmf_gpu_b200 — B200 (sm_100a) drop-in for the update_div hot path of recoord/nmf-gpu.

The product is libnmf_b200.so (csrc/: hand-written tcgen05/TMA kernels behind the C ABI of
include/nmf.h); this package is its Python host mirror.  The directory is spelled with an
underscore because "nmf-gpu_b200" is not an importable Python identifier.
"""
from .api import (EXACT, REFCOMPAT, ENGINE_SIMT, ENGINE_TCGEN05, NmfError, Session, matrix, nmf_options,
                  nmf_result, nccl_unique_id, options, read_matrix, update_div, write_matrix, lib, lib_path)

__all__ = ["EXACT", "REFCOMPAT", "ENGINE_SIMT", "ENGINE_TCGEN05", "NmfError", "Session", "matrix", "nmf_options",
           "nmf_result", "nccl_unique_id", "options", "read_matrix", "update_div", "write_matrix", "lib", "lib_path"]
