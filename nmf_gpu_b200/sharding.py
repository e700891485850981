"""Host-side sharding rules of the multi-GPU path: X and H are split into contiguous
column blocks, W is replicated, and one packed fp32 buffer [numerator (M*K, column-major) | row sums
of H (K)] is sum-allreduced per iteration.  Pure Python: no device code here."""
import numpy as np


def column_shard(N, rank, world, multiple=256):
    """(first column, column count) of `rank`'s block.  Blocks are multiples of the GEMM tile width
    (so only the last rank sees a ragged tile) and may be empty for tiny N."""
    per = -(-N // world)
    per = -(-per // multiple) * multiple
    lo = min(N, rank * per)
    return lo, min(N, lo + per) - lo


def packed_size(M, K):
    return M * K + K


def pack(numerator, rowsum):
    """[P | s] as one float32 vector (P column-major), the allreduce payload."""
    return np.concatenate([np.asarray(numerator, np.float32).reshape(-1, order="F"),
                           np.asarray(rowsum, np.float32).reshape(-1)])


def unpack(buf, M, K):
    return buf[:M * K].reshape((M, K), order="F"), buf[M * K:M * K + K]
