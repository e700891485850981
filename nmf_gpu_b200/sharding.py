"""Host-side sharding rules of the multi-GPU path: X and H are split into contiguous
column blocks, W is replicated, and one packed fp32 buffer [numerator (M*K, column-major) | row sums
of H (K)] is sum-allreduced per iteration.  Pure Python: no device code here."""
import numpy as np


def column_shard(N, rank, world, multiple=256):
    """(first column, column count) of `rank`'s block.  The ceil(N / multiple) tile-wide column blocks
    are dealt out as evenly as possible (the first ranks get one more), so block boundaries fall on
    GEMM tile boundaries, only the last non-empty rank sees a ragged tile, and a shard is empty only
    when there are fewer blocks than ranks."""
    nblocks = -(-N // multiple)
    base, extra = divmod(nblocks, world)
    first = rank * base + min(rank, extra)
    count = base + (1 if rank < extra else 0)
    lo = min(N, first * multiple)
    return lo, min(N, (first + count) * multiple) - lo


def packed_size(M, K):
    return M * K + K


def pack(numerator, rowsum):
    """[P | s] as one float32 vector (P column-major), the allreduce payload."""
    return np.concatenate([np.asarray(numerator, np.float32).reshape(-1, order="F"),
                           np.asarray(rowsum, np.float32).reshape(-1)])


def unpack(buf, M, K):
    return buf[:M * K].reshape((M, K), order="F"), buf[M * K:M * K + K]
