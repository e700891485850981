"""nmf_rowsum_noalloc (include/nmf.h): the row sums of H through the entry point that allocates nothing — its partial sums
live in a device array built into the library — against the existing entry point for the same sums (nmf_dbg_sums: a
session with its own cudaMalloc'ed scratch), on calls in a row that share the array.  Tolerance as test_sums
(tests/test_gpu_parity.py): rtol 1e-5.  Collected last on purpose (the file name).
The device buffers of the TEST come from torch; the entry point itself gets raw device addresses."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from util import rand_problem


@pytest.fixture(scope="module")
def api():
    import nmf_gpu_b200.api as a
    a.lib()
    return a


def run_noalloc(api, torch, H):
    K, N = H.shape
    Hd = torch.from_numpy(np.ascontiguousarray(H.T)).cuda()          # (N, K) row-major == K x N column-major, ld = K
    out = torch.full((K,), float("nan"), dtype=torch.float32, device="cuda")
    api.rowsum_noalloc(Hd.data_ptr(), K, N, out.data_ptr())            # default stream; only enqueues
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_rowsum_noalloc_matches_session_sums_on_calls_in_a_row(api):
    torch = pytest.importorskip("torch")
    # a large call first, then smaller ones: the shared array still holds the first call's partial sums beyond what the
    # later calls use (and, for the different K, in a different layout)
    for M, N, K, seed in [(200, 3000, 64, 5), (200, 700, 64, 6), (200, 1500, 96, 7), (200, 3000, 64, 5)]:
        W, H, _ = rand_problem(M, N, K, seed=seed)
        _, ref = api.dbg_sums(W, H, api.options())                     # existing entry point (session scratch)
        got = run_noalloc(api, torch, H)
        assert np.isfinite(got).all()
        assert np.allclose(got, ref, rtol=1e-5), (M, N, K)
        assert np.allclose(got, H.astype(np.float64).sum(1), rtol=1e-5)


def test_rowsum_noalloc_rejects_sizes_beyond_its_scratch(api):
    torch = pytest.importorskip("torch")
    buf = torch.zeros(1024, dtype=torch.float32, device="cuda")
    with pytest.raises(api.NmfError, match="exceed the built-in scratch"):
        api.rowsum_noalloc(buf.data_ptr(), 1024, 600000, buf.data_ptr())     # ceil(600000/512) * 1024 > 2^20: nothing is launched
    with pytest.raises(api.NmfError, match="multiple of 4"):
        api.rowsum_noalloc(buf.data_ptr(), 6, 10, buf.data_ptr())
