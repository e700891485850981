"""The N>1 host logic on CPU: world_size 2 over gloo.  Each rank owns a column block of X and H,
computes its partial numerator Y*H' and partial row sums (numpy stands in for the GPU kernels), packs
them as the library does, allreduces ONE buffer, and applies the W update; the result must equal the
unsharded oracle half-steps.  Also checks the sharding rule itself."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
import numpy as np
import torch, torch.distributed as dist
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from nmf_gpu_b200 import sharding
from oracle import oracle
from util import rand_problem, max_rel
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
M, N, K = 96, 700, 24
W, H, X = rand_problem(M, N, K, seed=3)
lo, n = sharding.column_shard(N, rank, world, multiple=256)
Hl, Xl = H[:, lo:lo + n], X[:, lo:lo + n]
EPS = np.float32(2.2204e-16)
# H half-step is local to the shard (no communication)
Hl = oracle.h_step(W, Hl, Xl)
# W half-step: partials -> one packed allreduce -> identical update on every rank
Y = Xl / np.maximum(W @ Hl, EPS)
buf = torch.from_numpy(sharding.pack(Y @ Hl.T, Hl.sum(axis=1)))
assert buf.numel() == sharding.packed_size(M, K)
dist.all_reduce(buf)
P, s = sharding.unpack(buf.numpy(), M, K)
Wn = W * (P / np.maximum(s, EPS)[None, :])
# reference: the unsharded oracle
Hfull = oracle.h_step(W, H, X)
Wfull = oracle.w_step(W, Hfull, X)
assert max_rel(Hl, Hfull[:, lo:lo + n]) < 1e-6, "H shard mismatch"
assert max_rel(Wn, Wfull) < 2e-5, max_rel(Wn, Wfull)
# replicas stay in lock-step: every rank holds the same bits
chk = torch.tensor([float(np.abs(Wn).sum())], dtype=torch.float64)
lst = [torch.zeros_like(chk) for _ in range(world)]
dist.all_gather(lst, chk)
assert all(float(t) == float(lst[0]) for t in lst)
dist.destroy_process_group()
print("RANK_OK", rank)
'''


def test_column_shard_rule():
    from nmf_gpu_b200.sharding import column_shard
    for N in (1, 255, 256, 700, 262144, 1000000):
        for world in (1, 2, 4, 8):
            blocks = [column_shard(N, r, world) for r in range(world)]
            assert sum(n for _, n in blocks) == N
            pos = 0
            for lo, n in blocks:
                assert lo == pos or n == 0
                pos += n
            assert all(n % 256 == 0 for _, n in blocks[:-1] if n and _ + n < N)


def test_world_size_2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(f"ROOT = {ROOT!r}\n" + WORKER)
    env = dict(os.environ, OMP_NUM_THREADS="2")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert r.stdout.count("RANK_OK") == 2
