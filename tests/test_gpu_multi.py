"""Multi-GPU correctness as `-m gpu` tests (skipped on a box with fewer than 2 GPUs): the column-sharded run against the
single-GPU run and the fp64 oracle under torchrun (one process per GPU), and update_div on 2 GPUs from one process
(NMF_NUM_GPUS: a host thread per GPU) including the verbose / thresh combinations whose cost collectives every rank
must join."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def gpu_count():
    try:
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=60).stdout
        return sum(1 for l in out.splitlines() if l.startswith("GPU "))
    except Exception:
        return 0


@pytest.fixture(scope="module")
def two_gpus():
    if gpu_count() < 2:
        pytest.skip("needs 2 GPUs")


def parity_line(out):
    lines = [l for l in out.splitlines() if l.startswith("PARITY ")]
    assert lines, out
    return json.loads(lines[-1][len("PARITY "):])


@pytest.mark.parametrize("fuse", ["1", "0"])
def test_torchrun_sharded_vs_single_gpu_and_oracle(two_gpus, fuse):
    env = dict(os.environ, NMF_UMMA_FUSE=fuse)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29631", os.path.join(ROOT, "tests", "dist_gpu_parity.py")],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    p = parity_line(r.stdout)
    assert p["world"] == 2 and p["exact"]["W_vs_oracle"] < 1e-4 and p["exact"]["H_vs_oracle"] < 1e-4


@pytest.mark.parametrize("K,verbose", [(128, 0), (128, 1), (320, 1), (512, 0)])
def test_update_div_num_gpus_2(two_gpus, K, verbose):
    """README-style call on 2 GPUs from one process; verbose = 1 used to leave rank 0 alone in the cost allreduce"""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "multi_gpu_update_div.py"), "2", str(K), str(verbose)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    p = parity_line(r.stdout)
    assert p["gpus"] == 2 and p["exact"]["iters"][0] == p["exact"]["iters"][1]


def test_more_gpus_than_column_blocks_is_clamped(two_gpus):
    """N = 256 is one 256-column block: NMF_NUM_GPUS = 2 used to leave rank 1 without columns and rank 0 waiting for it"""
    code = ("import os, sys, numpy as np; sys.path.insert(0, %r); sys.path.insert(0, %r);"
            "import nmf_gpu_b200.api as api; from util import rand_problem;"
            "W, H, X = rand_problem(300, 256, 64, 1); os.environ['NMF_NUM_GPUS'] = '2';"
            "r = api.update_div(W, H, X, 0.0, 5); print('DONE', r.iterations)") % (ROOT, os.path.join(ROOT, "tests"))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "DONE 5" in r.stdout, r.stdout + r.stderr
