"""Multi-GPU parity (run under torchrun on N GPUs): the column-sharded run must reproduce the
single-GPU run of the same problem (differences come only from reduction order) AND the fp64 CPU oracle
(exact semantics, the north-star tolerance for this iteration count).
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/dist_gpu_parity.py
Prints one PARITY line on rank 0 and exits non-zero on failure."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import nmf_gpu_b200.api as api
from nmf_gpu_b200.sharding import column_shard
from util import max_rel, rand_problem


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    # every rank owns columns (shards are 256-column blocks); K a multiple of 64 so that the fused half-iteration
    # kernels run on every rank (NMF_UMMA_FUSE=0 takes the two-kernel path instead)
    M, N, K, iters = 1024, max(4096, 1024 * world) + 300, 128, 30
    W, H, X = rand_problem(M, N, K, seed=11)
    out = {}
    for mode in ("exact", "refcompat"):
        opt = api.options(semantics=mode, device=local)
        lo, n = column_shard(N, rank, world)
        s = api.Session(M, n, K, n_offset=lo, n_global=N, opt=opt)
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid = torch.frombuffer(bytearray(api.nccl_unique_id()), dtype=torch.uint8).cuda()
        dist.broadcast(uid, 0)
        s.join(rank, world, uid.cpu().numpy().tobytes())
        s.upload(W, H[:, lo:lo + n], X[:, lo:lo + n])
        fused = s.fused_steps()
        res = s.run(1e-9, iters)                      # thresh > 0 exercises the allreduced cost too
        Wd, Hd = s.download()
        cost = s.cost()
        # a second session of this process on the shared communicator reproduces the run bit for bit
        s2 = api.Session(M, n, K, n_offset=lo, n_global=N, opt=opt)
        s2.share_comm(s)
        s2.upload(W, H[:, lo:lo + n], X[:, lo:lo + n]); s2.run(1e-9, iters)
        W2, H2 = s2.download(); s2.close()
        assert np.array_equal(W2, Wd) and np.array_equal(H2, Hd), "shared-communicator session differs"
        s.close()
        # gather the H shards on every rank
        Hfull = torch.zeros((N, K), dtype=torch.float32, device="cuda")
        Hfull[lo:lo + n] = torch.from_numpy(np.ascontiguousarray(Hd.T)).cuda()
        dist.all_reduce(Hfull)
        Hm = Hfull.cpu().numpy().T
        if rank == 0:
            with api.Session(M, N, K, opt=opt) as s1:             # the same problem on one GPU
                s1.upload(W, H, X)
                r1 = s1.run(1e-9, iters)
                W1, H1 = s1.download()
                cost1 = s1.cost()
            out[mode] = dict(W=max_rel(Wd, W1), H=max_rel(Hm, H1), cost_rel=abs(cost - cost1) / abs(cost1),
                             iters=(res.iterations, r1.iterations), fused_steps=fused)
            if mode == "exact":                                   # and against the oracle, not only against ourselves
                from oracle import oracle
                Wo, Ho, _, costo = oracle.update_div(W, H, X, 0.0, iters, "exact", acc64=True)
                out[mode].update(W_vs_oracle=max_rel(Wd, Wo), H_vs_oracle=max_rel(Hm, Ho),
                                 cost_vs_oracle=abs(cost - costo) / abs(costo))
    ok = True
    if rank == 0:
        print("PARITY " + json.dumps(dict(world=world, **out)), flush=True)
        ok = all(v["W"] < 2e-5 and v["H"] < 2e-5 and v["cost_rel"] < 1e-5 and v["iters"][0] == v["iters"][1]
                 and v["W_vs_oracle"] < 1e-4 and v["H_vs_oracle"] < 1e-4 and v["cost_vs_oracle"] < 1e-5
                 for k, v in out.items() if k == "exact")
        ok &= out["refcompat"]["cost_rel"] < 1e-5
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
