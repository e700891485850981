"""update_div on several GPUs from ONE process (NMF_NUM_GPUS=g: a host thread per GPU inside the library,
README-style call, no launcher): must reproduce the single-GPU result up to reduction order.
    python tests/multi_gpu_update_div.py [g [K [verbose]]]    (needs g visible GPUs; prints one PARITY line; K = 320 takes
the two-kernel path, K <= 256 the fused half-iterations; verbose = 1 exercises the cost collectives of a verbose run —
every rank must take part in them — with thresh = 0 and thresh > 0, and an iteration count that is not a multiple of 25)"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import nmf_gpu_b200.api as api
from util import max_rel, rand_problem


def main():
    g = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    M, N, K, iters = 1024, max(4096, 1024 * g) + 300, (int(sys.argv[2]) if len(sys.argv) > 2 else 96), 30
    verbose = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    W, H, X = rand_problem(M, N, K, seed=11)
    out = {}
    for mode in ("exact", "refcompat"):
        opt = api.options(semantics=mode)
        os.environ["NMF_NUM_GPUS"] = "1"
        W1 = W.copy(order="F"); H1 = H.copy(order="F")
        r1 = api.update_div(W1, H1, X, 1e-9, iters, verbose=verbose, opt=opt)   # same verbose: a verbose run closes with one more cost
        os.environ["NMF_NUM_GPUS"] = str(g)
        Wg = W.copy(order="F"); Hg = H.copy(order="F")
        t = np.zeros(10)
        rg = api.update_div(Wg, Hg, X, 1e-9, iters, t=t, verbose=verbose, opt=opt)
        out[mode] = dict(W=max_rel(Wg, W1), H=max_rel(Hg, H1), iters=(rg.iterations, r1.iterations),
                         cost_rel=abs(rg.cost - r1.cost) / abs(r1.cost), allreduce_s=float(t[7]))
        if verbose:                                               # thresh = 0 + verbose: the cost is evaluated for printing only
            Wv = W.copy(order="F"); Hv = H.copy(order="F")
            rv = api.update_div(Wv, Hv, X, 0.0, iters, verbose=1, opt=opt)
            out[mode]["verbose_thresh0"] = dict(W=max_rel(Wv, Wg), iters=rv.iterations)
    os.environ["NMF_NUM_GPUS"] = "1"
    print("PARITY " + json.dumps(dict(gpus=g, K=K, **out)), flush=True)
    ok = all(v["W"] < 2e-5 and v["H"] < 2e-5 and v["cost_rel"] < 1e-5 and v["iters"][0] == v["iters"][1]
             for k, v in out.items() if k == "exact") and out["refcompat"]["cost_rel"] < 1e-5
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
