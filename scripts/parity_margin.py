"""Prints the actual parity margins of the C1 / C2-shape 100-iteration runs against the fp64 oracle
(tolerance: 1e-4 on W and H, 1e-5 on the divergence) for a few engine settings."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import nmf_gpu_b200.api as api
from oracle import oracle
from util import max_rel, rand_problem

sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from make_golden import c1_inputs
cases = {"C1 4096x350 K=128": c1_inputs(), "513x2000 K=40": rand_problem(513, 2000, 40, seed=1),
         "600x1300 K=256": rand_problem(600, 1300, 256, seed=2)}
for name, (W, H, X) in cases.items():
    W = np.asfortranarray(W); H = np.asfortranarray(H); X = np.asfortranarray(X)
    Wo, Ho, _, co = oracle.update_div(W, H, X, 0.0, 100, "exact", acc64=True)
    Wf, Hf, _, cf = oracle.update_div(W, H, X, 0.0, 100, "exact", acc64=False)
    print(f"{name}: oracle fp32 vs fp64: W {max_rel(Wf, Wo):.2e} H {max_rel(Hf, Ho):.2e}")
    for chunk in (os.environ.get("MARGIN_CHUNKS", "1,4").split(",")):
        os.environ["NMF_UMMA_CHUNK"] = chunk
        Wd = W.copy(order="F"); Hd = H.copy(order="F")
        api.update_div(Wd, Hd, X, 0.0, 100)
        c = oracle.kl(Wd, Hd, X)
        print(f"   tcgen05 chunk {chunk}: W {max_rel(Wd, Wo):.2e} H {max_rel(Hd, Ho):.2e} KL rel {abs(c - co) / co:.2e}")
