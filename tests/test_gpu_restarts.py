"""Batched random restarts (the paper's "several random initialisations", nmf_ismir_2009.pdf p.2
section 3.2): R factor pairs that share X run in the same kernel launches, stacked along k.  Each restart must come out
as its own update_div call would (the restarts never interact), the per-restart costs must be right, and the returned
index must be the lowest-cost one."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from util import max_rel, rand_problem


@pytest.fixture(scope="module")
def api():
    import nmf_gpu_b200.api as a
    a.lib()
    return a


@pytest.fixture(scope="module")
def oracle():
    from oracle import oracle as o
    o.lib()
    return o


def inits(M, N, K, R, seed):
    rng = np.random.RandomState(seed)
    Ws = [np.asfortranarray(rng.rand(M, K).astype(np.float32)) for _ in range(R)]
    Hs = [np.asfortranarray(rng.rand(K, N).astype(np.float32)) for _ in range(R)]
    return Ws, Hs


@pytest.mark.parametrize("shape,R", [((513, 2000, 40), 4), ((300, 700, 64), 3), ((1100, 900, 128), 2), ((600, 1300, 256), 2),
                                     ((2100, 600, 96), 5)])
def test_batch_matches_independent_runs(api, oracle, shape, R):
    """thresh = 0: every restart of the stacked run against (a) the same restart run alone through update_div (only the
    order of the split reductions may differ) and (b) the fp64 oracle"""
    M, N, K = shape
    _, _, X = rand_problem(M, N, K, seed=K)
    Ws, Hs = inits(M, N, K, R, seed=R)
    Wb = [w.copy(order="F") for w in Ws]; Hb = [h.copy(order="F") for h in Hs]
    best, its, conv, cost = api.update_div_batch(Wb, Hb, X, 0.0, 20)
    assert its == [20] * R and conv == [0] * R
    costs_ref = []
    for r in range(R):
        W1, H1 = Ws[r].copy(order="F"), Hs[r].copy(order="F")
        api.update_div(W1, H1, X, 0.0, 20)
        assert max_rel(Wb[r], W1) < 1e-5 and max_rel(Hb[r], H1) < 1e-5, f"restart {r} differs from its own run"
        Wo, Ho, _, co = oracle.update_div(Ws[r], Hs[r], X, 0.0, 20, "exact", acc64=True)
        assert max_rel(Wb[r], Wo) < 1e-4 and max_rel(Hb[r], Ho) < 1e-4
        assert abs(cost[r] - co) / co < 1e-5                       # evaluated once at the end (keep-best needs it)
        costs_ref.append(co)
    assert best == int(np.argmin(costs_ref))


def test_batch_restarts_converge_independently(api, oracle):
    """thresh > 0 on a planted low-rank target: restarts meet the threshold at different checks; each one is returned
    with the factors, iteration count and cost its own update_div call stops with, while the others went on"""
    M, N, K, R = 300, 700, 64, 4
    rng = np.random.RandomState(5)
    # a noisy planted target: the divergence levels off, so the relative decrease does fall below the threshold
    # (oracle: restarts 0-2 stop after 250 iterations, restart 3 after 200)
    X = np.asfortranarray(((rng.rand(M, 8) @ rng.rand(8, N)) * (1 + 0.2 * (2 * rng.rand(M, N) - 1))).astype(np.float32))
    Ws, Hs = inits(M, N, K, R, seed=17)
    # make the restarts differ in how fast they converge: one starts from another's factors a few iterations further on
    W3, H3, _, _ = oracle.update_div(Ws[0], Hs[0], X, 0.0, 60, "exact")
    Ws[3], Hs[3] = np.asfortranarray(W3), np.asfortranarray(H3)
    thresh = 2e-2
    single = []
    for r in range(R):
        W1, H1 = Ws[r].copy(order="F"), Hs[r].copy(order="F")
        res = api.update_div(W1, H1, X, thresh, 400)
        single.append((W1, H1, res))
    assert len({s[2].iterations for s in single}) > 1, "the test wants restarts that stop at different checks"
    Wb = [w.copy(order="F") for w in Ws]; Hb = [h.copy(order="F") for h in Hs]
    best, its, conv, cost = api.update_div_batch(Wb, Hb, X, thresh, 400)
    for r in range(R):
        W1, H1, res = single[r]
        assert its[r] == res.iterations and conv[r] == res.converged
        assert abs(cost[r] - res.cost) / res.cost < 1e-5
        assert max_rel(Wb[r], W1) < 1e-4 and max_rel(Hb[r], H1) < 1e-4
    assert best == int(np.argmin([s[2].cost for s in single]))


def test_batch_session_api_and_fallback(api, oracle):
    """the session form (opt.restarts): device-generated restarts, per-restart costs; and the shapes the stacked kernels do
    not cover (K > 256 here) run the restarts one after another behind the same call"""
    M, N, K, R = 513, 4096, 40, 4
    with api.Session(M, N, K, opt=api.options(restarts=R)) as s:
        assert s.fused_steps() == 3
        s.fill_uniform(1)
        c0 = s.cost()
        assert len(c0) == R and len(set(c0)) == R                  # different initial factors
        res = s.run(0.0, 10)
        c1 = s.cost()
        assert all(b < a for a, b in zip(c0, c1))
        W0, H0 = s.download_restart(0)
    with api.Session(M, N, K) as s1:                                # restart 0 = what a single session of this seed does
        s1.fill_uniform(1)
        s1.run(0.0, 10)
        W1, H1 = s1.download()
    assert max_rel(W0, W1) < 1e-5 and max_rel(H0, H1) < 1e-5
    M, N, K, R = 200, 600, 320, 2
    _, _, X = rand_problem(M, N, K, seed=3)
    Ws, Hs = inits(M, N, K, R, seed=4)
    Wb = [w.copy(order="F") for w in Ws]; Hb = [h.copy(order="F") for h in Hs]
    best, its, conv, cost = api.update_div_batch(Wb, Hb, X, 0.0, 10)
    for r in range(R):
        Wo, Ho, _, co = oracle.update_div(Ws[r], Hs[r], X, 0.0, 10, "exact", acc64=True)
        assert max_rel(Wb[r], Wo) < 1e-4 and max_rel(Hb[r], Ho) < 1e-4 and abs(cost[r] - co) / co < 1e-5
