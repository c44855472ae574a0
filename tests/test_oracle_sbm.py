"""T1 (vectorised) oracle == T0 (literal loops) on fresh random cases; NIW / discrete helpers."""
import numpy as np
import pytest
from scipy.stats import multivariate_normal

from oracle import niw, pg, sbm


def _problem(seed, N=13, K=3, B=2):
    rng = np.random.RandomState(seed)
    c = rng.randint(0, K, N); m = rng.dirichlet(np.ones(K)); p = rng.beta(1, 1, (K, K))
    mus = rng.randn(K, K, B); sig = np.zeros((K, K, B, B))
    for a in range(K):
        for b in range(K):
            X = rng.randn(B, B); sig[a, b] = X @ X.T + 0.5 * np.eye(B)
    A = (rng.rand(N, N) < 0.4).astype(float); W = rng.randn(N, N, B)
    return A, W, c, m, p, mus, sig, rng.rand(N)


@pytest.mark.parametrize("seed,B", [(0, 1), (1, 2), (2, 3)])
def test_sweep_vectorised_equals_loops(seed, B):
    A, W, c, m, p, mus, sig, u = _problem(seed, B=B)
    c0, l0 = sbm.gibbs_sbm_resample_c_loops(A, W, c, m, p, mus, sig, u)
    c1, l1 = sbm.sbm_assign_sweep(A, W, c, m, p, mus, sig, u, sbm.SELF_OLD)
    assert np.array_equal(c0, c1) and np.allclose(l0, l1, rtol=1e-11, atol=1e-10)


def test_stats_vectorised_equals_masks():
    A, W, c, m, p, mus, sig, u = _problem(5, N=17, K=4)
    s0 = sbm.block_pair_stats_loops(A, W, c, 4); s1 = sbm.block_pair_stats(A, W, c, 4, row_chunk=5)
    for k in s0:
        assert np.allclose(s0[k], s1[k], atol=1e-12), k


def test_empty_blocks_and_K1():
    A, W, c, m, p, mus, sig, u = _problem(6, K=3)
    c[:] = 1                                             # blocks 0 and 2 empty
    s1 = sbm.block_pair_stats(A, W, c, 3)
    assert s1["n_pair"][1, 1] == 13 * 12 and s1["n_pair"].sum() == 13 * 12
    c2, lp = sbm.sbm_assign_sweep(A, W, c, np.ones(1), p[:1, :1], mus[:1, :1], sig[:1, :1], u)
    assert np.array_equal(c2, c)                          # C == 1 -> no-op (adjacency.py:657-658)


def test_sample_discrete_from_log():
    lp = np.log(np.array([0.1, 0.2, 0.3, 0.4]))
    assert [niw.sample_discrete_from_log(lp, u) for u in (0.05, 0.1, 0.29, 0.31, 0.999)] == [0, 0, 1, 2, 3]


def test_gaussian_loglik_and_stats_form():
    rng = np.random.RandomState(3)
    X = rng.randn(50, 3); mu = rng.randn(3); S = np.cov(rng.randn(3, 20))
    ll = niw.gaussian_log_likelihood(X, mu, S)
    assert np.allclose(ll, multivariate_normal(mu, S).logpdf(X))
    assert np.allclose(ll.sum(), niw.gaussian_loglik_from_stats(50, X.sum(0), X.T @ X, mu, S))


def test_niw_posterior_moments():
    rng = np.random.RandomState(4)
    X = rng.randn(400, 2) * [1.0, 2.0] + [0.5, -1.0]
    draws = [niw.niw_resample(np.zeros(2), np.eye(2), 1.0, 4, data=X, rng=rng) for _ in range(300)]
    mu = np.mean([d[0] for d in draws], 0); Sg = np.mean([d[1] for d in draws], 0)
    assert np.allclose(mu, X.mean(0), atol=0.05) and np.allclose(np.diag(Sg), X.var(0), rtol=0.1)


def test_pg_moments():
    for z in (0.0, 1.5, 6.0):
        x = pg.pgdraw(np.full(400000, z), seed=11)
        assert abs(x.mean() - pg.pg1_mean(z)) < 5 * np.sqrt(pg.pg1_var(z) / x.size)
        assert abs(x.var() / pg.pg1_var(z) - 1) < 0.03
