"""CPU checks of the latent-distance oracle: the hand-derived gradient against finite differences of the
reference's literal log probability (adjacency.py:244-275), the golden fixture, and the row-block form."""
import os

import numpy as np

from oracle import latent


def _problem(N=23, dim=2, seed=0):
    rng = np.random.RandomState(seed)
    L = rng.randn(N, dim)
    A = (rng.rand(N, N) < 0.3).astype(float)
    return L, 0.3, -0.4, A, 2.0


def test_gradient_matches_finite_differences():
    L, mu0, mus, A, sigma = _problem()
    lp, gL, g0, gs = latent.log_probability_and_grad(L, mu0, mus, A, sigma)
    assert np.isclose(lp, latent.hmc_log_probability(L, mu0, mus, A, sigma), rtol=1e-12)
    h = 1e-6
    for (n, d) in [(0, 0), (5, 1), (22, 0)]:
        Lp, Lm = L.copy(), L.copy()
        Lp[n, d] += h; Lm[n, d] -= h
        fd = (latent.hmc_log_probability(Lp, mu0, mus, A, sigma) - latent.hmc_log_probability(Lm, mu0, mus, A, sigma)) / (2 * h)
        assert np.isclose(gL[n, d], fd, rtol=1e-6, atol=1e-7)
    fd0 = (latent.hmc_log_probability(L, mu0 + h, mus, A, sigma) - latent.hmc_log_probability(L, mu0 - h, mus, A, sigma)) / (2 * h)
    fds = (latent.hmc_log_probability(L, mu0, mus + h, A, sigma) - latent.hmc_log_probability(L, mu0, mus - h, A, sigma)) / (2 * h)
    assert np.isclose(g0, fd0, rtol=1e-6) and np.isclose(gs, fds, rtol=1e-6)


def test_row_blocks_sum_to_the_full_evaluation():
    L, mu0, mus, A, sigma = _problem(N=41, dim=3, seed=2)
    lp, gL, g0, gs = latent.log_probability_and_grad(L, mu0, mus, A, sigma)
    ll = 0.0; G = np.zeros_like(L); s0 = 0.0; sd = 0.0
    for r0, r1 in [(0, 7), (7, 30), (30, 41)]:
        a, b, c, d = latent.rows_value_and_grad(L, mu0, mus, A[r0:r1], A[:, r0:r1].T, r0, sigma)
        ll += a; G[r0:r1] = b; s0 += c; sd += d
    assert np.isclose(ll - 0.5 * np.sum(L * L) / sigma - 0.5 * mu0 ** 2 - 0.5 * mus ** 2, lp, rtol=1e-12)
    assert np.allclose(G, gL, rtol=1e-11, atol=1e-12)
    assert np.isclose(s0 - mu0, g0, rtol=1e-11) and np.isclose(sd - mus, gs, rtol=1e-11)


def test_golden_fixture(golden_dir):
    g = np.load(os.path.join(golden_dir, "latent_small.npz"))
    keys = set(g.files)
    assert {"L", "A"} <= keys
    lp, gL, g0, gs = latent.log_probability_and_grad(g["L"], float(g["mu_0"]), float(g["mu_self"]), g["A"], float(g["sigma"]))
    assert np.isclose(lp, float(g["lp"]), rtol=1e-12)
    assert np.allclose(gL, g["gL"], rtol=1e-11, atol=1e-12)
