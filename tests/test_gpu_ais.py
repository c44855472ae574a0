"""GPU parity of the AIS path (a17: internals/graph_models.py + the driver of examples/demo_ais.py) through the C ABI.

The tempered feature sweep (beta-weighted likelihood, diagonal counted in both directions, hips log_sum_exp_sample)
must give bit-identical labels to the oracle's loop transcription for the same uniforms; the tempered Beta/Dirichlet
posteriors and the whole AIS log weight must equal the reference arithmetic driven by the same np.random stream."""
import numpy as np
import pytest
from scipy.special import betaln

from oracle import sbm

pytestmark = pytest.mark.gpu


def _sbm_data(N, R, seed):
    from graphistician_b200.internals.graph_models import StochasticBlockModel, sample_network
    np.random.seed(seed)
    model = StochasticBlockModel(R=R, b0=1.0, b1=1.0, alpha0=10 * np.ones(R))
    A, f, theta = sample_network(model, N)
    return model, A.astype(np.uint8), np.array(f), theta


@pytest.mark.parametrize("N,R,beta", [(40, 3, 1.0), (130, 3, 0.35), (257, 5, 0.0101), (64, 2, 0.0)])
def test_tempered_feature_sweep_matches_reference_loop(N, R, beta):
    model, A, f, theta = _sbm_data(N, R, 3 + N)
    rng = np.random.RandomState(5)
    z0 = rng.randint(0, R, N)
    u = rng.rand(N)
    z = model.sample_fs(theta, A, z0, beta=beta, uniforms=u)
    z_ref, _ = sbm.ais_naive_sample_f_loops(A.astype(float), z0, theta[0], theta[1], beta, u)
    assert np.array_equal(z, z_ref)


def _ref_sample_theta(A, z, R, b0, b1, alpha0, beta):
    """internals/graph_models.py:337-371 literally (np.ix_ gathers, row-major Beta draws)."""
    alpha_post = np.array([alpha0[r] + np.sum(z == r) for r in range(R)], dtype=float)
    pi = np.random.dirichlet(alpha_post)
    B = np.zeros((R, R), dtype=np.float32)
    for r1 in range(R):
        for r2 in range(R):
            b0p, b1p = b0, b1
            Ar = A[np.ix_(z == r1, z == r2)]
            if np.size(Ar) > 0:
                b0p += beta * np.sum(1 - Ar)
                b1p += beta * np.sum(Ar)
            B[r1, r2] = np.random.beta(b1p, b0p)
    return B.astype(np.float64), pi


@pytest.mark.parametrize("beta", [1.0, 0.2])
def test_tempered_theta_and_likelihood_match_reference(beta):
    N, R = 150, 4
    model, A, f, theta = _sbm_data(N, R, 11)
    Af = A.astype(float)
    np.random.seed(42)
    B, pi = model.sample_theta((A, f), beta=beta)
    np.random.seed(42)
    B_ref, pi_ref = _ref_sample_theta(Af, f, R, 1.0, 1.0, 10 * np.ones(R), beta)
    assert np.array_equal(pi, pi_ref)
    assert np.allclose(B, B_ref, rtol=1e-12)
    P = B[np.ix_(f, f)]
    ll_ref = np.sum(Af * np.log(P) + (1 - Af) * np.log(1 - P))
    assert np.isclose(model.log_lkhd(A, f, (B, pi)), ll_ref, rtol=1e-11)
    lp_ref = np.sum(-betaln(1.0, 1.0) * np.ones((R, R))) + R * np.sum(9.0 * np.log(pi)) + np.sum(np.log(pi[f])) + beta * ll_ref
    assert np.isclose(model.logpr(A, f, (B, pi), beta=beta), lp_ref, rtol=1e-11)


def test_ais_log_weights_match_reference_driver():
    """examples/demo_ais.py:10-62 re-run on the CPU with the oracle's loops and the same np.random stream."""
    from graphistician_b200.internals.graph_models import estimate_marginal_likelihood
    N, R, NB, steps, M = 48, 3, 7, 2, 2
    model, A, f0, theta0 = _sbm_data(N, R, 21)
    Af = A.astype(float)
    np.random.seed(77)
    logZ, w = estimate_marginal_likelihood(A, f0, theta0, model, N_samples=M, B=NB, steps_per_B=steps, return_weights=True)
    np.random.seed(77)
    betas = np.linspace(0, 1, NB)
    w_ref = np.zeros(M)
    for m in range(M):
        f, theta = f0.copy(), theta0
        for b in range(1, NB):
            for s in range(steps):
                theta = _ref_sample_theta(Af, f, R, 1.0, 1.0, 10 * np.ones(R), betas[b])
                u = np.random.rand(N)
                f, _ = sbm.ais_naive_sample_f_loops(Af, f, theta[0], theta[1], betas[b], u)
            P = theta[0][np.ix_(f, f)]
            ll = np.sum(Af * np.log(P) + (1 - Af) * np.log(1 - P))
            w_ref[m] += (betas[b] - betas[b - 1]) * ll
    assert np.allclose(w, w_ref, rtol=1e-10)


def test_erdos_renyi_ais_recovers_the_exact_marginal_likelihood():
    from graphistician_b200.internals.graph_models import ErdosRenyiNetwork, estimate_marginal_likelihood, sample_network
    N = 64
    np.random.seed(4)
    model = ErdosRenyiNetwork(x=(1.0, 1.0))
    A, f0, theta0 = sample_network(ErdosRenyiNetwork(rho=0.2), N)
    A = A.astype(np.uint8)
    nnz = float(A.sum())
    exact = betaln(1.0 + nnz, 1.0 + N * N - nnz) - betaln(1.0, 1.0)
    _, f0, theta0 = sample_network(model, N)
    # The reference's ladder scores each sample under the temperature it was just drawn at (demo_ais.py:44-52), which
    # over-estimates log Z by O(1/B); the estimate must close in on the exact value as the ladder gets finer.
    est = estimate_marginal_likelihood(A, f0, theta0, model, N_samples=40, B=200, steps_per_B=1)
    fine = estimate_marginal_likelihood(A, f0, theta0, model, N_samples=40, B=2000, steps_per_B=1)
    assert -0.5 < est - exact < 3.0, (est, exact)
    assert abs(fine - exact) < 1.0 and abs(fine - exact) < abs(est - exact), (fine, est, exact)


def test_sbm_ais_prefers_the_block_model_on_block_data():
    from graphistician_b200.internals.graph_models import (ErdosRenyiNetwork, StochasticBlockModel, sample_network,
                                                           estimate_marginal_likelihood, invariant_sbm_order)
    N, R = 96, 3
    np.random.seed(8)
    sbm_model = StochasticBlockModel(R=R, b0=1.0, b1=1.0, alpha0=10 * np.ones(R))
    theta = (np.array([[0.8, 0.05, 0.05], [0.05, 0.7, 0.1], [0.1, 0.05, 0.9]]), np.ones(R) / R)
    f = [sbm_model.sample_f(theta) for _ in range(N)]
    A = sbm_model.sample_A(f, theta).astype(np.uint8)
    order = invariant_sbm_order(f, R)
    assert sorted(order) == list(range(N))
    er = ErdosRenyiNetwork(x=(1.0, 1.0))
    _, f0, t0 = sample_network(er, N)
    z_er = estimate_marginal_likelihood(A, f0, t0, er, N_samples=4, B=60, steps_per_B=1)
    _, f0, t0 = sample_network(sbm_model, N)
    z_sbm = estimate_marginal_likelihood(A, np.array(f0), t0, sbm_model, N_samples=4, B=60, steps_per_B=2)
    assert np.isfinite(z_er) and np.isfinite(z_sbm)
    assert z_sbm > z_er + 50.0


def test_collapsed_sample_f_matches_reference_loop():
    """internals/graph_models.py:299-335 (B integrated out): the count-table form against the literal np.ix_ loops."""
    N, R = 60, 3
    model, A, f, theta = _sbm_data(N, R, 31)
    rng = np.random.RandomState(2)
    z = rng.randint(0, R, N)
    for n in (0, 7, N - 1):
        u = rng.rand()
        zn, lp = model.collapsed_sample_f(theta, n, A, z, uniform=u, return_lp=True)
        lp_ref = sbm.ais_collapsed_ln_post_loops(A, z, n, theta[1], model.b0, model.b1)
        assert np.allclose(lp, lp_ref, rtol=1e-10, atol=1e-9)
        assert zn == sbm.log_sum_exp_sample(lp_ref, u)
        z[n] = zn
