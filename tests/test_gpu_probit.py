"""GPU parity of the probit eigenmodel path (K7 truncated-normal draw + K4-K6 in probit mode) through the C ABI.

* the inverse-CDF map against the golden vector made by the REFERENCE'S OWN ``sample_truncnorm``
  (internals/utils.py:53-76, tests/golden/ref_internals.npz) -- this pins K7's arithmetic;
* with injected uniforms / normals the conditionals (prec_n, h_n, mu_0 and lmbda posteriors) against the oracle's
  loop transcription of eigenmodel.py:289-379: 1e-5 (fp64 storage) / 1e-3 (fp32);
* the device generator distributionally (truncation respected, moments vs the reference's expected_truncnorm)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_truncnorm_map_matches_reference_golden(golden_dir):
    from graphistician_b200.device import truncnorm_draw
    g = np.load(os.path.join(golden_dir, "ref_internals.npz"))
    z = truncnorm_draw(g["tn_mu"], g["tn_A"], g["tn_u"])
    assert np.allclose(z, g["tn_z"], rtol=1e-9, atol=1e-12)
    # edge cases the reference's clip decides: far tails on the "wrong" side, u at the ends
    mu = np.array([-9.0, 9.0, -40.0, 40.0, 0.0, 0.0, 3.0, -3.0])
    a = np.array([1, 0, 1, 0, 1, 0, 1, 0], dtype=np.uint8)
    u = np.array([0.5, 0.5, 0.999999, 1e-9, 1e-300, 1.0 - 1e-16, 0.3, 0.7])
    from oracle import eigenmodel as eg
    lb = np.where(a == 1, 0.0, -np.inf); ub = np.where(a == 1, np.inf, 0.0)
    ref = eg.sample_truncnorm(mu, lb, ub, u)
    out = truncnorm_draw(mu, a, u)
    assert np.all(np.isfinite(out))
    assert np.allclose(out, ref, rtol=1e-9, atol=1e-9)


def _problem(N, D, seed):
    from oracle import eigenmodel as eg
    rng = np.random.RandomState(seed)
    F = 0.6 * rng.randn(N, D)
    lm = rng.randn(D)
    mu0 = -0.4
    A = (rng.rand(N, N) < eg.probit_P(F, mu0, lm)).astype(np.uint8)
    return A, F, lm, mu0, rng.rand(N, N), rng.randn(N, D)


@pytest.mark.parametrize("N,D,dtype,tol", [(40, 3, "f64", 1e-8), (200, 2, "f64", 1e-8), (150, 8, "f32", 1e-3), (129, 1, "f64", 1e-8)])
def test_probit_sweep_matches_oracle(N, D, dtype, tol):
    from oracle import eigenmodel as eg
    from graphistician_b200.eigenmodel import ProbitEigenmodel
    A, F, lm, mu0, u, z = _problem(N, D, 5 + N)
    np.random.seed(1)
    model = ProbitEigenmodel(N, D, p=0.3, sigma_F=1.5, sigma_mu0=2.0, mu_lmbda=0.2, sigma_lmbda=1.7, dtype=dtype)
    model.F, model.mu_0, model.lmbda = F.copy(), mu0, lm.copy()
    np.random.seed(2)
    model.resample(A, uniforms=u, normals=z, want_conditionals=True)
    cond = model._last_conditionals
    # oracle: eigenmodel.py:268-379 with the same draws
    Z = eg.probit_resample_Z(A, F, mu0, lm, u)
    F_ref, precs, hs = eg.probit_resample_F_loops(Z, F, mu0, lm, 1.5, z)
    scale = lambda x: np.max(np.abs(x))
    assert np.max(np.abs(cond["prec"] - precs)) <= tol * scale(precs)
    assert np.max(np.abs(cond["h"] - hs)) <= tol * scale(hs)
    assert np.max(np.abs(model.F - F_ref)) <= 50 * tol * max(scale(F_ref), 1.0)
    pp, ph = eg.probit_mu0_conditional(Z, F_ref, lm, model.mu_mu_0, 2.0)
    assert abs(cond["mu0"][0] - pp) <= 1e-12 * pp
    assert abs(cond["mu0"][1] - ph) <= 50 * tol * max(abs(ph), np.abs(Z).sum() * 1e-3)
    # lmbda conditional (eigenmodel.py:349-379) with the NEW mu_0 the model drew
    S2, R2 = eg.probit_sym(Z, model.mu_0)
    np.fill_diagonal(S2, 0.0); np.fill_diagonal(R2, 0.0)
    lp, lh = eg.lmbda_conditional(S2, R2, F_ref, 0.2 * np.ones(D), 1.7)
    assert np.max(np.abs(cond["lmbda"][0] - lp)) <= 50 * tol * scale(lp)
    assert np.max(np.abs(cond["lmbda"][1] - lh)) <= 50 * tol * max(scale(lh), 1.0)


def test_device_truncnorm_generator_distribution(golden_dir):
    """Device-generated Z: truncation sides respected, Z + Z^T symmetric with zero diagonal, mean of Z matches
    E[truncnorm] (the reference's expected_truncnorm formula) within Monte Carlo error, draws change per sweep."""
    from scipy.stats import norm
    from graphistician_b200 import DeviceNetwork
    from graphistician_b200.device import DeviceEigenmodel
    N, D = 384, 3
    A, F, lm, mu0, _, _ = _problem(N, D, 99)
    net = DeviceNetwork(A, None, dtype="f64")
    eig = DeviceEigenmodel(net, D)
    eig.set_params(F, mu0, lm)
    s1 = eig.tn_draw(1234)
    Zs1 = eig.get_S()
    s2 = eig.tn_draw(1234)
    Zs2 = eig.get_S()
    assert np.allclose(Zs1, Zs1.T) and np.all(np.diag(Zs1) == 0)
    assert not np.allclose(Zs1, Zs2)                          # the sweep counter advances the stream
    Mu = (F * lm) @ F.T + mu0
    off = ~np.eye(N, dtype=bool)
    # E[z | a=1] = mu + phi(mu)/Phi(mu);  E[z | a=0] = mu - phi(mu)/Phi(-mu)
    Ez = np.where(A == 1, Mu + norm.pdf(Mu) / norm.cdf(Mu), Mu - norm.pdf(Mu) / norm.cdf(-Mu))
    expect_sum = Ez[off].sum()
    n = off.sum()
    assert abs(s1 - expect_sum) < 6.0 * np.sqrt(n)            # Var[z] <= 1 per entry
    assert abs(s2 - expect_sum) < 6.0 * np.sqrt(n)
    assert abs(Zs1[off].sum() - 2 * s1) <= 1e-8 * abs(s1) + 1e-6
    # both edges present -> both draws positive; both absent -> both negative
    both = (A == 1) & (A.T == 1) & off
    none = (A == 0) & (A.T == 0) & off
    assert np.all(Zs1[both] > 0) and np.all(Zs1[none] < 0)


def test_probit_model_runs_and_improves_fit():
    from graphistician_b200.eigenmodel import ProbitEigenmodel, Eigenmodel
    assert Eigenmodel is ProbitEigenmodel
    N, D = 256, 2
    np.random.seed(11)
    truth = ProbitEigenmodel(N, D, p=0.2, sigma_F=2.0)
    A = truth.rvs().astype(np.uint8)
    model = ProbitEigenmodel(N, D, p=0.2, sigma_F=2.0)
    ll0 = model.log_likelihood(A)
    for _ in range(30):
        model.resample(A)
    ll1 = model.log_likelihood(A)
    from oracle import eigenmodel as eg
    ref = eg.log_likelihood(A.astype(float), eg.probit_P(model.F, model.mu_0, model.lmbda))
    assert np.isclose(ll1, ref, rtol=1e-8)
    assert ll1 > ll0
