"""Mean-field logistic eigenmodel on the GPU through the C ABI against oracle/mf_eigen.py.
Deterministic parts (F update, pair sums) with E_O injected: 1e-5 relative (fp64 storage) / 1e-3 (fp32).
The Monte Carlo E[omega] kernel: against Gauss-Hermite quadrature within Monte Carlo error."""
import os

import numpy as np
import pytest

from oracle import mf_eigen as M

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "mf_eigen_small.npz"))


def _problem(N, D, rng, p=0.3):
    A = (rng.rand(N, N) < p).astype(np.uint8)
    mu = rng.randn(N, D) * 0.6
    X = rng.randn(N, D, D) * 0.3
    Sg = X @ X.transpose(0, 2, 1) + 0.1 * np.eye(D)
    lam = rng.randn(D)
    Y = rng.randn(D, D) * 0.2
    SL = Y @ Y.T + 0.05 * np.eye(D)
    return A, mu, Sg, lam, SL, rng.randn(N, D)


def _dev(A, D, dtype):
    from graphistician_b200.old_abstractions import FixedGaussianNetwork
    from graphistician_b200.device import DeviceEigenmodel
    net = FixedGaussianNetwork(A, None).device(dtype)
    return DeviceEigenmodel(net, D)


@pytest.mark.parametrize("tr", [0, 1])
@pytest.mark.parametrize("tag,step", [("full", 1.0), ("step", 0.4)])
def test_update_F_matches_golden_loops(tr, tag, step):
    A, D = G["A"], G["mu"].shape[1]
    eig = _dev(A, D, "f64")
    S = G["E_O"] + G["E_O"].T
    np.fill_diagonal(S, 0.0)
    eig.set_S(S)
    llT = M.expected_llT(G["lam"], G["SigL"])
    mu, Sig, prec, h = eig.mf_update_F(G["mu"], G["Sig"], None if tr else G["Fg"], llT, G["lam"], float(G["mu0"]), 1.5, step, True)
    for nm, v in zip(("mu", "Sig", "prec", "h"), (mu, Sig, prec, h)):
        ref = G["F_%s_%s_tr%d" % (nm, tag, tr)]
        assert np.abs(v - ref).max() <= 1e-8 * max(1.0, np.abs(ref).max()), nm


@pytest.mark.parametrize("N,D,dtype,tr", [(200, 3, "f64", 0), (333, 8, "f64", 1), (300, 8, "f32", 0), (150, 1, "f64", 0), (260, 5, "f32", 1)])
def test_update_F_and_pair_stats_match_oracle(N, D, dtype, tr):
    rng = np.random.RandomState(N + D)
    A, mu, Sg, lam, SL, Fg = _problem(N, D, rng)
    mu0, s0 = -0.7, 0.1
    EO = M.expected_omega_quad(M.expected_mu(mu0, mu, lam), M.variance_mu(mu0, s0, mu, Sg, lam, SL), order=40)
    eig = _dev(A, D, dtype)
    S = EO + EO.T
    np.fill_diagonal(S, 0.0)
    eig.set_S(S)
    llT = M.expected_llT(lam, SL)
    Af = A.astype(float)
    tol = 1e-5 if dtype == "f64" else 1e-3
    for step in (1.0, 0.5):
        ref = M.meanfieldupdate_F(Af, EO, Fg, mu, Sg, mu0, lam, llT, 1.3, step, bool(tr))
        out = eig.mf_update_F(mu, Sg, None if tr else Fg, llT, lam, mu0, 1.3, step, True)
        for nm, v, r in zip(("mu", "Sig", "prec", "h"), out, ref):
            assert np.abs(v - r).max() <= tol * max(1.0, np.abs(r).max()), (nm, np.abs(v - r).max())
        mu, Sg = ref[0], ref[1]
    Pq, hS, hK = M.pair_stats(Af, EO, mu, Sg)
    st = eig.mf_pair_stats(mu, Sg)
    for v, r in ((st["Pq"], Pq), (st["hS"], hS), (st["hK"], hK)):
        assert np.abs(v - r).max() <= tol * max(1.0, np.abs(r).max())


@pytest.mark.parametrize("N,D,dtype", [(130, 3, "f64"), (200, 8, "f32")])
def test_omega_kernel_matches_quadrature(N, D, dtype):
    rng = np.random.RandomState(5)
    A, mu, Sg, lam, SL, _ = _problem(N, D, rng)
    mu0, s0 = -0.4, 0.3
    Em = M.expected_mu(mu0, mu, lam)
    V = M.variance_mu(mu0, s0, mu, Sg, lam, SL)
    EO = M.expected_omega_quad(Em, V)
    eig = _dev(A, D, dtype)
    ns = 4000
    sum_EO, sum_EOF = eig.mf_omega(mu, Sg, M.expected_llT(lam, SL), lam, mu0, s0, ns, 12345)
    S = eig.get_S()
    Sref = EO + EO.T
    np.fill_diagonal(Sref, 0.0)
    assert np.allclose(S, S.T) and np.all(np.diag(S) == 0)
    # per-entry Monte Carlo error: std of tanh(z/2)/(2z) is < 0.05 here; 2*ns samples per pair -> 6 sigma bound
    assert np.abs(S - Sref).max() < 6 * 2 * 0.05 / np.sqrt(2 * ns) + 1e-4
    assert abs(S.mean() - Sref.mean()) < 2e-4
    FdotF = Em - mu0
    np.fill_diagonal(FdotF, 0.0)
    assert abs(sum_EO - EO.sum()) < 2e-4 * N * N
    assert abs(sum_EOF - (EO * FdotF).sum()) < 5e-4 * np.abs(FdotF).sum()


def test_model_meanfieldupdate_matches_oracle_sequence():
    """LogisticEigenmodel.meanfieldupdate (:998-1022) with E_O taken from the device after the omega step."""
    from graphistician_b200.eigenmodel import LogisticEigenmodel
    from graphistician_b200.old_abstractions import FixedGaussianNetwork
    rng = np.random.RandomState(9)
    np.random.seed(4)
    N, D = 180, 4
    A = (rng.rand(N, N) < 0.25).astype(np.uint8)
    net = FixedGaussianNetwork(A, None)
    model = LogisticEigenmodel(N, D, p=0.25, sigma_F=1.0, dtype="f64")
    model.mf_n_samples = 200
    Af = A.astype(float)
    mu, Sg = model.mf_mu_F.copy(), model.mf_Sigma_F.copy()
    lam, SL = model.mf_mu_lmbda.copy(), model.mf_Sigma_lmbda.copy()
    mu0, s0 = model.mf_mu_mu0, model.mf_sigma_mu0
    Fg = model.F.copy()
    # one update on the device, replayed on the host from the device's own E_O + E_O^T
    sums = model._meanfieldupdate_Omega(net)
    S = model._device_state(net).get_S()
    EO = 0.5 * S                                        # any split with E_O + E_O^T = S gives the same F / lmbda updates
    model._meanfieldupdate_mu0(net, sums)
    diagEO = sums[0] - S.sum()                          # the diagonal entries enter the mu0 precision only
    post_prec = 1.0 / model.sigma_mu0 + S.sum() + diagEO
    FdotF = (mu * lam[None, :]).dot(mu.T); np.fill_diagonal(FdotF, 0)
    h0 = model.mu_mu_0 / model.sigma_mu0 + (Af - 0.5).sum() - (EO * FdotF).sum()
    assert abs(model.mf_sigma_mu0 - 1.0 / post_prec) < 1e-12
    assert abs(model.mf_mu_mu0 - h0 / post_prec) < 1e-5 * max(1.0, abs(h0 / post_prec))
    mu0 = model.mf_mu_mu0
    model._meanfieldupdate_F(net)
    ref = M.meanfieldupdate_F(Af, EO, Fg, mu, Sg, mu0, lam, M.expected_llT(lam, SL), model.sigma_F, 1.0, False)
    assert np.abs(model.mf_mu_F - ref[0]).max() < 1e-6 and np.abs(model.mf_Sigma_F - ref[1]).max() < 1e-6
    model._meanfieldupdate_lmbda(net)
    r = M.meanfieldupdate_lmbda(Af, EO, ref[0], ref[1], mu0, model.mu_lmbda, model.sigma_lmbda, lam, SL)
    assert np.abs(model.mf_mu_lmbda - r[0]).max() < 1e-6 and np.abs(model.mf_Sigma_lmbda - r[1]).max() < 1e-6
    model.meanfieldupdate(net)
    model.resample_from_mf()
    assert np.all(np.isfinite(model.F)) and np.isfinite(model.mu_0)
