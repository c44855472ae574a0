"""CPU: the mean-field eigenmodel oracle (oracle/mf_eigen.py): vectorised forms against the frozen loop transcriptions
(tests/golden/mf_eigen_small.npz), Monte Carlo E[omega] against quadrature."""
import os

import numpy as np
import pytest

from oracle import mf_eigen as M

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "mf_eigen_small.npz"))


def test_expectations_and_omega():
    Em = M.expected_mu(float(G["mu0"]), G["mu"], G["lam"])
    V = M.variance_mu(float(G["mu0"]), float(G["s0"]), G["mu"], G["Sig"], G["lam"], G["SigL"])
    np.testing.assert_allclose(Em, G["E_mu"], rtol=1e-12)
    np.testing.assert_allclose(V, G["Var_mu"], rtol=1e-10)
    assert V.min() > 0
    np.testing.assert_allclose(M.expected_omega_mc(Em, V, G["z"]), G["E_O"], rtol=1e-12)
    # E[omega] = E[tanh(z/2) / (2z)]: the PG(1, z) mean identity the reference relies on (eigenmodel.py:1062-1063)
    rng = np.random.RandomState(0)
    mc = M.expected_omega_mc(Em, V, rng.randn(20000, *Em.shape))
    assert np.abs(mc - G["E_O_quad"]).max() < 2e-3
    np.testing.assert_allclose(M.expected_omega_quad(Em, V), G["E_O_quad"], rtol=1e-12)


@pytest.mark.parametrize("tr", [0, 1])
@pytest.mark.parametrize("tag,step", [("full", 1.0), ("step", 0.4)])
def test_update_F_vectorised_matches_loops(tr, tag, step):
    A = G["A"].astype(float)
    llT = M.expected_llT(G["lam"], G["SigL"])
    r = M.meanfieldupdate_F(A, G["E_O"], G["Fg"], G["mu"], G["Sig"], float(G["mu0"]), G["lam"], llT, 1.5, step, bool(tr))
    for nm, v in zip(("mu", "Sig", "prec", "h"), r):
        np.testing.assert_allclose(v, G["F_%s_%s_tr%d" % (nm, tag, tr)], rtol=1e-9, atol=1e-11)


def test_quirk_changes_the_result():
    assert np.abs(G["F_mu_full_tr0"] - G["F_mu_full_tr1"]).max() > 1e-3


def test_mu0_and_lmbda_updates():
    A = G["A"].astype(float)
    m0 = M.meanfieldupdate_mu0(A, G["E_O"], G["mu"], G["lam"], -1.2, 0.9, float(G["mu0"]), float(G["s0"]), 0.8)
    np.testing.assert_allclose([m0[0], m0[1], m0[2], m0[3]], [G["mu0_new"], G["s0_new"], G["mu0_prec"], G["mu0_h"]], rtol=1e-12)
    r = M.meanfieldupdate_lmbda(A, G["E_O"], G["mu"], G["Sig"], float(G["mu0"]), 0.0, 2.0, G["lam"], G["SigL"], 0.7)
    for v, k in zip(r, ("lam_new", "SigL_new", "lam_prec", "lam_h")):
        np.testing.assert_allclose(v, G[k], rtol=1e-9, atol=1e-11)


# ---- get_vlb (eigenmodel.py:1262-1298) -------------------------------------------------------------------------------
def test_vlb_matches_reference_distribution_classes():
    """golden values made with the reference's own ScalarGaussian / Gaussian classes (make_golden.py:make_mf_eigen)"""
    D = G["mu"].shape[1]
    args = (-1.2, 0.9, 1.5, np.zeros(D), 2.0, float(G["mu0"]), float(G["s0"]), G["mu"], G["Sig"], G["lam"], G["SigL"])
    assert np.isclose(M.get_vlb_loops(*args), float(G["vlb"]), rtol=1e-13)
    assert np.isclose(M.get_vlb_loops(*args, lmbda_given=True), float(G["vlb_lmbda_given"]), rtol=1e-13)


@pytest.mark.parametrize("given", [False, True])
def test_model_get_vlb_matches_oracle(given):
    """host class: no device work in get_vlb, as in the reference"""
    from graphistician_b200.eigenmodel import LogisticEigenmodel
    rng = np.random.RandomState(9)
    N, D = 11, 3
    model = LogisticEigenmodel(N, D, p=0.2, sigma_mu0=0.9, sigma_F=1.5, sigma_lmbda=2.0, lmbda=np.ones(D) if given else None)
    model.mf_mu_F = rng.randn(N, D)
    X = rng.randn(N, D, D) * 0.4
    model.mf_Sigma_F = X @ X.transpose(0, 2, 1) + 0.1 * np.eye(D)
    model.mf_mu_mu0, model.mf_sigma_mu0 = -0.4, 0.3
    if not given:
        Y = rng.randn(D, D) * 0.3
        model.mf_mu_lmbda, model.mf_Sigma_lmbda = rng.randn(D), Y @ Y.T + 0.05 * np.eye(D)
    mu_l, sig_l = (np.zeros(D), 1.0) if given else (np.asarray(model.mu_lmbda, float) * np.ones(D), model.sigma_lmbda)
    ref = M.get_vlb_loops(model.mu_mu_0, model.sigma_mu0, model.sigma_F, mu_l,
                          sig_l, model.mf_mu_mu0, model.mf_sigma_mu0, model.mf_mu_F, model.mf_Sigma_F,
                          model.mf_mu_lmbda, model.mf_Sigma_lmbda, lmbda_given=given)
    assert np.isclose(model.get_vlb(), ref, rtol=1e-11)
