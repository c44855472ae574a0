"""Mean-field PROBIT eigenmodel (eigenmodel.py:382-770) on the GPU through the C ABI, against
tests/golden/ref_mf_eigen.npz -- values produced by executing the reference's OWN method bodies
(tests/golden/make_ref_golden.py: _meanfieldupdate_Z / mf_expected_Z / _meanfieldupdate_mu0 / _F / _lmbda,
expected_log_likelihood, get_vlb).  fp conditionals: 1e-5 relative with fp64 device storage, 1e-3 with fp32."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_mf_eigen.npz"))


def _model(given, dtype):
    from graphistician_b200.eigenmodel import ProbitEigenmodel
    t = "probit_g%d_" % given
    A = G[t + "A"]
    N, D = G[t + "F"].shape
    np.random.seed(0)
    kw = dict(lmbda=G[t + "lmbda"].copy()) if given else dict(mu_lmbda=0.2, sigma_lmbda=1.7)
    m = ProbitEigenmodel(N, D, p=0.3, sigma_mu0=0.9, sigma_F=1.3, dtype=dtype, **kw)
    m.F = G[t + "F"].copy()
    m.mf_mu_F, m.mf_Sigma_F = G[t + "mf_mu_F"].copy(), G[t + "mf_Sigma_F"].copy()
    m.mf_mu_mu0, m.mf_sigma_mu0 = float(G[t + "mf_mu_mu0"]), float(G[t + "mf_sigma_mu0"])
    if given:
        m.mf_mu_lmbda = G[t + "lmbda"].copy(); m.mf_Sigma_lmbda = 1e-16 * np.eye(D)
    else:
        m.mf_mu_lmbda, m.mf_Sigma_lmbda = G[t + "mf_mu_lmbda"].copy(), G[t + "mf_Sigma_lmbda"].copy()
    return m, A, t


@pytest.mark.parametrize("given", [0, 1])
@pytest.mark.parametrize("dtype,tol", [("f64", 1e-5), ("f32", 1e-3)])
def test_probit_meanfield_matches_reference(given, dtype, tol):
    from graphistician_b200.old_abstractions import FixedGaussianNetwork
    m, A, t = _model(given, dtype)
    net = FixedGaussianNetwork(A, None)
    N, D = m.N, m.D
    # Z: mf_mu_Z and E[Z] (the device's E[Z] + E[Z]^T and its off-diagonal sum against the reference's N x N arrays)
    sum_EZ = m._meanfieldupdate_Z(net)
    np.testing.assert_allclose(m.mf_mu_Z, G[t + "mu_Z"], rtol=1e-12)
    EZ = G[t + "E_Z"].copy()
    np.testing.assert_allclose(m.mf_expected_Z(A), EZ, rtol=1e-10)              # host restatement == reference
    np.fill_diagonal(EZ, 0.0)
    assert abs(sum_EZ - EZ.sum()) <= tol * np.abs(EZ).sum()
    S = m._device_state(net).get_S()
    np.testing.assert_allclose(S, EZ + EZ.T, rtol=tol, atol=tol * 1e-2)
    # mu0
    m._meanfieldupdate_mu0(sum_EZ)
    np.testing.assert_allclose([m.mf_mu_mu0, m.mf_sigma_mu0], G[t + "full_mu0"], rtol=tol)
    # F (serial in n): conditionals, then the variational parameters
    prec, h = m._meanfieldupdate_F(net, want_conditionals=True)
    np.testing.assert_allclose(prec, G[t + "full_F_prec"], rtol=tol, atol=tol * np.abs(G[t + "full_F_prec"]).max() * 1e-2)
    np.testing.assert_allclose(m.mf_Sigma_F, G[t + "full_Sigma_F"], rtol=10 * tol, atol=tol * 1e-2)
    np.testing.assert_allclose(m.mf_mu_F, G[t + "full_mu_F"], rtol=10 * tol, atol=10 * tol * np.abs(G[t + "full_mu_F"]).max())
    if not given:
        m._meanfieldupdate_lmbda(net)
        np.testing.assert_allclose(m.mf_Sigma_lmbda, G[t + "full_Sigma_lmbda"], rtol=20 * tol, atol=tol * 1e-2)
        np.testing.assert_allclose(m.mf_mu_lmbda, G[t + "full_mu_lmbda"], rtol=20 * tol, atol=20 * tol * np.abs(G[t + "full_mu_lmbda"]).max())
    if dtype == "f64":
        np.testing.assert_allclose(m.expected_log_likelihood(net), float(G[t + "ell"]), rtol=1e-8)
        key = t + "full_vlb"
        if key in G.files:
            np.testing.assert_allclose(m.get_vlb(net), float(G[key]), rtol=1e-6)


def test_probit_meanfieldupdate_runs_end_to_end():
    """model.meanfieldupdate(network) (:412-438) == the pieces above in the reference's order; ProbitEigenmodel now carries
    both the Gibbs and the mean-field halves (eigenmodel.py:772)."""
    from graphistician_b200.old_abstractions import FixedGaussianNetwork
    m, A, t = _model(0, "f64")
    m2, _, _ = _model(0, "f64")
    net = FixedGaussianNetwork(A, None)
    m.meanfieldupdate(net)
    s = m2._meanfieldupdate_Z(net); m2._meanfieldupdate_mu0(s); m2._meanfieldupdate_F(net); m2._meanfieldupdate_lmbda(net)
    np.testing.assert_allclose(m.mf_mu_F, m2.mf_mu_F, rtol=1e-12)
    np.testing.assert_allclose(m.mf_mu_lmbda, m2.mf_mu_lmbda, rtol=1e-12)
    m.resample_from_mf()
    assert np.all(np.isfinite(m.F))
    m.resample(A)                       # the Gibbs half still works on the same object
    assert np.all(np.isfinite(m.F))
