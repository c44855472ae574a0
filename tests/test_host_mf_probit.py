"""Host arithmetic of the mean-field probit eigenmodel class (graphistician_b200/eigenmodel.py) against values produced by
the reference's own code (tests/golden/ref_mf_eigen.npz): mf_mu_Z, mf_expected_Z (with the reference's literal
expected_truncnorm normaliser, internals/utils.py:93), expected_log_likelihood and get_vlb (:688-756).  No device call."""
import os

import numpy as np
import pytest

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_mf_eigen.npz"))


@pytest.mark.parametrize("given", [0, 1])
def test_probit_mf_host_terms_match_reference(given):
    from graphistician_b200.eigenmodel import ProbitEigenmodel
    from graphistician_b200.old_abstractions import FixedGaussianNetwork
    t = "probit_g%d_" % given
    A = G[t + "A"]
    N, D = G[t + "F"].shape
    np.random.seed(0)
    kw = dict(lmbda=G[t + "lmbda"].copy()) if given else dict(mu_lmbda=0.2, sigma_lmbda=1.7)
    m = ProbitEigenmodel(N, D, p=0.3, sigma_mu0=0.9, sigma_F=1.3, dtype="f64", **kw)
    assert np.array_equal(m.mf_mu_Z, np.zeros((N, N)))                      # :388 before the first update
    lam0 = G[t + "lmbda"] if given else G[t + "mf_mu_lmbda"]
    m._mf_Z_snapshot = (float(G[t + "mf_mu_mu0"]), G[t + "mf_mu_F"].copy(), lam0.copy())
    np.testing.assert_allclose(m.mf_mu_Z, G[t + "mu_Z"], rtol=1e-13)
    np.testing.assert_allclose(m.mf_expected_Z(A), G[t + "E_Z"], rtol=1e-12)
    m.literal_truncnorm_normalizer = False                                   # the truncated-normal mean differs from the quirk
    assert np.abs(m.mf_expected_Z(A) - G[t + "E_Z"]).max() > 1e-3
    m.literal_truncnorm_normalizer = True
    m.mf_mu_mu0, m.mf_sigma_mu0 = G[t + "full_mu0"]
    m.mf_mu_F, m.mf_Sigma_F = G[t + "full_mu_F"].copy(), G[t + "full_Sigma_F"].copy()
    if given:
        m.mf_mu_lmbda = G[t + "lmbda"].copy(); m.mf_Sigma_lmbda = 1e-16 * np.eye(D)
    else:
        m.mf_mu_lmbda, m.mf_Sigma_lmbda = G[t + "full_mu_lmbda"].copy(), G[t + "full_Sigma_lmbda"].copy()
    net = FixedGaussianNetwork(A, None)
    np.testing.assert_allclose(m.get_vlb(net), float(G[t + "full_vlb"]), rtol=1e-12)
    np.testing.assert_allclose(m.expected_log_likelihood(net), float(G[t + "ell"]), rtol=1e-12)


def test_network_cache_is_bounded():
    """network_for must not pin every array it ever saw (Geweke / AIS loops draw a fresh A per iteration)."""
    from graphistician_b200 import old_abstractions as oa
    oa.drop_cached_networks()
    first = None
    for i in range(20):
        A = (np.random.rand(8, 8) < 0.3).astype(np.uint8)
        net = oa.network_for(A)
        assert oa.network_for(A) is net                                     # identity cache hit
        first = first or net
    assert len(oa._NETWORKS) <= oa.MAX_CACHED_NETWORKS
    assert all(n is not first for n in oa._NETWORKS.values())                # the oldest entry was evicted
    net._sum_A = 3.0
    net.invalidate()
    assert not hasattr(net, "_sum_A")                                       # host-side caches of A go stale with it
