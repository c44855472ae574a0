"""Old-API factorised weighted models: an adjacency model and an independent NIW Gaussian weight model.

* ``FactorizedWeightedNetworkDistribution`` -- the base class the reference imports from the missing
  ``graphistician.old_abstractions`` (eigenmodel.py:1319-1320); contract reconstructed from its use sites (SURVEY.md
  App. B; examples/demo_logistic_eigenmodel_gibbs.py:39,57-58, demo_logistic_eigenmodel_mf.py:45,58,62-68): every call
  is delegated to ``adjacency_dist`` and ``weight_dist``.
* ``GaussianWeightedEigenmodel`` -- eigenmodel.py:1319-1416: LogisticEigenmodel adjacency x GaussianWeights.
* ``GaussianErdosRenyiFixedSparsity`` -- the class examples/demo_gaussian_er_gibbs.py:7,27-41 and demo_gaussian_er_mf.py
  import (absent from the tree): Bernoulli(p) adjacency with FIXED p x GaussianWeights (nearest in-tree relative:
  networks.GaussianBernoulliNetwork, networks.py:78-88).

Every O(N^2) reduction runs in libgk (K1 statistics with K = 1, the eigenmodel kernels); the NIW / Gaussian draws are
O(B^2) host arithmetic on the global ``np.random`` stream, as in the reference."""
import numpy as np

from .old_abstractions import (NetworkDistribution, GaussianWeightedNetworkDistribution, FixedGaussianNetwork,
                               as_network)
from .internals.weights import GaussianWeights
from .eigenmodel import LogisticEigenmodel


def _edge_stats(network, dtype="f32"):
    """(n, sum_w, sum_ww) over ALL entries with A == 1, diagonal included (internals/weights.py:93-99)."""
    dev = network.device(dtype)
    zeros = np.zeros(network.N, np.int32)
    ne, s1, s2 = dev.block_stats(zeros, 1)
    d = dev.diag_stats(zeros, 1)
    return int(ne[0, 0] + d["d_edge"][0]), s1[0, 0] + d["d_sum_w"][0], s2[0, 0] + d["d_sum_ww"][0]


class FactorizedWeightedNetworkDistribution(NetworkDistribution):
    def __init__(self, N, B):
        self.N, self.B = N, B

    @property
    def adjacency_dist(self):
        raise NotImplementedError

    @property
    def weight_dist(self):
        raise NotImplementedError

    dtype = "f32"

    # ---- densities -------------------------------------------------------------------------------------
    def log_prior(self):
        return self.adjacency_dist.log_prior() + self.weight_dist.log_prior()

    def log_likelihood(self, network):
        """Bernoulli term on A (the adjacency model's own fused pass) + Gaussian term on W where A == 1, from the device
        statistics (successor formulas abstractions.py:111-115, 257-274)."""
        network = as_network(network)
        ll = self.adjacency_dist.log_likelihood(network)
        if self.B > 0 and network.W is not None:
            n, sw, sww = _edge_stats(network, self.dtype)
            if n:
                ll += self.weight_dist.log_likelihood_from_stats(n, sw, sww)
        return ll

    def rvs(self, size=[]):
        A = np.asarray(self.adjacency_dist.rvs())
        if isinstance(A, FixedGaussianNetwork):
            A = A.A
        W = self.weight_dist.rvs(size=(self.N, self.N)).reshape(self.N, self.N, self.B) if self.B > 0 else None
        return FixedGaussianNetwork(np.asarray(A, dtype=np.uint8), W)

    # ---- Gibbs -------------------------------------------------------------------------------------------
    def resample(self, network):
        network = as_network(network)
        self.adjacency_dist.resample(network)
        if self.B > 0 and network.W is not None:
            n, sw, sww = _edge_stats(network, self.dtype)
            self.weight_dist.resample_from_stats(n, sw, sww)

    # ---- mean field ----------------------------------------------------------------------------------------
    def meanfieldupdate(self, network):
        network = as_network(network)
        self.adjacency_dist.meanfieldupdate(network)
        if self.B > 0 and network.W is not None:
            n, sw, sww = _edge_stats(network, self.dtype)          # fixed network: E_A = A, E_W = W, E_WWT = W W^T
            self.weight_dist.meanfieldupdate_from_stats(n, sw, sww)

    def get_vlb(self):
        return self.adjacency_dist.get_vlb() + self.weight_dist.get_vlb()

    def expected_log_likelihood(self, network):
        network = as_network(network)
        ell = self.adjacency_dist.expected_log_likelihood(network)
        if self.B > 0 and network.W is not None:
            n, sw, sww = _edge_stats(network, self.dtype)
            c0, c1, C2 = self.weight_dist.expected_log_likelihood_coefficients()
            ell += n * c0 + c1.dot(sw) + np.sum(C2 * sww)
        return ell

    def resample_from_mf(self):
        self.adjacency_dist.resample_from_mf()
        self.weight_dist.resample_from_mf()


class GaussianWeightedEigenmodel(FactorizedWeightedNetworkDistribution, GaussianWeightedNetworkDistribution):
    """eigenmodel.py:1319-1416."""

    def __init__(self, N, B, D=2, p=0.5, sigma_mu0=1.0, sigma_F=1.0, lmbda=None, mu_lmbda=0, sigma_lmbda=1.0,
                 mu_0=None, Sigma_0=None, nu_0=None, kappa_0=None, dtype="f32"):
        super(GaussianWeightedEigenmodel, self).__init__(N, B)
        self.D = D
        self.dtype = dtype
        self._adjacency_dist = LogisticEigenmodel(N, D, p=p, sigma_mu0=sigma_mu0, sigma_F=sigma_F, lmbda=lmbda,
                                                  mu_lmbda=mu_lmbda, sigma_lmbda=sigma_lmbda, dtype=dtype)
        self._weight_dist = GaussianWeights(self.B, mu_0=mu_0, kappa_0=kappa_0, Sigma_0=Sigma_0, nu_0=nu_0)

    @property
    def adjacency_dist(self):
        return self._adjacency_dist

    @property
    def weight_dist(self):
        return self._weight_dist

    # latent variables of the eigenmodel (:1351-1391)
    @property
    def F(self):
        return self.adjacency_dist.F

    @property
    def P(self):
        return self.adjacency_dist.P

    @property
    def Mu(self):
        return np.tile(self.weight_dist.mu[None, None, :], [self.N, self.N, 1])

    @property
    def Sigma(self):
        return np.tile(self.weight_dist.sigma[None, None, :, :], [self.N, self.N, 1, 1])

    @property
    def mu_0(self):
        return self.adjacency_dist.mu_0

    @property
    def lmbda(self):
        return self.adjacency_dist.lmbda

    @property
    def mu_w(self):
        return self.weight_dist.mu

    @property
    def Sigma_w(self):
        return self.weight_dist.sigma

    # network-level expectations (:1393-1413)
    def mf_expected_log_p(self):
        return self.adjacency_dist.mf_expected_log_p()

    def mf_expected_log_notp(self):
        return self.adjacency_dist.mf_expected_log_notp()

    def mf_expected_mu(self):
        return np.tile(self.weight_dist.mf_expected_mu()[None, None, :], [self.N, self.N, 1])

    def mf_expected_mumuT(self):
        return np.tile(self.weight_dist.mf_expected_mumuT()[None, None, :, :], [self.N, self.N, 1, 1])

    def mf_expected_Sigma_inv(self):
        return np.tile(self.weight_dist.mf_expected_Sigma_inv()[None, None, :, :], [self.N, self.N, 1, 1])

    def mf_expected_logdet_Sigma(self):
        return self.weight_dist.mf_expected_logdet_Sigma() * np.ones((self.N, self.N))


class _FixedSparsityAdjacency(NetworkDistribution):
    """Bernoulli(p) edges with p FIXED: nothing to resample; the likelihood needs the edge count only (K1, K = 1)."""

    def __init__(self, N, p, dtype="f32"):
        assert 0.0 <= p <= 1.0, "p must be a probability"
        self.N, self.p, self.dtype = N, p, dtype

    @property
    def P(self):
        return self.p * np.ones((self.N, self.N))

    def log_prior(self):
        return 0.0

    def log_likelihood(self, network):
        network = as_network(network)
        dev = network.device(self.dtype)
        ne, _, _ = dev.block_stats(np.zeros(self.N, np.int32), 1)
        n1 = float(ne[0, 0]) + float(np.count_nonzero(dev.diagA))
        n0 = float(self.N) * self.N - n1
        with np.errstate(divide="ignore", invalid="ignore"):
            return float(np.nan_to_num(n1 * np.log(self.p)) + np.nan_to_num(n0 * np.log1p(-self.p)))

    def rvs(self, size=[]):
        return np.random.rand(self.N, self.N) < self.p

    def resample(self, network=None):
        pass

    def meanfieldupdate(self, network=None):
        pass

    def get_vlb(self):
        return 0.0

    def expected_log_likelihood(self, network):
        return self.log_likelihood(network)

    def resample_from_mf(self):
        pass


class GaussianErdosRenyiFixedSparsity(FactorizedWeightedNetworkDistribution, GaussianWeightedNetworkDistribution):
    """The model of examples/demo_gaussian_er_gibbs.py / demo_gaussian_er_mf.py (BASELINE config 1): N = 10, B = 2 there."""

    def __init__(self, N, B, p=0.5, mu_0=None, Sigma_0=None, nu_0=None, kappa_0=None, dtype="f32"):
        super(GaussianErdosRenyiFixedSparsity, self).__init__(N, B)
        self.dtype = dtype
        self._adjacency_dist = _FixedSparsityAdjacency(N, p, dtype)
        self._weight_dist = GaussianWeights(self.B, mu_0=mu_0, kappa_0=kappa_0, Sigma_0=Sigma_0, nu_0=nu_0)

    @property
    def adjacency_dist(self):
        return self._adjacency_dist

    @property
    def weight_dist(self):
        return self._weight_dist

    @property
    def p(self):
        return self._adjacency_dist.p

    @property
    def P(self):
        return self._adjacency_dist.P

    @property
    def Mu(self):
        return np.tile(self.weight_dist.mu[None, None, :], [self.N, self.N, 1])

    @property
    def Sigma(self):
        return np.tile(self.weight_dist.sigma[None, None, :, :], [self.N, self.N, 1, 1])
