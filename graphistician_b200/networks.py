"""Factorised network distributions: adjacency model x weight model (mirrors ``graphistician.networks``)."""
import numpy as np

from .abstractions import NetworkDistribution, AdjacencyDistribution, WeightDistribution, GibbsSampling
from . import adjacency
from . import weights
from .old_abstractions import network_for


class FactorizedNetworkDistribution(NetworkDistribution, GibbsSampling):
    """networks.py:9-75."""

    def __init__(self, N, adjacency_class, adjacency_kwargs, weight_class, weight_kwargs):
        assert issubclass(adjacency_class, AdjacencyDistribution)
        self._adjacency = adjacency_class(N, **adjacency_kwargs)
        assert issubclass(weight_class, WeightDistribution)
        self._weights = weight_class(N, **weight_kwargs)
        self.N = N

    @property
    def adjacency(self):
        return self._adjacency

    @property
    def weights(self):
        return self._weights

    def log_likelihood(self, network):
        A, W = network
        network_for(A, W)                    # one device copy serves both factors
        return self.adjacency.log_likelihood(A) + self.weights.log_likelihood((A, W))

    def log_prior(self):
        return self.adjacency.log_prior() + self.weights.log_prior()

    def rvs(self, *args, **kwargs):
        return self.adjacency.rvs(), self.weights.rvs()

    def sample_predictive_parameters(self):
        """networks.py:42-48 (the reference asks the adjacency model twice, :47; the weight model is meant)."""
        return self.adjacency.sample_predictive_parameters(), self.weights.sample_predictive_parameters()

    def sample_predictive_distribution(self):
        Arow, Acol = self.adjacency.sample_predictive_distribution()
        Wrow, Wcol = self.weights.sample_predictive_distribution()
        return Arow, Acol, Wrow, Wcol

    def approx_predictive_ll(self, Arow, Acol, Wrow, Wcol, M=100):
        return self.adjacency.approx_predictive_ll(Arow, Acol, M=M) \
            + self.weights.approx_predictive_ll(Arow, Acol, Wrow, Wcol, M=M)

    def resample(self, network):
        """networks.py:69-71."""
        A, W = network
        network_for(A, W)                    # upload (A, W) once; both factors find it through the cache
        self.adjacency.resample(A)
        self.weights.resample((A, W))

    def meanfieldupdate(self, expected_suff_stats):
        pass                                 # networks.py:73-75 (TODO in the reference)


class GaussianBernoulliNetwork(FactorizedNetworkDistribution):
    """networks.py:78-88."""

    def __init__(self, N, B, p=0.5, mu=None, sigma=None):
        mu = np.zeros(B) if mu is None else mu
        sigma = np.eye(B) if sigma is None else sigma
        super(GaussianBernoulliNetwork, self).__init__(
            N, adjacency.BernoulliAdjacencyDistribution, {"p": p},
            weights.FixedGaussianWeightDistribution, {"B": B, "mu": mu, "sigma": sigma})
