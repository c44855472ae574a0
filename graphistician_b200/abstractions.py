"""New-API base classes: adjacency and weight distributions over (A, W) tuples.

Mirrors ``graphistician.abstractions`` (abstractions.py:9-357).  Distributions hold parameters, never data.
The O(N^2) reductions (log likelihoods) of the concrete classes run in libgk; what stays here is the O(N)
predictive machinery and the API contract.
"""
import numpy as np
from scipy.special import logsumexp


class Distribution(object):
    """pybasicbayes.abstractions.Distribution (not in the reference tree): rvs / log_likelihood contract."""

    def rvs(self, size=[]):
        raise NotImplementedError

    def log_likelihood(self, x):
        raise NotImplementedError

    def copy_sample(self):
        import copy
        return copy.deepcopy(self)


class GibbsSampling(Distribution):
    def resample(self, data=[]):
        raise NotImplementedError


class NetworkDistribution(Distribution):
    """abstractions.py:9-71."""

    @property
    def adjacency(self):
        raise NotImplementedError

    @property
    def weights(self):
        raise NotImplementedError

    def log_prior(self):
        return np.nan

    def log_probability(self, network):
        return self.log_likelihood(network) + self.log_prior()

    def initialize_hypers(self, A, W):
        self.adjacency.initialize_hypers(A)
        self.weights.initialize_hypers(W)

    def initialize_from_prior(self):
        self.adjacency.initialize_from_prior()
        self.weights.initialize_from_prior()


class AdjacencyDistribution(Distribution):
    """abstractions.py:74-182.  Must expose an N x N matrix of connection probabilities ``P``."""

    def __init__(self, N):
        self.N = N

    @property
    def P(self):
        return np.nan

    @property
    def safe_P(self):
        return np.clip(self.P, 1e-64, 1 - 1e-64)

    @property
    def is_deterministic(self):
        P = self.P
        return np.all(np.isclose(P, 0) | np.isclose(P, 1))

    def log_prior(self):
        return np.nan

    def log_likelihood(self, A):
        """abstractions.py:111-115 (host form, O(N^2) temporaries; the concrete classes override it with a
        device reduction).  Note the reference's asymmetry: the A term uses the unclipped P."""
        assert A.shape == (self.N, self.N)
        P = self.safe_P
        with np.errstate(divide="ignore", invalid="ignore"):
            return np.sum(A * np.log(self.P) + (1 - A) * np.log(1 - P))

    def log_probability(self, A):
        return self.log_likelihood(A) + self.log_prior()

    def rvs(self, size=[]):
        P = self.P
        return np.random.rand(*P.shape) < P

    def initialize_from_prior(self):
        raise NotImplementedError

    def initialize_hypers(self, A):
        pass

    def sample_predictive_parameters(self):
        raise NotImplementedError

    def sample_predictive_distribution(self):
        """abstractions.py:141-155."""
        N = self.N
        Prow, Pcol = self.sample_predictive_parameters()
        assert Prow[-1] == Pcol[-1]
        Arow = np.random.rand(N + 1) < Prow
        Acol = np.random.rand(N + 1) < Pcol
        Acol[-1] = Arow[-1]
        return Arow, Acol

    def approx_predictive_ll(self, Arow, Acol, M=100):
        """abstractions.py:157-182."""
        N = self.N
        assert Arow.shape == Acol.shape == (N + 1,)
        Acol = Acol[:-1]
        lps = np.zeros(M)
        for m in range(M):
            Prow, Pcol = self.sample_predictive_parameters()
            Prow = np.clip(Prow, 1e-64, 1 - 1e-64)
            Pcol = np.clip(Pcol, 1e-64, 1 - 1e-64)[:-1]
            lps[m] += (Arow * np.log(Prow) + (1 - Arow) * np.log(1 - Prow)).sum()
            lps[m] += (Acol * np.log(Pcol) + (1 - Acol) * np.log(1 - Pcol)).sum()
        return -np.log(M) + logsumexp(lps)


class WeightDistribution(Distribution):
    """abstractions.py:185-229."""

    def __init__(self, N):
        self.N = N

    def initialize_from_prior(self):
        raise NotImplementedError

    def initialize_hypers(self, W):
        pass

    def log_prior(self):
        raise NotImplementedError


class GaussianWeightDistribution(WeightDistribution):
    """abstractions.py:232-357: exposes Mu (N x N x B) and Sigma (N x N x B x B)."""

    def __init__(self, N, B=1):
        super(GaussianWeightDistribution, self).__init__(N)
        self.B = B

    @property
    def Mu(self):
        return np.nan

    @property
    def Sigma(self):
        return np.nan

    def log_likelihood(self, network):
        """abstractions.py:257-274: sum over edges of log N(W[m,n]; Mu[m,n], Sigma[m,n]) (host form, batched
        instead of N^2 scipy constructions; the concrete classes override it with device statistics)."""
        A, W = network
        N = self.N
        assert A.shape == (N, N) and W.shape == (N, N, self.B)
        e = np.asarray(A) != 0
        x = W[e] - self.Mu[e]
        Sig = self.Sigma[e]
        sol = np.linalg.solve(Sig, x[..., None])[..., 0]
        _, logdet = np.linalg.slogdet(Sig)
        return float(np.sum(-0.5 * (x * sol).sum(-1) - 0.5 * logdet - 0.5 * self.B * np.log(2 * np.pi)))

    def rvs(self, size=[]):
        """abstractions.py:276-287 (row-major draws, one multivariate normal per entry)."""
        N = self.N
        W = np.zeros((N, N, self.B))
        Mu, Sig = self.Mu, self.Sigma
        for m in range(N):
            for n in range(N):
                W[m, n] = np.random.multivariate_normal(Mu[m, n], Sig[m, n])
        return W

    def sample_predictive_distribution(self):
        """abstractions.py:289-311."""
        N = self.N
        Murow, Mucol, Lrow, Lcol = self.sample_predictive_parameters()
        assert np.all(Murow[-1] == Mucol[-1]) and np.all(Lrow[-1] == Lcol[-1])
        Wrow = np.zeros((N + 1, self.B))
        Wcol = np.zeros((N + 1, self.B))
        for n in range(N + 1):
            Wrow[n] = np.random.multivariate_normal(Murow[n], Lrow[n].dot(Lrow[n].T))
            if n < N:
                Wcol[n] = np.random.multivariate_normal(Mucol[n], Lcol[n].dot(Lcol[n].T))
            else:
                Wcol[n] = Wrow[n]
        return Wrow, Wcol

    def approx_predictive_ll(self, Arow, Acol, Wrow, Wcol, M=100):
        """abstractions.py:313-357, with the per-entry triangular solves batched."""
        N, B = self.N, self.B
        assert Arow.shape == Acol.shape == (N + 1,)

        def _ll(L, x):
            if x.shape[0] == 0:
                return 0.0
            xs = np.linalg.solve(L, x[..., None])[..., 0]          # L lower triangular
            ld = np.log(np.diagonal(L, axis1=-2, axis2=-1)).sum(-1)
            return float(np.sum(-0.5 * (xs * xs).sum(-1) - B / 2. * np.log(2 * np.pi) - ld))

        lps = np.zeros(M)
        er = np.asarray(Arow, dtype=bool)
        ec = np.asarray(Acol, dtype=bool).copy()
        ec[-1] = False
        for m in range(M):
            Murow, Mucol, Lrow, Lcol = self.sample_predictive_parameters()
            lps[m] += _ll(Lrow[er], Wrow[er] - Murow[er])
            lps[m] += _ll(Lcol[ec], Wcol[ec] - Mucol[ec])
        return -np.log(M) + logsumexp(lps)
