"""New-API Gaussian weight distributions (mirrors ``graphistician.weights``).

The O(N^2) part of every ``resample((A, W))`` / ``log_likelihood((A, W))`` -- the ``W[mask]`` gathers of
weights.py:154-162, 308-331, 399-413 and the per-node masked gathers of :415-488 -- runs in libgk
(gk_block_stats, gk_sbm_sweep) on the device copy of (A, W).  The NIW draws stay on the host.
"""
import numpy as np
from scipy.special import gammaln, multigammaln

from . import _lib
from .abstractions import GaussianWeightDistribution, GibbsSampling
from .internals.gaussian import Gaussian, niw_resample_batched
from .old_abstractions import network_for


def normal_inverse_wishart_log_prob(g):
    """internals/utils.py:33-51: log IW(sigma; nu_0, sigma_0) + log N(mu; mu_0, sigma / kappa_0)."""
    p = g.sigma.shape[0]
    nu, Psi = float(g.nu_0), np.asarray(g.sigma_0, dtype=np.float64)
    ld_psi, ld_sig = np.linalg.slogdet(Psi)[1], np.linalg.slogdet(g.sigma)[1]
    lp = 0.5 * nu * ld_psi - 0.5 * nu * p * np.log(2.0) - multigammaln(0.5 * nu, p) \
        - 0.5 * (nu + p + 1) * ld_sig - 0.5 * np.trace(np.linalg.solve(g.sigma, Psi))
    d = g.mu - g.mu_0
    cov = g.sigma / g.kappa_0
    lp += -0.5 * d.dot(np.linalg.solve(cov, d)) - 0.5 * np.linalg.slogdet(cov)[1] - 0.5 * p * np.log(2 * np.pi)
    return lp


def _dirichlet_logpdf(m, pi):
    return gammaln(pi.sum()) - gammaln(pi).sum() + ((pi - 1) * np.log(m)).sum()


def _uniform_stats(A, W, dtype):
    """(off-diagonal (n, sum w, sum w w^T), diagonal (n, sum w, sum w w^T)) over the edges of A, from the device."""
    net = network_for(A, W)
    dev = net.device(dtype)
    z = np.zeros(net.N, np.int32)
    n, sw, sww = dev.block_stats(z, 1)
    d = dev.diag_stats(z, 1)
    return (int(n[0, 0]), sw[0, 0], sww[0, 0]), (int(d["d_edge"][0]), d["d_sum_w"][0], d["d_sum_ww"][0])


class _UniformGaussianWeights(GaussianWeightDistribution, GibbsSampling):
    """Shared body of the two models with one Gaussian off the diagonal and one on it."""
    dtype = "f32"

    @property
    def Mu(self):
        Mu = np.tile(self._gaussian.mu[None, None, :], (self.N, self.N, 1))
        idx = np.arange(self.N)
        Mu[idx, idx, :] = self._self_gaussian.mu
        return Mu

    @property
    def Sigma(self):
        Sig = np.tile(self._gaussian.sigma[None, None, :, :], (self.N, self.N, 1, 1))
        idx = np.arange(self.N)
        Sig[idx, idx, :, :] = self._self_gaussian.sigma
        return Sig

    def log_likelihood(self, network):
        """abstractions.py:257-274 from the device statistics (no N^2 logpdf calls)."""
        A, W = network
        assert A.shape == (self.N, self.N) and W.shape == (self.N, self.N, self.B)
        off, dg = _uniform_stats(A, W, self.dtype)
        ll = self._gaussian.log_likelihood_from_stats(*off) if off[0] else 0.0
        ll += self._self_gaussian.log_likelihood_from_stats(*dg) if dg[0] else 0.0
        return ll

    def sample_predictive_parameters(self):
        Murow = np.tile(self._gaussian.mu[None, :], (self.N + 1, 1))
        Lrow = np.tile(self._gaussian.sigma_chol[None, :, :], (self.N + 1, 1, 1))
        Murow[-1, :] = self._self_gaussian.mu
        Lrow[-1, :, :] = self._self_gaussian.sigma_chol
        return Murow, Murow.copy(), Lrow, Lrow.copy()


class FixedGaussianWeightDistribution(_UniformGaussianWeights):
    """weights.py:8-62."""

    def __init__(self, N, B, mu, sigma, mu_self=None, sigma_self=None, dtype="f32"):
        super(FixedGaussianWeightDistribution, self).__init__(N, B)
        assert mu.shape == (B,)
        assert sigma.shape == (B, B)
        self.mu, self.sigma, self.dtype = mu, sigma, dtype
        self._gaussian = Gaussian(mu=mu, sigma=sigma)
        if mu_self is not None and sigma_self is not None:
            self._self_gaussian = Gaussian(mu=mu_self, sigma=sigma_self)
        else:
            self._self_gaussian = self._gaussian

    def initialize_from_prior(self):
        pass

    def initialize_hypers(self, W=None):
        pass

    def log_prior(self):
        return 0

    def resample(self, network=None):
        pass


class NIWGaussianWeightDistribution(_UniformGaussianWeights):
    """weights.py:65-162: one NIW Gaussian for off-diagonal edges, one for self edges."""

    def __init__(self, N, B=1, mu_0=None, Sigma_0=None, nu_0=None, kappa_0=None, dtype="f32"):
        super(NIWGaussianWeightDistribution, self).__init__(N, B)
        self.dtype = dtype
        mu_0 = np.zeros(B) if mu_0 is None else mu_0
        Sigma_0 = np.eye(B) if Sigma_0 is None else Sigma_0
        nu_0 = B + 2 if nu_0 is None else nu_0
        kappa_0 = 1.0 if kappa_0 is None else kappa_0
        self._gaussian = Gaussian(mu_0=mu_0, sigma_0=Sigma_0, nu_0=nu_0, kappa_0=kappa_0)
        self._self_gaussian = Gaussian(mu_0=mu_0, sigma_0=Sigma_0, nu_0=nu_0, kappa_0=kappa_0)

    def initialize_from_prior(self):
        self._gaussian.resample()
        self._self_gaussian.resample()

    def initialize_hypers(self, W):
        """weights.py:123-136."""
        self._gaussian.mu_0 = W.mean(axis=(0, 1))
        self._gaussian.sigma_0 = np.diag(W.var(axis=(0, 1)))
        self._gaussian.resample()
        W_self = W[np.arange(self.N), np.arange(self.N)]
        self._self_gaussian.mu_0 = W_self.mean(axis=0)
        self._self_gaussian.sigma_0 = np.diag(W_self.var(axis=0))
        self._self_gaussian.resample()

    def log_prior(self):
        return normal_inverse_wishart_log_prob(self._gaussian) + normal_inverse_wishart_log_prob(self._self_gaussian)

    def resample(self, network):
        """weights.py:154-162: NIW posterior draws given W[A offdiag] and W[A ondiag]; the gathers and their
        reduction to (n, sum w, sum w w^T) are K1 with a single block."""
        A, W = network
        off, dg = _uniform_stats(A, W, self.dtype)
        self._last_stats = (off, dg)
        self._gaussian.resample_from_stats(*off)
        self._self_gaussian.resample_from_stats(*dg)


class LowRankGaussianWeightDistribution(GaussianWeightDistribution, GibbsSampling):
    """weights.py:165-170."""

    def __init__(self, N, dim):
        raise NotImplementedError


class SBMGaussianWeightDistribution(GaussianWeightDistribution, GibbsSampling):
    """weights.py:172-498: block-structured Gaussian weights with its own block vector ``c``."""

    def __init__(self, N, B=1, C=2, pi=10.0, mu_0=None, Sigma_0=None, nu_0=None, kappa_0=None,
                 special_case_self_conns=True, dtype="f32"):
        super(SBMGaussianWeightDistribution, self).__init__(N, B)
        assert isinstance(C, (int, np.integer)) and C >= 1, "C must be a positive integer number of blocks"
        self.C = int(C)
        self.dtype = dtype
        if isinstance(pi, (int, float)):
            self.pi = pi * np.ones(C)
        else:
            assert isinstance(pi, np.ndarray) and pi.shape == (C,), "pi must be a sclar or a C-vector"
            self.pi = pi
        self.m = np.random.dirichlet(self.pi)
        self.c = np.random.choice(self.C, p=self.m, size=(self.N))
        mu_0 = np.zeros(B) if mu_0 is None else mu_0
        Sigma_0 = np.eye(B) if Sigma_0 is None else Sigma_0
        nu_0 = B + 2 if nu_0 is None else nu_0
        kappa_0 = 1.0 if kappa_0 is None else kappa_0
        self._gaussians = [[Gaussian(mu_0=mu_0, nu_0=nu_0, kappa_0=kappa_0, sigma_0=Sigma_0) for _ in range(C)]
                           for _ in range(C)]
        self.special_case_self_conns = special_case_self_conns
        if special_case_self_conns:
            self._self_gaussian = Gaussian(mu_0=mu_0, sigma_0=Sigma_0, nu_0=nu_0, kappa_0=kappa_0)

    @property
    def _Mu(self):
        return np.array([[g.mu for g in row] for row in self._gaussians])

    @property
    def _Sigma(self):
        return np.array([[g.sigma for g in row] for row in self._gaussians])

    @property
    def Mu(self):
        Mu = self._Mu[np.ix_(self.c, self.c)]
        if self.special_case_self_conns:
            idx = np.arange(self.N)
            Mu[idx, idx] = self._self_gaussian.mu
        return Mu

    @property
    def Sigma(self):
        Sigma = self._Sigma[np.ix_(self.c, self.c)]
        if self.special_case_self_conns:
            idx = np.arange(self.N)
            Sigma[idx, idx] = self._self_gaussian.sigma
        return Sigma

    def initialize_from_prior(self):
        self.m = np.random.dirichlet(self.pi)
        self.c = np.random.choice(self.C, p=self.m, size=(self.N))
        for row in self._gaussians:
            for g in row:
                g.resample()
        if self.special_case_self_conns:
            self._self_gaussian.resample()

    def initialize_hypers(self, W):
        """weights.py:281-306 (host-side KMeans initialiser)."""
        mu_0 = W.mean(axis=(0, 1))
        sigma_0 = np.diag(W.var(axis=(0, 1)))
        for row in self._gaussians:
            for g in row:
                g.mu_0 = mu_0
                g.sigma_0 = sigma_0 * (g.nu_0 - self.B - 1) / self.C
                g.resample()
        if self.special_case_self_conns:
            W_self = W[np.arange(self.N), np.arange(self.N)]
            self._self_gaussian.mu_0 = W_self.mean(axis=0)
            self._self_gaussian.sigma_0 = np.diag(W_self.var(axis=0))
            self._self_gaussian.resample()
        from sklearn.cluster import KMeans
        km = KMeans(n_clusters=self.C)
        km.fit(np.hstack((W[:, :, 0], W[:, :, 0].T)))
        self.c = km.labels_.astype(int)

    # ---- device statistics ---------------------------------------------------------------------------
    def _cell_stats(self, A, W):
        """Per-cell (n, sum w, sum w w^T) over edges, plus the diagonal's; the diagonal is folded into cell (k,k)
        unless self connections are special-cased (the masks of weights.py:308-314)."""
        dev = network_for(A, W).device(self.dtype)
        n, sw, sww = dev.block_stats(self.c, self.C)
        d = dev.diag_stats(self.c, self.C)
        if not self.special_case_self_conns:
            k = np.arange(self.C)
            n[k, k] += d["d_edge"]
            sw[k, k] += d["d_sum_w"]
            sww[k, k] += d["d_sum_ww"]
        return n, sw, sww, d

    def log_likelihood(self, network):
        """weights.py:316-331."""
        A, W = network
        assert A.shape == (self.N, self.N) and W.shape == (self.N, self.N, self.B)
        n, sw, sww, d = self._cell_stats(A, W)
        ll = 0.0
        for c1 in range(self.C):
            for c2 in range(self.C):
                if n[c1, c2]:
                    ll += self._gaussians[c1][c2].log_likelihood_from_stats(n[c1, c2], sw[c1, c2], sww[c1, c2])
        if self.special_case_self_conns and d["d_edge"].sum():
            ll += self._self_gaussian.log_likelihood_from_stats(d["d_edge"].sum(), d["d_sum_w"].sum(0),
                                                                d["d_sum_ww"].sum(0))
        return ll

    def log_prior(self):
        """weights.py:333-358."""
        lp = _dirichlet_logpdf(self.m, self.pi)
        for row in self._gaussians:
            for g in row:
                lp += normal_inverse_wishart_log_prob(g)
        if self.special_case_self_conns:
            lp += normal_inverse_wishart_log_prob(self._self_gaussian)
        lp += np.log(self.m)[self.c].sum()
        return lp

    def rvs(self, size=[]):
        """weights.py:361-373."""
        W = np.zeros((self.N, self.N, self.B))
        for c1 in range(self.C):
            for c2 in range(self.C):
                blk = (self.c == c1)[:, None] * (self.c == c2)[None, :]
                W[blk] = self._gaussians[c1][c2].rvs(size=blk.sum())
        if self.special_case_self_conns:
            for n in range(self.N):
                W[n, n] = self._self_gaussian.rvs()
        return W

    def sample_predictive_parameters(self):
        """weights.py:375-389."""
        c2 = np.random.choice(self.C, p=self.m)
        cext = np.concatenate((self.c, [c2]))
        Murow = np.array([self._gaussians[c2][c1].mu for c1 in cext])
        Lrow = np.array([self._gaussians[c2][c1].sigma_chol for c1 in cext])
        Mucol = np.array([self._gaussians[c1][c2].mu for c1 in cext])
        Lcol = np.array([self._gaussians[c1][c2].sigma_chol for c1 in cext])
        if self.special_case_self_conns:
            Murow[-1] = Mucol[-1] = self._self_gaussian.mu
            Lrow[-1] = Lcol[-1] = self._self_gaussian.sigma_chol
        return Murow, Mucol, Lrow, Lcol

    # ---- Gibbs -----------------------------------------------------------------------------------------
    def resample(self, network):
        """weights.py:394-397."""
        A, W = network
        self.resample_mu_and_Sig(A, W)
        self.resample_c(A, W)
        self.resample_m()

    def resample_mu_and_Sig(self, A, W):
        """weights.py:399-413.  The C^2 NIW posterior draws are made in one vectorised pass (chi-square Bartlett
        factors, as the old-API class does), then the self Gaussian."""
        n, sw, sww, d = self._cell_stats(A, W)
        self._last_stats = (n, sw, sww, d)
        mus, sigmas = niw_resample_batched(self._gaussians[0][0].hypparams, n, sw, sww)
        for c1 in range(self.C):
            for c2 in range(self.C):
                g = self._gaussians[c1][c2]
                g.mu, g.sigma = mus[c1, c2], sigmas[c1, c2]
        if self.special_case_self_conns:
            self._self_gaussian.resample_from_stats(int(d["d_edge"].sum()), d["d_sum_w"].sum(0), d["d_sum_ww"].sum(0))

    def resample_c(self, A, W, uniforms=None, want_lp=False):
        """weights.py:415-488: sequential in n1, Gaussian edge terms only, one uniform per node."""
        if self.C == 1:
            return None
        dev = network_for(A, W).device(self.dtype)
        u = np.random.random(self.N) if uniforms is None else uniforms
        with np.errstate(divide="ignore"):
            logm = np.log(self.m)
        mode = _lib.SELF_SKIP if self.special_case_self_conns else _lib.SELF_ONCE
        c, lp, self._n_changed = dev.sbm_sweep(self.c, logm, None, self._Mu, self._Sigma, u, self_mode=mode,
                                               sampler=_lib.SAMPLER_CUMSUM, want_lp=want_lp)
        self.c = c
        return lp

    def resample_m(self):
        """weights.py:493-498."""
        self.m = np.random.dirichlet(self.pi + np.bincount(self.c, minlength=self.C))
