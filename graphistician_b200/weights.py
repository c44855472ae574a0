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


class SBMGaussianWeightSharedCov(SBMGaussianWeightDistribution):
    """weights.py:501-611: the SBM weight model with ONE covariance shared by all C x C cells.  Cell means are conjugate
    given the covariance (GaussianFixedCov: prior N(mu_0, I), :523-526), the covariance is inverse-Wishart given the
    residuals of all off-diagonal edges (GaussianFixedMean around 0: prior IW(Sigma_0, nu_0), :520-521, 545-549); the self
    Gaussian keeps the parent's NIW model (:541-543).  Label sweep, m, predictive parameters and log likelihood are
    inherited: the per-cell (mu, Sigma) tables repeat the shared covariance.

    Device work per ``resample_mu_and_Sig``: the parent's single K1 pass (n, sum w, sum w w^T per cell); both conjugate
    updates are host O(C^2 B^2) arithmetic on those statistics (the reference gathers W[mask] per cell and W - Mu, :530-549):
      mean of cell (a, b):  precision = I + n Sigma^-1,  h = mu_0 + Sigma^-1 sum w
      shared covariance:    S = sum over cells of  sum ww^T - mu sw^T - sw mu^T + n mu mu^T   over off-diagonal edges"""

    def __init__(self, N, B=1, C=3, pi=10.0, mu_0=None, Sigma_0=None, nu_0=None, special_case_self_conns=True, dtype="f32"):
        super(SBMGaussianWeightSharedCov, self).__init__(N, B=B, C=C, pi=pi, mu_0=mu_0, Sigma_0=Sigma_0, nu_0=nu_0,
                                                         special_case_self_conns=special_case_self_conns, dtype=dtype)
        from .internals.gaussian import GaussianFixedMean
        self._prior_mu = np.zeros(B) if mu_0 is None else np.array(mu_0, dtype=np.float64)
        Sigma_0 = np.eye(B) if Sigma_0 is None else np.array(Sigma_0, dtype=np.float64)
        self._cov_model = GaussianFixedMean(mu=np.zeros(B), nu_0=(B + 2 if nu_0 is None else nu_0), lmbda_0=Sigma_0)
        for c1 in range(self.C):
            for c2 in range(self.C):          # GaussianFixedCov(mu_0, sigma_0 = I, sigma = shared): mean drawn from its prior
                self._gaussians[c1][c2].mu = np.random.multivariate_normal(self._prior_mu, np.eye(B))
        self._set_shared_cov()

    def _set_shared_cov(self):
        for row in self._gaussians:
            for g in row:
                g.sigma = self._cov_model.sigma.copy()

    @property
    def shared_sigma(self):
        return self._cov_model.sigma

    def initialize_from_prior(self):
        self.m = np.random.dirichlet(self.pi)
        self.c = np.random.choice(self.C, p=self.m, size=self.N)
        self._cov_model.resample()
        for row in self._gaussians:
            for g in row:
                g.mu = np.random.multivariate_normal(self._prior_mu, np.eye(self.B))
        self._set_shared_cov()
        if self.special_case_self_conns:
            self._self_gaussian.resample()

    def log_prior(self):
        """weights.py:553-575: the Dirichlet on m and the assignment terms only (the Gaussian terms are commented out there)."""
        return _dirichlet_logpdf(self.m, self.pi) + float(np.log(self.m)[self.c].sum())

    def resample_mu_and_Sig(self, A, W, draws=None):
        """weights.py:528-551.  ``draws`` (parity tests) = dict(z_mu = C*C x B standard normals) fixes the mean draws
        through mean + chol(cov) z; the self Gaussian and the covariance draw from ``np.random``."""
        n, sw, sww, d = self._cell_stats(A, W)
        self._last_stats = (n, sw, sww, d)
        B, C = self.B, self.C
        Lam = np.linalg.inv(self._cov_model.sigma)
        for c1 in range(C):
            for c2 in range(C):
                cov = np.linalg.inv(np.eye(B) + n[c1, c2] * Lam)
                mean = cov.dot(self._prior_mu + Lam.dot(sw[c1, c2]))
                if draws is not None:
                    self._gaussians[c1][c2].mu = mean + np.linalg.cholesky(cov).dot(draws["z_mu"][c1 * C + c2])
                else:
                    self._gaussians[c1][c2].mu = np.random.multivariate_normal(mean, cov)
        if self.special_case_self_conns:
            self._self_gaussian.resample_from_stats(int(d["d_edge"].sum()), d["d_sum_w"].sum(0), d["d_sum_ww"].sum(0))
        # residuals of the OFF-DIAGONAL edges around the new means (:545-549)
        n, sw, sww = n.copy(), sw.copy(), sww.copy()
        if not self.special_case_self_conns:
            k = np.arange(C)
            n[k, k] -= d["d_edge"]; sw[k, k] -= d["d_sum_w"]; sww[k, k] -= d["d_sum_ww"]
        S = np.zeros((B, B))
        for c1 in range(C):
            for c2 in range(C):
                mu = self._gaussians[c1][c2].mu
                S += sww[c1, c2] - np.outer(mu, sw[c1, c2]) - np.outer(sw[c1, c2], mu) + n[c1, c2] * np.outer(mu, mu)
        self._cov_model.resample_from_stats(int(n.sum()), 0.5 * (S + S.T))
        self._set_shared_cov()


class LatentDistanceGaussianWeightDistribution(GaussianWeightDistribution, GibbsSampling):
    """weights.py:613-901:  l_n ~ N(0, eta I),  W_ij | A_ij = 1 ~ N(-||l_i - l_j||^2 + b, Sigma) for i != j (B = 1),
    self weights from their own NIW Gaussian.

    ``resample((A, W))`` keeps the reference's five moves (:757-763):
      L      hips HMC, step 0.005, 10 leapfrog steps (:765-780): 11 row passes of k_latw_rows over (A, W) and (A^T, W^T)
             on the device (gk_latent_set_weight_mode + gk_latent_hmc), momentum and log-uniform from np.random in hips' order
      b      exact Gaussian conditional (:838-860) from one statistics pass: n, sum r, sum r^2 with r = W + d^2 - b
      Sigma  inverse-Wishart given the residuals around the NEW b (:863-867): sum (r - delta)^2 from the same three numbers
      self   NIW posterior draw from the diagonal statistics (:869-872)
      eta    inverse gamma (:874-890)."""

    def __init__(self, N, B=1, dim=2, b=0.5, sigma=None, Sigma_0=None, nu_0=None, mu_self=0.0, eta=0.01, dtype="f32"):
        super(LatentDistanceGaussianWeightDistribution, self).__init__(N, B)
        from .internals.gaussian import GaussianFixedMean
        self.dim, self.b, self.eta, self.dtype = dim, b, eta, dtype
        self.L = np.sqrt(eta) * np.random.randn(N, dim)
        Sigma_0 = np.eye(B) if Sigma_0 is None else Sigma_0
        nu_0 = B + 2 if nu_0 is None else nu_0
        self.cov = GaussianFixedMean(mu=np.zeros(B), sigma=sigma, lmbda_0=Sigma_0, nu_0=nu_0)
        self._self_gaussian = Gaussian(mu_0=mu_self * np.ones(B), sigma_0=Sigma_0, nu_0=nu_0, kappa_0=1.0)
        self._dev = None

    @property
    def D(self):
        return ((self.L[:, None, :] - self.L[None, :, :]) ** 2).sum(2)

    @property
    def Mu(self):
        Mu = np.tile((-self.D + self.b)[:, :, None], (1, 1, self.B))
        idx = np.arange(self.N)
        Mu[idx, idx, :] = self._self_gaussian.mu
        return Mu

    @property
    def Sigma(self):
        Sig = np.tile(self.cov.sigma[None, None, :, :], (self.N, self.N, 1, 1))
        idx = np.arange(self.N)
        Sig[idx, idx] = self._self_gaussian.sigma
        return Sig

    def initialize_from_prior(self):
        self.L = np.sqrt(self.eta) * np.random.randn(self.N, self.dim)
        self.cov.resample()

    def log_prior(self):
        """weights.py:679-695 (note: the L term there is a standard normal, not N(0, eta))."""
        from scipy.stats import norm, invgamma, invwishart
        lp = invgamma.logpdf(self.eta, 1, 1)
        lp += norm.logpdf(self.b, 0, 1)
        lp += norm.logpdf(self.L, 0, 1).sum()
        lp += invwishart.logpdf(self.cov.sigma, self.cov.nu_0, self.cov.lmbda_0)
        lp += normal_inverse_wishart_log_prob(self._self_gaussian)
        return lp

    # ---- device state ----------------------------------------------------------------------------------
    def _device_state(self, A, W):
        from .device import DeviceLatent
        assert self.B == 1, "the distance model of the weights is one-dimensional (weights.py:705)"
        net = network_for(A, W).device(self.dtype)
        if self._dev is None or self._dev[0] is not net:
            if self._dev is not None:
                self._dev[1].close()
            self._dev = (net, DeviceLatent(net, self.dim))
        dev = self._dev[1]
        dev.set_weight_mode(True, 1.0 / float(self.cov.sigma[0, 0]))
        dev.set_params(self.L, self.b, 0.0, self.eta)
        return dev

    def __deepcopy__(self, memo):
        import copy
        new = self.__class__.__new__(self.__class__)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            setattr(new, k, None if k == "_dev" else copy.deepcopy(v, memo))
        return new

    def hmc_log_probability_and_grad(self, A, W):
        """(lp, d lp / d L, d lp / d b) of weights.py:697-731 and of its autograd gradient."""
        ll, lp, gL, gb, _ = self._device_state(A, W).lp_grad()
        return lp, gL, gb

    def log_likelihood(self, network):
        """abstractions.py:257-274 with Mu / Sigma of :649-668: Gaussian log density of the off-diagonal edges from
        (n, sum r, sum r^2) of one device pass, plus the self edges from the diagonal statistics."""
        A, W = network
        dev = self._device_state(A, W)
        n, sr, srr = dev.weight_stats()
        sig = float(self.cov.sigma[0, 0])
        ll = -0.5 * srr / sig - 0.5 * n * np.log(2 * np.pi * sig)
        d = dev.net.diag_stats(np.zeros(self.N, np.int32), 1)
        if d["d_edge"].sum():
            ll += self._self_gaussian.log_likelihood_from_stats(d["d_edge"].sum(), d["d_sum_w"].sum(0), d["d_sum_ww"].sum(0))
        return float(ll)

    def sample_predictive_parameters(self):
        """weights.py:735-761, O(N) instead of an (N+1)^2 distance matrix."""
        lnew = np.sqrt(self.eta) * np.random.randn(1, self.dim)
        mu = -((self.L - lnew) ** 2).sum(1) + self.b
        Mu_row = np.tile(np.concatenate((mu, [self.b]))[:, None], (1, self.B))
        Mu_row[-1] = self._self_gaussian.mu
        Lc = np.linalg.cholesky(self.cov.sigma)
        L_row = np.tile(Lc[None, :, :], (self.N + 1, 1, 1))
        L_row[-1] = np.linalg.cholesky(self._self_gaussian.sigma)
        return Mu_row, Mu_row.copy(), L_row, L_row.copy()

    # ---- Gibbs / HMC -----------------------------------------------------------------------------------
    def resample(self, network, draws=None):
        """weights.py:757-763.  ``draws`` (parity tests) = dict(L=(p0, log_u), b=z) injects the HMC and b draws."""
        A, W = network
        dev = self._device_state(A, W)
        self._resample_L(dev, None if draws is None else draws.get("L"))
        stats = dev.weight_stats()
        delta = self._resample_b(stats, None if draws is None else draws.get("b"))
        self._resample_cov(stats, delta)
        self._resample_self_gaussian(dev)
        self._resample_eta()

    def _resample_L(self, dev, draws=None, stepsz=0.005, nsteps=10):
        if draws is None:
            p0 = np.random.randn(self.N, self.dim)
            log_u = np.log(np.random.rand())
        else:
            p0, log_u = draws
        accepted, self._last_L_ratio = dev.hmc(0, stepsz, nsteps, p0, log_u)
        if accepted:
            self.L = dev.get_params()[0]
        return accepted

    def _resample_b(self, stats, z=None):
        """weights.py:838-860: X = W + D on the off-diagonal edges, sum X = sum r + n b."""
        n, sr, srr = stats
        sigma = float(self.cov.sigma[0, 0])
        sigma_post = 1.0 / (1.0 + n / sigma)
        mu_post = sigma_post * ((sr + n * self.b) / sigma)
        b_old = self.b
        self.b = mu_post + np.sqrt(sigma_post) * (np.random.randn() if z is None else z)
        return self.b - b_old

    def _resample_cov(self, stats, delta):
        """weights.py:863-867: residuals around the new mean, sum (r - delta)^2 = sum r^2 - 2 delta sum r + n delta^2."""
        n, sr, srr = stats
        S = srr - 2.0 * delta * sr + n * delta * delta
        self.cov.resample_from_stats(int(round(n)), np.array([[max(S, 0.0)]]))

    def _resample_self_gaussian(self, dev):
        d = dev.net.diag_stats(np.zeros(self.N, np.int32), 1)
        self._self_gaussian.resample_from_stats(int(d["d_edge"].sum()), d["d_sum_w"].sum(0), d["d_sum_ww"].sum(0))

    def _resample_eta(self):
        """weights.py:874-890: eta ~ IG(1 + N dim / 2, 1 + sum L^2 / 2)."""
        from scipy.stats import invgamma
        self.eta = invgamma.rvs(a=1.0 + self.L.size / 2.0, scale=1.0 + (self.L ** 2).sum() / 2.0)
