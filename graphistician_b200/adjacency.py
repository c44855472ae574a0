"""New-API adjacency distributions (mirrors ``graphistician.adjacency``).

Every O(N^2) step of ``resample(A)`` / ``log_likelihood(A)`` runs in libgk on the device copy of A:

  BetaBernoulliAdjacencyDistribution.resample   <- gk_block_stats with K = 1        (adjacency.py:153-181)
  SBMAdjacencyDistribution.resample_p / _c      <- gk_block_stats / gk_sbm_sweep    (adjacency.py:620-695)
  LatentDistanceAdjacencyDistribution.resample  <- gk_latent_hmc (K8 pair passes)   (adjacency.py:356-421)

The O(1) / O(C^2) conjugate draws stay on the host and consume ``np.random`` in the reference's order.
"""
import numpy as np
from scipy.special import gammaln

from . import _lib
from .abstractions import AdjacencyDistribution, GibbsSampling
from .device import DeviceLatent
from .old_abstractions import network_for


def logistic(x):
    """internals/utils.py:5-6."""
    return 1.0 / (1 + np.exp(-x))


def _beta_logpdf(x, a, b):
    return gammaln(a + b) - gammaln(a) - gammaln(b) + (a - 1) * np.log(x) + (b - 1) * np.log1p(-x)


def _edge_counts(A, dtype="f32"):
    """(off-diagonal edges, diagonal edges) of A from the device: K1 with a single block + the O(N) diagonal."""
    net = network_for(A)
    dev = net.device(dtype)
    n_edge, _, _ = dev.block_stats(np.zeros(net.N, np.int32), 1)
    return int(n_edge[0, 0]), int(np.count_nonzero(dev.diagA))


class FixedAdjacencyDistribution(AdjacencyDistribution, GibbsSampling):
    """adjacency.py:13-40."""

    def __init__(self, N, A):
        super(FixedAdjacencyDistribution, self).__init__(N)
        assert A.shape == (N, N)
        assert np.all((A == 1) | (A == 0))
        self._P = A

    @property
    def P(self):
        return self._P

    def initialize_from_prior(self):
        pass

    def log_prior(self):
        return 0

    def resample(self, data=[]):
        pass

    def sample_predictive_parameters(self):
        raise Exception("Cannot sample parameters for FixedAdjacencyDistribution")


class CompleteAdjacencyDistribution(FixedAdjacencyDistribution):
    """adjacency.py:43-54."""

    def __init__(self, N):
        super(CompleteAdjacencyDistribution, self).__init__(N, np.ones((N, N)))

    def sample_predictive_parameters(self):
        return np.ones(self.N + 1), np.ones(self.N + 1)

    def log_likelihood(self, A):
        return 0 if np.all(A == 1) else -np.inf


class EmptyAdjacencyDistribution(FixedAdjacencyDistribution):
    """adjacency.py:57-62."""

    def __init__(self, N):
        super(EmptyAdjacencyDistribution, self).__init__(N, np.zeros((N, N)))

    def sample_predictive_parameters(self):
        return np.zeros(self.N + 1), np.zeros(self.N + 1)


class _ScalarBernoulliLikelihood(object):
    """Shared device log likelihood for models whose P is p off the diagonal and p_diag on it: two counts."""

    def _ll_from_counts(self, A, p, p_diag):
        assert A.shape == (self.N, self.N)
        n_off, n_diag = _edge_counts(A, getattr(self, "dtype", "f32"))
        N = self.N
        with np.errstate(divide="ignore", invalid="ignore"):
            def term(n1, n, pr):
                # abstractions.py:111-115: log P unclipped on the A term, log(1 - safe_P) on the other
                a = n1 * np.log(pr) if n1 else 0.0
                b = (n - n1) * np.log(1 - np.clip(pr, 1e-64, 1 - 1e-64)) if n - n1 else 0.0
                return a + b
            return term(n_off, N * (N - 1), p) + term(n_diag, N, p_diag)


class BernoulliAdjacencyDistribution(AdjacencyDistribution, GibbsSampling, _ScalarBernoulliLikelihood):
    """adjacency.py:65-111: fixed connection probability."""

    def __init__(self, N, p=None, P=None, p_self=None):
        super(BernoulliAdjacencyDistribution, self).__init__(N)
        self.p = None
        if P is not None:
            assert P.shape == (N, N)
            assert np.all(P >= 0) and np.all(P <= 1)
            self._P = P.copy()
        else:
            assert p is not None
            assert p > 0 and p < 1
            self.p = p
            self.p_self = p_self if p_self else p
            self._P = p * np.ones((N, N))
            np.fill_diagonal(self._P, self.p_self)

    @property
    def P(self):
        return self._P

    def log_prior(self):
        return 0

    def log_likelihood(self, A):
        if self.p is None:                       # arbitrary P matrix: host form
            return super(BernoulliAdjacencyDistribution, self).log_likelihood(A)
        return self._ll_from_counts(A, self.p, self.p_self)

    def initialize_from_prior(self):
        pass

    def resample(self, data=[]):
        pass

    def sample_predictive_parameters(self):
        p_row, p_col = self.p * np.ones(self.N + 1), self.p * np.ones(self.N + 1)
        p_row[-1] = self.p_self
        p_col[-1] = self.p_self
        return p_row, p_col


class BetaBernoulliAdjacencyDistribution(AdjacencyDistribution, GibbsSampling, _ScalarBernoulliLikelihood):
    """adjacency.py:114-187."""

    def __init__(self, N, tau1=1.0, tau0=1.0, tau1_self=None, tau0_self=None, dtype="f32"):
        super(BetaBernoulliAdjacencyDistribution, self).__init__(N)
        assert tau1 > 0 and tau0 > 0
        self.tau1, self.tau0 = tau1, tau0
        self.dtype = dtype
        self.p = np.random.beta(tau1, tau0)
        if tau1_self is not None and tau0_self is not None:
            self.self_connection = True
            self.tau1_self, self.tau0_self = tau1_self, tau0_self
            self.p_self = np.random.beta(tau1_self, tau0_self)
        else:
            self.self_connection = False

    @property
    def P(self):
        P = self.p * np.ones((self.N, self.N))
        if self.self_connection:
            np.fill_diagonal(P, self.p_self)
        return P

    def initialize_from_prior(self):
        self.p = np.random.beta(self.tau1, self.tau0)
        if self.self_connection:
            self.p_self = np.random.beta(self.tau1_self, self.tau0_self)

    def rvs(self, size=[]):
        return np.random.rand(self.N, self.N) < self.P

    def log_prior(self):
        lp = _beta_logpdf(self.p, self.tau1, self.tau0)
        if self.self_connection:
            lp += _beta_logpdf(self.p_self, self.tau1_self, self.tau0_self)
        return lp

    def log_likelihood(self, A):
        return self._ll_from_counts(A, self.p, self.p_self if self.self_connection else self.p)

    def resample(self, A):
        """adjacency.py:153-181.  Counts from the device (K1, K = 1); the reference's quirk of updating p_self
        with (tau1, tau0) rather than (tau1_self, tau0_self) (:172-173) is kept."""
        N = self.N
        n_off, n_diag = _edge_counts(A, self.dtype)
        if self.self_connection:
            self.p = np.random.beta(self.tau1 + n_off, self.tau0 + N * (N - 1) - n_off)
            self.p_self = np.random.beta(self.tau1 + n_diag, self.tau0 + N - n_diag)
        else:
            n1 = n_off + n_diag
            self.p = np.random.beta(self.tau1 + n1, self.tau0 + N * N - n1)

    def sample_predictive_parameters(self):
        p_row, p_col = self.p * np.ones(self.N + 1), self.p * np.ones(self.N + 1)
        if self.self_connection:
            p_row[-1] = p_col[-1] = self.p_self
        return p_row, p_col


class LatentDistanceAdjacencyDistribution(AdjacencyDistribution, GibbsSampling):
    """adjacency.py:189-437:  l_n ~ N(0, sigma I),  A_ij ~ Bern(logistic(-||l_i - l_j||^2 + mu_0 + mu_self d_ij)).

    ``resample(A)`` keeps the reference's three hips-HMC moves (L adaptive, mu_0, mu_self; 10 leapfrog steps each,
    :381,402,419).  The L and mu_0 trajectories run on the device (11 K8 pair passes each); the mu_self target
    depends on mu_self only through the N diagonal terms, all equal, so its trajectory is O(1) host arithmetic
    on the self-edge count (the pair terms are the same constant at both ends of the accept test).
    Random draws per move, in hips' order: momentum ``np.random.randn(*q.shape)``, then ``np.log(np.random.rand())``.
    """

    def __init__(self, N, dim=2, sigma=2.0, mu0=0.0, mu_self=0.0, dtype="f32"):
        self.N = N
        self.dim = dim
        self.sigma = sigma
        self.mu_0 = mu0
        self.mu_self = mu_self
        self.dtype = dtype
        self.L = np.sqrt(self.sigma) * np.random.randn(N, dim)
        self._L_step_sz = 0.01
        self._L_accept_rate = 0.9
        self._dev = None

    @property
    def D(self):
        Mu = -((self.L[:, None, :] - self.L[None, :, :]) ** 2).sum(2)
        Mu += self.mu_0
        Mu += self.mu_self * np.eye(self.N)
        return Mu

    @property
    def P(self):
        return logistic(self.D)

    def initialize_from_prior(self):
        self.mu_0 = np.random.randn()
        self.mu_self = np.random.randn()
        self.L = np.sqrt(self.sigma) * np.random.randn(self.N, self.dim)

    def log_prior(self):
        """adjacency.py:230-242 (normal log densities, with their constants)."""
        def nlp(x, var):
            return -0.5 * np.log(2 * np.pi * var) - 0.5 * np.asarray(x) ** 2 / var
        return nlp(self.L, self.sigma).sum() + nlp(self.mu_0, 1.0) + nlp(self.mu_self, 1.0)

    # ---- device state ----------------------------------------------------------------------------------
    def _device_state(self, A):
        net = network_for(A).device(self.dtype)
        if self._dev is None or self._dev[0] is not net:
            if self._dev is not None:
                self._dev[1].close()
            self._dev = (net, DeviceLatent(net, self.dim))
        dev = self._dev[1]
        dev.set_params(self.L, self.mu_0, self.mu_self, self.sigma)
        return dev

    def __deepcopy__(self, memo):
        import copy
        new = self.__class__.__new__(self.__class__)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            setattr(new, k, None if k == "_dev" else copy.deepcopy(v, memo))
        return new

    def log_likelihood(self, A):
        """abstractions.py:111-115 with P of :216-219, as one fused pair pass (P is never materialised)."""
        assert A.shape == (self.N, self.N)
        return self._device_state(A).lp_grad()[0]

    def hmc_log_probability_and_grad(self, A):
        """(lp, dlp/dL, dlp/dmu_0, dlp/dmu_self) of adjacency.py:244-275 and of its autograd gradients."""
        ll, lp, gL, g0, gs = self._device_state(A).lp_grad()
        return lp, gL, g0, gs

    def sample_predictive_parameters(self):
        """adjacency.py:291-303, O(N) instead of an (N+1)^2 distance matrix."""
        lnew = np.sqrt(self.sigma) * np.random.randn(1, self.dim)
        d = -((self.L - lnew) ** 2).sum(1) + self.mu_0
        Prow = logistic(np.concatenate((d, [self.mu_0 + self.mu_self])))
        return Prow, Prow.copy()

    # ---- Gibbs / HMC -----------------------------------------------------------------------------------
    def resample(self, A, draws=None):
        """adjacency.py:356-367.  ``draws`` (parity tests) = dict(L=(p0, log_u), mu_0=(p0, log_u),
        mu_self=(p0, log_u)) injects the random numbers."""
        dev = self._device_state(A)
        self._resample_L(dev, None if draws is None else draws["L"])
        self._resample_mu_0(dev, None if draws is None else draws["mu_0"])
        self._resample_mu_self(dev, None if draws is None else draws["mu_self"])

    @staticmethod
    def _hmc_draws(shape, draws):
        if draws is not None:
            return draws
        p0 = np.random.randn(*shape)
        return p0, np.log(np.random.rand())

    def _resample_L(self, dev, draws=None, nsteps=10):
        """adjacency.py:370-385 (adaptive step size: hips hmc's 0.95 moving average, x1.02 / x0.98, clipped)."""
        p0, log_u = self._hmc_draws((self.N, self.dim), draws)
        accepted, self._last_L_ratio = dev.hmc(0, self._L_step_sz, nsteps, p0, log_u)
        if accepted:
            self.L = dev.get_params()[0]
        rate = 0.95 * self._L_accept_rate + 0.05 * float(accepted)
        step = self._L_step_sz * (1.02 if rate > 0.9 else 0.98)
        self._L_step_sz = min(max(step, 0.00005), 1.0)
        self._L_accept_rate = rate
        return accepted

    def _resample_mu_0(self, dev, draws=None, stepsz=0.005, nsteps=10):
        """adjacency.py:389-404."""
        p0, log_u = self._hmc_draws((), draws)
        accepted, self._last_mu0_ratio = dev.hmc(1, stepsz, nsteps, p0, log_u)
        if accepted:
            self.mu_0 = float(dev.get_params()[1])
        return accepted

    def _resample_mu_self(self, dev, draws=None, stepsz=0.005, nsteps=10):
        """adjacency.py:408-421.  lp(mu_self) = n_d eta - N softplus(eta) - mu_self^2 / 2 + const, eta = mu_0 + mu_self."""
        p0, log_u = self._hmc_draws((), draws)
        nd, N, mu0 = float(np.count_nonzero(dev.net.diagA)), float(self.N), self.mu_0

        def lp(ms):
            return nd * (mu0 + ms) - N * np.logaddexp(0.0, mu0 + ms) - 0.5 * ms * ms

        def dlp(ms):
            return nd - N * logistic(mu0 + ms) - ms

        q, p = float(self.mu_self), float(p0)
        p_curr = p
        p = p + stepsz * dlp(q) / 2.0
        for i in range(nsteps):
            q = q + stepsz * p
            if i < nsteps - 1:
                p = p + stepsz * dlp(q)
        p = p + stepsz * dlp(q) / 2.0
        ratio = lp(q) - lp(self.mu_self) + 0.5 * p_curr ** 2 - 0.5 * p ** 2
        accepted = bool(log_u < ratio)
        if accepted:
            self.mu_self = q
        return accepted

    def _resample_sigma(self):
        """adjacency.py:423-437 (sigma ~ IG(1, 1); not called by the reference's resample)."""
        a_post = 1.0 + self.L.size / 2.0
        b_post = 1.0 + (self.L ** 2).sum() / 2.0
        self.sigma = b_post / np.random.gamma(a_post)


class SBMAdjacencyDistribution(AdjacencyDistribution, GibbsSampling):
    """adjacency.py:440-702: stochastic block model over A with block vector c, block probabilities m and
    connection probabilities p[c, c'] (Dirichlet(pi) / Beta(tau1, tau0) priors)."""

    def __init__(self, N, C=2, c=None, m=None, pi=1.0, p=None, tau0=1.0, tau1=1.0, allow_self_connections=True,
                 special_case_self_conns=True, dtype="f32"):
        super(SBMAdjacencyDistribution, self).__init__(N)
        assert isinstance(C, (int, np.integer)) and C >= 1, "C must be a positive integer number of blocks"
        self.C = int(C)
        self.dtype = dtype
        if isinstance(pi, (int, float)):
            self.pi = pi * np.ones(C)
        else:
            assert isinstance(pi, np.ndarray) and pi.shape == (C,), "pi must be a sclar or a C-vector"
            self.pi = pi
        self.tau0, self.tau1 = tau0, tau1
        self.allow_self_connections = allow_self_connections
        if m is not None:
            assert isinstance(m, np.ndarray) and m.shape == (C,) and np.allclose(m.sum(), 1.0) \
                and np.amin(m) >= 0.0, "m must be a length C probability vector"
            self.m = m
        else:
            self.m = np.random.dirichlet(self.pi)
        if c is not None:
            assert isinstance(c, np.ndarray) and c.shape == (self.N,) and np.issubdtype(c.dtype, np.integer) \
                and np.amin(c) >= 0 and np.amax(c) <= self.C - 1, "c must be a length K-vector of block assignments"
            self.c = c.copy()
        else:
            self.c = np.random.choice(self.C, p=self.m, size=(self.N))
        if p is not None:
            if np.isscalar(p):
                assert p >= 0 and p <= 1, "p must be a probability"
                self.p = p * np.ones((C, C))
            else:
                assert isinstance(p, np.ndarray) and p.shape == (C, C) and np.amin(p) >= 0 \
                    and np.amax(p) <= 1.0, "p must be a CxC matrix of probabilities"
                self.p = p
        else:
            self.p = np.random.beta(self.tau1, self.tau0, size=(self.C, self.C))
        self.special_case_self_conns = special_case_self_conns
        if special_case_self_conns:
            self.p_self = np.random.beta(self.tau1, self.tau0)

    @property
    def P(self):
        P = self.p[np.ix_(self.c, self.c)]
        if self.special_case_self_conns:
            np.fill_diagonal(P, self.p_self)
        if not self.allow_self_connections:
            np.fill_diagonal(P, 0.0)
        return P

    def initialize_from_prior(self):
        self.p = np.random.beta(self.tau1, self.tau0, size=(self.C, self.C))
        if self.special_case_self_conns:
            self.p_self = np.random.beta(self.tau1, self.tau0)

    def log_prior(self):
        """adjacency.py:540-558."""
        lp = gammaln(self.pi.sum()) - gammaln(self.pi).sum() + ((self.pi - 1) * np.log(self.m)).sum()
        lp += _beta_logpdf(self.p, self.tau1, self.tau0).sum()
        if self.special_case_self_conns:
            lp += _beta_logpdf(self.p_self, self.tau1, self.tau0)
        lp += np.log(self.m)[self.c].sum()
        return lp

    def rvs(self, size=[]):
        return np.random.rand(self.N, self.N) < self.P

    def sample_predictive_parameters(self):
        """adjacency.py:562-573 in O(N)."""
        cnew = np.random.choice(self.C, p=self.m)
        cext = np.concatenate((self.c, [cnew]))
        Prow = self.p[cnew, cext].copy()
        Pcol = self.p[cext, cnew].copy()
        if not self.allow_self_connections:
            Prow[-1] = Pcol[-1] = 0.0
        return Prow, Pcol

    def invariant_sbm_order(self):
        """adjacency.py:575-592."""
        blocksz = np.bincount(self.c, minlength=self.C)
        corder = np.argsort(blocksz)
        newc = np.zeros(self.N)
        for c in np.arange(self.C):
            newc[self.c == corder[c]] = c
        return np.argsort(-newc)

    # ---- device statistics ---------------------------------------------------------------------------
    def _counts(self, A):
        """(edges, pairs) per cell -- diagonal pairs excluded iff self connections are special-cased -- and the
        diagonal's (edges, pairs)."""
        dev = network_for(A).device(self.dtype)
        n_edge = dev.block_stats(self.c, self.C)[0]
        d = dev.diag_stats(self.c, self.C)
        sz = d["d_cnt"]
        n_pair = np.outer(sz, sz)
        k = np.arange(self.C)
        if self.special_case_self_conns:
            n_pair[k, k] -= sz
        else:
            n_edge[k, k] += d["d_edge"]
        return n_edge, n_pair, int(d["d_edge"].sum()), self.N

    def log_likelihood(self, A):
        """abstractions.py:111-115 with P of :511-522, from block counts.  Where the reference's formula gives
        0 * log 0 = nan (no self connections allowed and none present) the term is taken as 0."""
        assert A.shape == (self.N, self.N)
        n_edge, n_pair, nd, N = self._counts(A)

        def cell(n1, n, pr):
            with np.errstate(divide="ignore", invalid="ignore"):
                a = np.where(n1 > 0, n1 * np.log(pr), 0.0)
                b = np.where(n - n1 > 0, (n - n1) * np.log(1 - np.clip(pr, 1e-64, 1 - 1e-64)), 0.0)
            return float(np.sum(a + b))

        k = np.arange(self.C)
        if not self.allow_self_connections:
            dev = network_for(A).device(self.dtype)
            d = dev.diag_stats(self.c, self.C)
            if not self.special_case_self_conns:        # take the diagonal back out of the cells
                n_edge = n_edge.copy(); n_edge[k, k] -= d["d_edge"]
                n_pair = n_pair.copy(); n_pair[k, k] -= d["d_cnt"]
            return cell(n_edge, n_pair, self.p) + (-np.inf if nd else 0.0)
        ll = cell(n_edge, n_pair, self.p)
        if self.special_case_self_conns:
            ll += cell(np.array(nd), np.array(N), self.p_self)
        return ll

    # ---- Gibbs -----------------------------------------------------------------------------------------
    def resample(self, A):
        """adjacency.py:615-618."""
        self.resample_p([A])
        self.resample_c(A)
        self.resample_m()

    def resample_p(self, As):
        """adjacency.py:620-648: the C^2 full-matrix masks become one K1 pass per network; one vectorised Beta call
        draws the cells in the loop's row-major order, then p_self."""
        n_edge = np.zeros((self.C, self.C)); n_pair = np.zeros((self.C, self.C)); nd = 0; ndp = 0
        for A in As:
            e, pr, d, dn = self._counts(A)
            n_edge += e; n_pair += pr; nd += d; ndp += dn
        self._last_counts = (n_edge, n_pair - n_edge)
        self.p = np.random.beta(self.tau1 + n_edge, self.tau0 + n_pair - n_edge)
        if self.special_case_self_conns:
            self.p_self = np.random.beta(self.tau1 + nd, self.tau0 + ndp - nd)

    def resample_c(self, A, uniforms=None, want_lp=False):
        """adjacency.py:650-695: sequential in n1, Bernoulli terms only, one uniform per node."""
        if self.C == 1:
            return None
        dev = network_for(A).device(self.dtype)
        u = np.random.random(self.N) if uniforms is None else uniforms
        with np.errstate(divide="ignore"):
            logm = np.log(self.m)
        mode = _lib.SELF_SKIP if self.special_case_self_conns else _lib.SELF_ONCE
        c, lp, self._n_changed = dev.sbm_sweep(self.c, logm, self.p, None, None, u, self_mode=mode,
                                               sampler=_lib.SAMPLER_CUMSUM, want_lp=want_lp)
        self.c = c
        return lp

    def resample_m(self):
        """adjacency.py:697-702."""
        self.m = np.random.dirichlet(self.pi + np.bincount(self.c, minlength=self.C))
