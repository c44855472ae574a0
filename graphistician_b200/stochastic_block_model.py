"""Old-API Gaussian stochastic block model with a block vector ``c`` shared by adjacency and weights.

Mirrors ``graphistician.stochastic_block_model.GaussianStochasticBlockModel`` (stochastic_block_model.py:589;
constructor :39-102, Gibbs :206-309).  Every O(N^2) step of ``resample(network)`` runs in libgk:

  weight models + p   <- gk_block_stats          (replaces the C^2 masks of :221-250)
  c                   <- gk_sbm_sweep            (replaces the N*C*N loop of :252-302, same serial semantics)

The O(C^2) Beta / NIW / Dirichlet draws stay on the host and consume the global ``np.random`` stream in the
reference's order: weight models (row-major cells) -> p (row-major) -> N uniforms for c -> m.
"""
import copy

import numpy as np
from scipy.special import gammaln, psi

from . import _lib
from .internals.weights import GaussianWeights
from .internals.gaussian import niw_resample_batched
from .old_abstractions import (NetworkDistribution, GaussianWeightedNetworkDistribution, FixedGaussianNetwork,
                               as_network)


class _StochasticBlockModelBase(NetworkDistribution):
    _weight_class = None
    _default_weight_hypers = {}

    def __init__(self, N, B, C=1, c=None, m=None, pi=1.0, p=None, tau0=1.0, tau1=1.0, weight_hypers={},
                 allow_self_connections=True, dtype="f32"):
        self.N, self.B = N, B
        assert isinstance(C, (int, np.integer)) and C >= 1, "C must be a positive integer number of blocks"
        self.C = int(C)
        if isinstance(pi, (int, float)):
            self.pi = pi * np.ones(C)
        else:
            assert isinstance(pi, np.ndarray) and pi.shape == (C,), "pi must be a sclar or a C-vector"
            self.pi = pi
        self.tau0, self.tau1 = tau0, tau1
        self.allow_self_connections = allow_self_connections
        self.dtype = dtype                       # device storage of W: "f32" (default) or "f64" (parity mode)
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
        self.weight_hypers = copy.deepcopy(self._default_weight_hypers)
        self.weight_hypers.update(weight_hypers)
        self.weight_models = [[self._weight_class(self.B, **self.weight_hypers) for _ in range(self.C)]
                              for _ in range(self.C)]

    def __deepcopy__(self, memo):
        """copy_sample(): everything but the handle on the device network of the last sweep."""
        new = self.__class__.__new__(self.__class__)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            if k != "_swept":
                setattr(new, k, copy.deepcopy(v, memo))
        return new

    # ---- N x N views (stochastic_block_model.py:104-113, 594-628) -------------------------------------
    @property
    def P(self):
        P = self.p[np.ix_(self.c, self.c)]
        if not self.allow_self_connections:
            np.fill_diagonal(P, 0.0)
        return P

    @property
    def mus(self):
        return np.array([[wm.mu for wm in row] for row in self.weight_models])

    @property
    def sigmas(self):
        return np.array([[wm.sigma for wm in row] for row in self.weight_models])

    @property
    def Mu(self):
        return self.mus[np.ix_(self.c, self.c)]

    @property
    def Sigma(self):
        return self.sigmas[np.ix_(self.c, self.c)]

    def log_prior(self):
        """stochastic_block_model.py:115-129: Dirichlet(m) + Beta(p) + sum log m[c]."""
        lp = gammaln(self.pi.sum()) - gammaln(self.pi).sum() + ((self.pi - 1) * np.log(self.m)).sum()
        t1, t0 = self.tau1, self.tau0
        lp += (gammaln(t1 + t0) - gammaln(t1) - gammaln(t0) + (t1 - 1) * np.log(self.p)
               + (t0 - 1) * np.log(1 - self.p)).sum()
        lp += np.log(self.m)[self.c].sum()
        return lp

    def rvs(self, size=[]):
        """stochastic_block_model.py:152-163 (block-wise draws, row-major cells)."""
        A = np.zeros((self.N, self.N))
        W = np.zeros((self.N, self.N, self.B))
        for c1 in range(self.C):
            for c2 in range(self.C):
                blk = (self.c == c1)[:, None] * (self.c == c2)[None, :]
                A[blk] = np.random.rand(blk.sum()) < self.p[c1, c2]
                W[blk, :] = self.weight_models[c1][c2].rvs(size=blk.sum())
        return FixedGaussianNetwork(A, W)

    def invariant_sbm_order(self):
        """stochastic_block_model.py:165-182."""
        blocksz = np.bincount(self.c, minlength=self.C)
        corder = np.argsort(blocksz)
        newc = np.zeros(self.N)
        for c in np.arange(self.C):
            newc[self.c == corder[c]] = c
        return np.argsort(-newc)

    # ---- device statistics ---------------------------------------------------------------------------
    def _cell_stats(self, network):
        """Per-cell (n_edge, n_noedge, sum_w, sum_ww) with the diagonal inside cell (k,k) (old API)."""
        dev = network.device(self.dtype)
        st = None
        last = getattr(self, "_swept", None)          # (device network, labels) the last resample_c left behind
        if last is not None and last[0] is dev and hasattr(dev, "sbm_sweep_stats") and np.array_equal(last[1], self.c):
            st = dev.sbm_sweep_stats(self.C)          # folded from that sweep's own statistics: no third pass over A / W
        n_edge, sw, sww = st if st is not None else dev.block_stats(self.c, self.C)
        d = dev.diag_stats(self.c, self.C)
        k = np.arange(self.C)
        n_edge[k, k] += d["d_edge"]
        sw[k, k] += d["d_sum_w"]
        sww[k, k] += d["d_sum_ww"]
        sz = d["d_cnt"]
        n_pair = np.outer(sz, sz)                 # ordered pairs incl. the diagonal
        return n_edge, n_pair - n_edge, sw, sww

    def log_likelihood(self, network):
        """Bernoulli term on A plus Gaussian term on W where A == 1 (successor formulas abstractions.py:111-115,
        257-274), evaluated from the device block statistics instead of N^2 logpdf calls."""
        network = as_network(network)
        n_edge, n_noedge, sw, sww = self._cell_stats(network)
        p = np.clip(self.p, 1e-300, 1.0)
        with np.errstate(divide="ignore", invalid="ignore"):
            ll = np.nansum(n_edge * np.log(p)) + np.nansum(n_noedge * np.log1p(-np.clip(self.p, 0, 1 - 1e-16)))
        for c1 in range(self.C):
            for c2 in range(self.C):
                if n_edge[c1, c2]:
                    ll += self.weight_models[c1][c2].log_likelihood_from_stats(n_edge[c1, c2], sw[c1, c2], sww[c1, c2])
        return ll


class _GibbsSBM(_StochasticBlockModelBase):
    def resample(self, networks=[]):
        """stochastic_block_model.py:206-230."""
        if not isinstance(networks, list):
            networks = [networks]
        assert len(networks) == 1                 # :213 (TODO in the reference)
        network = as_network(networks[0])
        n_edge, n_noedge, sw, sww = self._cell_stats(network)
        self._resample_weight_models(n_edge, sw, sww)      # :221-226
        self._resample_p_from_counts(n_edge, n_noedge)
        self.resample_c(network)
        self.resample_m()

    def _resample_weight_models(self, n_edge, sw, sww):
        """All C^2 NIW posterior draws in one vectorised pass (internals/weights.py:88-101 per cell).  Draw order on
        the np.random stream: chi-squares for every cell (row-major), then the Bartlett normals, then the means."""
        wm0 = self.weight_models[0][0]
        if self.B == 0:
            return
        mus, sigmas = niw_resample_batched(wm0.hypparams, n_edge, sw, sww)
        for c1 in range(self.C):
            for c2 in range(self.C):
                wm = self.weight_models[c1][c2]
                wm.mu, wm.sigma = mus[c1, c2], sigmas[c1, c2]
        self._mus_cache = (mus, sigmas)

    def _resample_p_from_counts(self, n_edge, n_noedge):
        # :236-250; one vectorised legacy-RandomState call draws the cells in row-major order, as the loop does
        self.p = np.random.beta(self.tau1 + n_edge, self.tau0 + n_noedge)

    def resample_p(self, As):
        n_edge = np.zeros((self.C, self.C)); n_noedge = np.zeros((self.C, self.C))
        for A in As:
            e, ne, _, _ = self._cell_stats(as_network(A))
            n_edge += e; n_noedge += ne
        self._resample_p_from_counts(n_edge, n_noedge)

    def resample_c(self, network, uniforms=None, want_lp=False):
        """stochastic_block_model.py:252-302 -- sequential in n1, one uniform per node (:302)."""
        if self.C == 1:
            return None
        network = as_network(network)
        u = np.random.random(self.N) if uniforms is None else uniforms
        dev = network.device(self.dtype)
        with np.errstate(divide="ignore"):
            logm = np.log(self.m)
        has_w = self.B > 0 and network.W is not None
        c, lp, self._n_changed = dev.sbm_sweep(self.c, logm, self.p, self.mus if has_w else None,
                                               self.sigmas if has_w else None, u, self_mode=_lib.SELF_ONCE,
                                               sampler=_lib.SAMPLER_CUMSUM, want_lp=want_lp)
        self.c = c
        self._swept = (dev, c.copy())
        return lp

    def resample_m(self):
        """stochastic_block_model.py:304-309."""
        self.m = np.random.dirichlet(self.pi + np.bincount(self.c, minlength=self.C))


class _MeanFieldSBM(_StochasticBlockModelBase):
    """stochastic_block_model.py:310-584.  The N^2 reductions run in libgk for FIXED networks (E_A = A, E_W = W,
    E_WWT = W W^T -- what FixedGaussianNetwork reports, the only network class the reference's drivers build):

      mf_update_p, weight-model statistics  <- gk_mf_sbm_soft_stats   (the C x C loops of :340-344, 366-389)
      mf_update_c                            <- gk_mf_sbm_update_c    (the N*C*N*C loop of :392-445, serial in n1)

    `literal_reverse_edge` (default True) keeps a reference quirk: the reverse-edge branch `if n2 != n1:` (:425) sits
    one indentation level outside the `for cn2` loop, so only cn2 == C-1 contributes to the reverse term; False gives
    the evidently intended sum over cn2."""
    literal_reverse_edge = True

    def __init__(self, N, B, **kwargs):
        super(_MeanFieldSBM, self).__init__(N, B, **kwargs)
        self.mf_pi = np.ones(self.C)
        self.mf_m = np.random.dirichlet(10 * np.ones(self.C), size=(self.N,))      # :323 symmetry breaking
        self.mf_tau0 = self.tau0 * np.ones((self.C, self.C))
        self.mf_tau1 = self.tau1 * np.ones((self.C, self.C))

    # ---- device statistics ---------------------------------------------------------------------------
    def _soft_stats(self, network):
        """sum_ij mf_m[i,c1] mf_m[j,c2] A_ij (1, W_ij, W_ij W_ij^T), diagonal pairs included (as :340-344, :378-384)."""
        dev = network.device(self.dtype)
        n, sw, sww = dev.mf_soft_stats(self.mf_m)
        e = dev.diagA != 0
        if e.any():
            me = self.mf_m[e]
            n += me.T.dot(me)
            if self.B > 0:
                w = dev.diagW[e]
                sw += np.einsum("ik,il,ib->klb", me, me, w)
                sww += np.einsum("ik,il,ib,ic->klbc", me, me, w, w)
        return n, sw, sww

    def meanfieldupdate(self, network):
        """:327-344"""
        network = as_network(network)
        self.mf_update_p(network)
        self.mf_update_c(network)
        self.mf_update_m()
        self._mf_update_weight_models(network)

    def meanfield_sgdstep(self, network, minibatchfrac, stepsize):
        """:346-356 (the reference's order: p, m, c, weights)"""
        network = as_network(network)
        self.mf_update_p(network, stepsize=stepsize)
        self.mf_update_m(stepsize=stepsize)
        self.mf_update_c(network, stepsize=stepsize)
        self._mf_update_weight_models(network, stepsize=stepsize)

    def _mf_update_weight_models(self, network, stepsize=1.0):
        if self.B == 0:
            return
        n, sw, sww = self._soft_stats(network)
        for c1 in range(self.C):
            for c2 in range(self.C):
                self.weight_models[c1][c2].meanfieldupdate_from_stats(n[c1, c2], sw[c1, c2], sww[c1, c2], stepsize)

    def mf_update_m(self, stepsize=1.0):
        """:358-364"""
        pi_hat = self.pi + self.mf_m.sum(axis=0)
        self.mf_pi = (1.0 - stepsize) * self.mf_pi + stepsize * pi_hat

    def mf_update_p(self, network, stepsize=1.0):
        """:366-389 -- tau1_hat = tau1 + mf_m^T E_A mf_m, tau0_hat = tau0 + mf_m^T (1 - E_A) mf_m."""
        network = as_network(network)
        n1, _, _ = self._soft_stats(network)
        ms = self.mf_m.sum(0)
        n0 = np.outer(ms, ms) - n1
        self.mf_tau1 = (1.0 - stepsize) * self.mf_tau1 + stepsize * (self.tau1 + n1)
        self.mf_tau0 = (1.0 - stepsize) * self.mf_tau0 + stepsize * (self.tau0 + n0)

    def _mf_cell_tables(self):
        """Per-cell coefficients of the pair terms of mf_update_c (:410-435) in the layout gk_mf_sbm_update_c takes."""
        C, B = self.C, self.B
        S = 1 + B + B * (B + 1) // 2
        E_ln_p = psi(self.mf_tau1) - psi(self.mf_tau0 + self.mf_tau1)          # Beta.expected_log_p
        E_ln_notp = psi(self.mf_tau0) - psi(self.mf_tau0 + self.mf_tau1)
        coef = np.zeros((C, C, S))
        coef[:, :, 0] = E_ln_p - E_ln_notp
        if B > 0:
            for c1 in range(C):
                for c2 in range(C):
                    c0, c1v, C2 = self.weight_models[c1][c2].expected_log_likelihood_coefficients()
                    coef[c1, c2, 0] += c0
                    coef[c1, c2, 1:1 + B] = c1v
                    j = 1 + B
                    for b in range(B):
                        for b2 in range(b, B):
                            coef[c1, c2, j] = C2[b, b] if b2 == b else C2[b, b2] + C2[b2, b]
                            j += 1
        mask = np.ones(C)
        if self.literal_reverse_edge:
            mask[:] = 0.0
            mask[C - 1] = 1.0
        base_out, coef_out = E_ln_notp, coef
        base_in = E_ln_notp.T * mask[None, :]
        coef_in = np.transpose(coef, (1, 0, 2)) * mask[None, :, None]
        return base_out, base_in, coef_out, coef_in

    def mf_update_c(self, network, stepsize=1.0, want_lp=False):
        """:392-445, sequential in n1."""
        network = as_network(network)
        dev = network.device(self.dtype)
        base_out, base_in, coef_out, coef_in = self._mf_cell_tables()
        self.mf_m, lp = dev.mf_update_c(self.mf_m, self.mf_expected_log_m(), base_out, base_in, coef_out, coef_in,
                                        stepsize=stepsize, want_lp=want_lp)
        return lp

    # ---- expectations (:457-531) ------------------------------------------------------------------------
    def _mf_mix(self, cell):
        """sum_{c1,c2} mf_m[:,c1] mf_m[:,c2]^T cell[c1,c2,...]"""
        return np.einsum("ik,jl,kl...->ij...", self.mf_m, self.mf_m, cell)

    def mf_expected_p(self):
        E_p = self._mf_mix(self.mf_tau1 / (self.mf_tau0 + self.mf_tau1))
        if not self.allow_self_connections:
            np.fill_diagonal(E_p, 0.0)
        return E_p

    def mf_expected_notp(self):
        return 1.0 - self.mf_expected_p()

    def mf_expected_log_p(self):
        E = self._mf_mix(psi(self.mf_tau1) - psi(self.mf_tau0 + self.mf_tau1))
        if not self.allow_self_connections:
            np.fill_diagonal(E, -np.inf)
        return E

    def mf_expected_log_notp(self):
        E = self._mf_mix(psi(self.mf_tau0) - psi(self.mf_tau0 + self.mf_tau1))
        if not self.allow_self_connections:
            np.fill_diagonal(E, 0.0)
        return E

    def mf_expected_m(self):
        return self.mf_pi / self.mf_pi.sum()

    def mf_expected_log_m(self):
        return psi(self.mf_pi) - psi(self.mf_pi.sum())

    def expected_log_likelihood(self, network):
        raise NotImplementedError()               # :531-532

    def get_vlb(self):
        """:536-568.  Host arithmetic over N*C + C*C terms, as in the reference (no N^2 work): cross entropies minus
        negentropies of q(c), q(p), q(m), plus the weight models' NIW terms.  Like the reference it holds no data term."""
        from scipy.special import gammaln
        E_ln_m = self.mf_expected_log_m()
        vlb = float(self.mf_m.dot(E_ln_m).sum())                                   # Discrete().negentropy(E_x, E_ln_p)
        with np.errstate(divide="ignore", invalid="ignore"):
            vlb -= float(np.nan_to_num((self.mf_m * np.log(self.mf_m)).sum(1)).sum())   # Discrete(mf_m[n]).negentropy()

        def dirichlet_negentropy(gamma, E_ln_g):                                   # internals/distributions.py:327-340
            return float(gammaln(gamma.sum(-1, keepdims=True)).sum() - gammaln(gamma).sum() + ((gamma - 1) * E_ln_g).sum())

        E_ln_p = psi(self.mf_tau1) - psi(self.mf_tau0 + self.mf_tau1)
        E_ln_notp = psi(self.mf_tau0) - psi(self.mf_tau0 + self.mf_tau1)
        E_q = np.stack([E_ln_p, E_ln_notp], -1)
        # Beta(tau1, tau0).negentropy(E_ln_p, E_ln_notp), literally: with scalar hyperparameters gamma has shape (1, 2), so
        # the log normaliser enters once while the (gamma - 1) E[ln g] term broadcasts over all C x C cells
        tau = np.concatenate((np.atleast_1d(self.tau1)[..., None], np.atleast_1d(self.tau0)[..., None]), -1).astype(float)
        vlb += dirichlet_negentropy(tau, E_q)
        vlb -= dirichlet_negentropy(np.stack([self.mf_tau1, self.mf_tau0], -1), E_q)   # Beta(mf_tau1, mf_tau0).negentropy()
        vlb += dirichlet_negentropy(np.asarray(self.pi, float), E_ln_m)
        vlb -= dirichlet_negentropy(np.asarray(self.mf_pi, float), E_ln_m)
        for c1 in range(self.C):
            for c2 in range(self.C):
                vlb += self.weight_models[c1][c2].get_vlb()
        return vlb

    def resample_from_mf(self):
        """:570-584"""
        self.m = np.random.dirichlet(self.mf_pi)
        self.p = np.random.beta(self.mf_tau1, self.mf_tau0)
        for c1 in range(self.C):
            for c2 in range(self.C):
                self.weight_models[c1][c2].resample_from_mf()
        self.c = np.zeros(self.N, dtype=np.int64)
        for k in range(self.N):
            self.c[k] = int(np.random.choice(self.C, p=self.mf_m[k, :]))
        self.__dict__.pop("_mus_cache", None)

    def svi_step(self, augmented_data, minibatchfrac, stepsize):
        raise NotImplementedError()               # :586-587


class GaussianStochasticBlockModel(_GibbsSBM, _MeanFieldSBM, GaussianWeightedNetworkDistribution):
    """stochastic_block_model.py:589-592."""
    _weight_class = GaussianWeights
    _default_weight_hypers = {}

    # ---- mean-field expectations of the weights (:646-678) ----------------------------------------------------
    def _wm_cells(self, fn):
        return np.array([[fn(self.weight_models[c1][c2]) for c2 in range(self.C)] for c1 in range(self.C)])

    def mf_expected_mu(self):
        return self._mf_mix(self._wm_cells(lambda w: w.mf_expected_mu()))

    def mf_expected_mumuT(self):
        return self._mf_mix(self._wm_cells(lambda w: w.mf_expected_mumuT()))

    def mf_expected_Sigma_inv(self):
        return self._mf_mix(self._wm_cells(lambda w: w.mf_expected_Sigma_inv()))

    def mf_expected_logdet_Sigma(self):
        return self._mf_mix(self._wm_cells(lambda w: w.mf_expected_logdet_Sigma()))
