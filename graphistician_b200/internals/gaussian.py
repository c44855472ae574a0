"""Host-side Normal-inverse-Wishart Gaussian used for the per-cell weight models.

Plays the role of ``pybasicbayes.distributions.Gaussian`` as graphistician uses it
(internals/weights.py:62-101, weights.py:118-162,403-413): hyper-parameters ``mu_0, sigma_0, kappa_0, nu_0``,
state ``mu, sigma``, ``resample(data)`` = NIW posterior draw, ``log_likelihood(x)`` = row-wise normal log-pdf.
The O(N^2) part -- gathering ``W[A==1]`` and reducing it to (n, sum w, sum w w^T) -- is done on the device
(gk_block_stats); this class consumes those statistics via ``resample_from_stats``.  Draws come from the
global ``np.random`` stream, as the reference's do (SURVEY.md 8b "Threading / RNG").
"""
import numpy as np

_LOG2PI = 1.8378770664093453


def _chol_solve_lower(L, Bm):
    """Solve L X = Bm for lower-triangular L (small matrices; forward substitution)."""
    n = L.shape[0]
    X = np.array(Bm, dtype=np.float64, copy=True)
    for i in range(n):
        if i:
            X[i] -= L[i, :i].dot(X[:i])
        X[i] /= L[i, i]
    return X


def invwishart_draw(S, nu, rng=np.random):
    """Sigma ~ IW(S, nu) by the Bartlett construction pybasicbayes uses: for small integer nu a (nu x n) normal
    matrix, otherwise chi-square diagonal + normal upper triangle; R from its QR; T = (R^-T chol(S)^T)^T; T T^T."""
    n = S.shape[0]
    chol = np.linalg.cholesky(S)
    if nu <= 81 + n and nu == round(nu):
        x = rng.randn(int(nu), n)
    else:
        x = np.diag(np.sqrt(np.atleast_1d(rng.chisquare(nu - np.arange(n)))))
        x[np.triu_indices_from(x, 1)] = rng.randn(n * (n - 1) // 2)
    R = np.linalg.qr(x, "r")
    T = _chol_solve_lower(R.T, chol.T).T
    return T.dot(T.T)


class Gaussian(object):
    """Multivariate normal with a NIW prior on (mu, sigma)."""

    def __init__(self, mu=None, sigma=None, mu_0=None, sigma_0=None, kappa_0=None, nu_0=None):
        self.mu_0 = None if mu_0 is None else np.array(mu_0, dtype=np.float64)
        self.sigma_0 = None if sigma_0 is None else np.array(sigma_0, dtype=np.float64)
        self.kappa_0, self.nu_0 = kappa_0, nu_0
        self.mu = None if mu is None else np.array(mu, dtype=np.float64)
        self.sigma = None if sigma is None else np.array(sigma, dtype=np.float64)
        if self.mu is None and self.mu_0 is not None:
            self.resample()                       # prior draw, as pybasicbayes does on construction

    @property
    def D(self):
        return (self.mu if self.mu is not None else self.mu_0).shape[0]

    @property
    def sigma_chol(self):
        return np.linalg.cholesky(self.sigma)

    @property
    def hypparams(self):
        return dict(mu_0=self.mu_0, sigma_0=self.sigma_0, kappa_0=self.kappa_0, nu_0=self.nu_0)

    # ---- NIW posterior ---------------------------------------------------------------------------
    def posterior_hypparams(self, n, sum_x, sum_xxT):
        """(mu_n, sigma_n, kappa_n, nu_n) from sufficient statistics; the prior when n == 0."""
        if n <= 0:
            return self.mu_0, self.sigma_0, self.kappa_0, self.nu_0
        n = float(n)
        xbar = np.asarray(sum_x, dtype=np.float64) / n
        scatter = np.asarray(sum_xxT, dtype=np.float64) - n * np.outer(xbar, xbar)
        kappa_n = self.kappa_0 + n
        dev = xbar - self.mu_0
        mu_n = (self.kappa_0 * self.mu_0 + n * xbar) / kappa_n
        sigma_n = self.sigma_0 + scatter + (self.kappa_0 * n / kappa_n) * np.outer(dev, dev)
        return mu_n, 0.5 * (sigma_n + sigma_n.T), kappa_n, self.nu_0 + n

    def resample_from_stats(self, n, sum_x, sum_xxT, rng=np.random):
        mu_n, sigma_n, kappa_n, nu_n = self.posterior_hypparams(n, sum_x, sum_xxT)
        self.sigma = invwishart_draw(sigma_n, nu_n, rng)
        self.mu = rng.multivariate_normal(mu_n, self.sigma / kappa_n)
        return self

    def resample(self, data=None, rng=np.random):
        """``Gaussian.resample(data)`` with data an (n, D) array (or empty -> prior draw)."""
        if data is None or len(data) == 0:
            return self.resample_from_stats(0, None, None, rng)
        data = np.asarray(data, dtype=np.float64).reshape(-1, self.D)
        return self.resample_from_stats(data.shape[0], data.sum(0), data.T.dot(data), rng)

    # ---- mean field: natural parameters of the NIW (pybasicbayes Gaussian, natural-parameter form) --------------
    @staticmethod
    def _standard_to_natural(mu, sigma, kappa, nu):
        D = sigma.shape[0]
        out = np.zeros((D + 2, D + 2))
        out[:D, :D] = sigma + kappa * np.outer(mu, mu)
        out[:D, -2] = out[-2, :D] = kappa * mu
        out[-2, -2] = kappa
        out[-1, -1] = nu + 2 + D
        return out

    @staticmethod
    def _natural_to_standard(nat):
        D = nat.shape[0] - 2
        A, b, c, d = nat[:D, :D], nat[:D, -2], nat[-2, -2], nat[-1, -1]
        return b / c, A - np.outer(b, b) / c, c, d - 2 - D

    @property
    def natural_hypparam(self):
        return self._standard_to_natural(self.mu_0, self.sigma_0, self.kappa_0, self.nu_0)

    @property
    def mf_natural_hypparam(self):
        if getattr(self, "_mf_nat", None) is None:
            self._mf_nat = self.natural_hypparam          # q starts at the prior
        return self._mf_nat

    @mf_natural_hypparam.setter
    def mf_natural_hypparam(self, nat):
        self._mf_nat = np.array(nat, dtype=np.float64)

    @property
    def mu_mf(self):
        return self._natural_to_standard(self.mf_natural_hypparam)[0]

    @property
    def sigma_mf(self):
        return self._natural_to_standard(self.mf_natural_hypparam)[1]

    @property
    def kappa_mf(self):
        return self._natural_to_standard(self.mf_natural_hypparam)[2]

    @property
    def nu_mf(self):
        return self._natural_to_standard(self.mf_natural_hypparam)[3]

    @staticmethod
    def _statistics_matrix(n, sum_x, sum_xxT):
        """layout of pybasicbayes' `_get_statistics` / the reference's `_get_weighted_statistics`
        (internals/weights.py:158-171)"""
        D = len(sum_x)
        out = np.zeros((D + 2, D + 2))
        out[:D, :D] = sum_xxT
        out[-2, :D] = out[:D, -2] = sum_x
        out[-2, -2] = out[-1, -1] = n
        return out

    def _loglmbdatilde(self):
        """E_q[log |Sigma^-1|]"""
        from scipy.special import psi
        D, sigma_mf, nu_mf = self.D, self.sigma_mf, self.nu_mf
        chol = np.linalg.cholesky(sigma_mf)
        return -2 * np.log(np.diag(chol)).sum() + D * np.log(2) + psi((nu_mf - np.arange(D)) / 2.0).sum()

    def get_vlb(self):
        """NIW part of the variational bound, E_q[ln p(mu, Sigma)] + H[q(mu, Sigma)] = -KL(q || prior) -- what
        pybasicbayes' ``Gaussian.get_vlb`` returns (Bishop 10.74 and 10.77 without the sum over components); the
        reference calls it per block pair at stochastic_block_model.py:564-566."""
        from scipy.special import gammaln, psi

        def iw_log_partition(sigma, nu):
            D = sigma.shape[0]
            half_logdet = np.log(np.diag(np.linalg.cholesky(sigma))).sum()
            return -(nu * half_logdet - (nu * D / 2.0 * np.log(2) + D * (D - 1) / 4.0 * np.log(np.pi)
                                         + gammaln((nu - np.arange(D)) / 2.0).sum()))

        def iw_entropy(sigma, nu):
            D = sigma.shape[0]
            half_logdet = np.log(np.diag(np.linalg.cholesky(sigma))).sum()
            E_logdet_lmbda = psi((nu - np.arange(D)) / 2.0).sum() + D * np.log(2) - 2 * half_logdet
            return iw_log_partition(sigma, nu) - (nu - D - 1) / 2.0 * E_logdet_lmbda + nu * D / 2.0

        D = self.D
        mu_mf, sigma_mf, kappa_mf, nu_mf = self._natural_to_standard(self.mf_natural_hypparam)
        sigma_mf = 0.5 * (sigma_mf + sigma_mf.T)
        llt = self._loglmbdatilde()
        q_entropy = -0.5 * (llt + D * (np.log(kappa_mf / (2 * np.pi)) - 1)) + iw_entropy(sigma_mf, nu_mf)
        dmu = mu_mf - self.mu_0
        p_avgengy = 0.5 * (D * np.log(self.kappa_0 / (2 * np.pi)) + llt - D * self.kappa_0 / kappa_mf
                           - self.kappa_0 * nu_mf * dmu.dot(np.linalg.solve(sigma_mf, dmu))) \
            - iw_log_partition(self.sigma_0, self.nu_0) + (self.nu_0 - D - 1) / 2.0 * llt \
            - 0.5 * nu_mf * np.trace(np.linalg.solve(sigma_mf, self.sigma_0))
        return float(p_avgengy + q_entropy)

    def _resample_from_mf(self, rng=np.random):
        mu_n, sigma_n, kappa_n, nu_n = self._natural_to_standard(self.mf_natural_hypparam)
        self.sigma = invwishart_draw(0.5 * (sigma_n + sigma_n.T), nu_n, rng)
        self.mu = rng.multivariate_normal(mu_n, self.sigma / kappa_n)
        return self

    # ---- likelihood / sampling ---------------------------------------------------------------------
    def log_likelihood(self, x):
        x = np.asarray(x, dtype=np.float64).reshape(-1, self.D)
        L = np.linalg.cholesky(self.sigma)
        r = _chol_solve_lower(L, (x - self.mu).T)
        return -0.5 * (r * r).sum(0) - 0.5 * self.D * _LOG2PI - np.log(np.diag(L)).sum()

    def log_likelihood_from_stats(self, n, sum_x, sum_xxT):
        """sum_i log N(x_i; mu, sigma) from (n, sum x, sum x x^T)."""
        Sinv = np.linalg.inv(self.sigma)
        logdet = np.linalg.slogdet(self.sigma)[1]
        M = sum_xxT - np.outer(self.mu, sum_x) - np.outer(sum_x, self.mu) + n * np.outer(self.mu, self.mu)
        return -0.5 * (n * (self.D * _LOG2PI + logdet) + np.sum(Sinv * M))

    def rvs(self, size=None):
        n = 1 if size is None else int(np.prod(size))
        out = self.mu + np.random.randn(n, self.D).dot(self.sigma_chol.T)
        return out[0] if size is None else out


# ------------------------------------------------------------------------------------------------------
# batched NIW draws for the K x K weight models of an SBM (one vectorised call instead of K^2 object calls)
# ------------------------------------------------------------------------------------------------------
def niw_posterior_batched(mu_0, sigma_0, kappa_0, nu_0, n, sum_x, sum_xxT):
    """Posterior hyper-parameters for a stack of cells.  n [...], sum_x [..., B], sum_xxT [..., B, B];
    cells with n == 0 keep the prior.  Same formulas as ``Gaussian.posterior_hypparams``."""
    n = np.asarray(n, dtype=np.float64)
    nn = np.maximum(n, 1.0)[..., None]
    xbar = np.where(n[..., None] > 0, sum_x / nn, 0.0)
    scatter = sum_xxT - n[..., None, None] * xbar[..., :, None] * xbar[..., None, :]
    kappa_n = kappa_0 + n
    dev = xbar - mu_0
    mu_n = (kappa_0 * mu_0 + n[..., None] * xbar) / kappa_n[..., None]
    sigma_n = sigma_0 + np.where(n[..., None, None] > 0, scatter, 0.0) \
        + (kappa_0 * n / kappa_n)[..., None, None] * dev[..., :, None] * dev[..., None, :]
    return mu_n, 0.5 * (sigma_n + np.swapaxes(sigma_n, -1, -2)), kappa_n, nu_0 + n


def niw_draw_batched(mu_n, sigma_n, kappa_n, nu_n, chi2, z_off, z_mu):
    """(mu, Sigma) for a stack of cells from injected draws.

    Bartlett construction (the chi-square branch of pybasicbayes ``sample_invwishart``): X upper triangular with
    X[i,i] = sqrt(chi2[..., i]), chi2[..., i] ~ chi^2(nu_n - i), strict upper triangle = z_off (row-major);
    T = chol(sigma_n) X^-1;  Sigma = T T^T;  mu = mu_n + chol(Sigma / kappa_n) z_mu.
    """
    B = mu_n.shape[-1]
    X = np.zeros(sigma_n.shape)
    idx = np.arange(B)
    X[..., idx, idx] = np.sqrt(chi2)
    iu = np.triu_indices(B, 1)
    if len(iu[0]):
        X[..., iu[0], iu[1]] = z_off
    chol = np.linalg.cholesky(sigma_n)
    T = np.swapaxes(np.linalg.solve(np.swapaxes(X, -1, -2), np.swapaxes(chol, -1, -2)), -1, -2)
    sigma = T @ np.swapaxes(T, -1, -2)
    sigma = 0.5 * (sigma + np.swapaxes(sigma, -1, -2))
    Lm = np.linalg.cholesky(sigma / kappa_n[..., None, None])
    mu = mu_n + np.einsum("...ab,...b->...a", Lm, z_mu)
    return mu, sigma


def niw_resample_batched(hyp, n, sum_x, sum_xxT, rng=np.random):
    """Draw order (one vectorised rng call each): chi-squares [cells, B] -> off-diagonal normals -> mean normals."""
    mu_n, sigma_n, kappa_n, nu_n = niw_posterior_batched(hyp["mu_0"], hyp["sigma_0"], hyp["kappa_0"], hyp["nu_0"],
                                                         n, sum_x, sum_xxT)
    B = mu_n.shape[-1]
    chi2 = rng.chisquare(nu_n[..., None] - np.arange(B))
    z_off = rng.standard_normal(n.shape + (B * (B - 1) // 2,))
    z_mu = rng.standard_normal(n.shape + (B,))
    return niw_draw_batched(mu_n, sigma_n, kappa_n, nu_n, chi2, z_off, z_mu)


class GaussianFixedMean(object):
    """pybasicbayes.distributions.GaussianFixedMean as weights.py:638,863-867 uses it: Sigma ~ IW(lmbda_0, nu_0) around a
    fixed mean; ``resample_from_stats(n, S)`` draws from IW(lmbda_0 + S, nu_0 + n), S = sum (x - mu)(x - mu)^T."""

    def __init__(self, mu=None, sigma=None, nu_0=None, lmbda_0=None):
        self.mu = None if mu is None else np.array(mu, dtype=np.float64)
        self.sigma = None if sigma is None else np.array(sigma, dtype=np.float64)
        self.nu_0 = nu_0
        self.lmbda_0 = None if lmbda_0 is None else np.array(lmbda_0, dtype=np.float64)
        if sigma is None and nu_0 is not None and lmbda_0 is not None:
            self.resample()

    @property
    def D(self):
        return self.mu.shape[0]

    @property
    def sigma_chol(self):
        return np.linalg.cholesky(self.sigma)

    def resample_from_stats(self, n, S, rng=np.random):
        self.sigma = invwishart_draw(self.lmbda_0 + S, self.nu_0 + n, rng)
        return self

    def resample(self, data=None, rng=np.random):
        if data is None or len(data) == 0:
            return self.resample_from_stats(0, 0.0, rng)
        d = np.asarray(data, dtype=np.float64).reshape(-1, self.D) - self.mu
        return self.resample_from_stats(d.shape[0], d.T.dot(d), rng)
