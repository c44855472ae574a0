"""
oracle/niw.py -- TEST INFRASTRUCTURE (see oracle/__init__.py).

Restatement of the pybasicbayes pieces the reference calls on the hot path.  pybasicbayes is a
git submodule (``.gitmodules:1-3``, ``deps/pybasicbayes``) that is absent from /root/reference and
whose pinned SHA is unknown; the functions below follow its published ~2015 code
("recollection" -- SURVEY.md App. C).  Call sites in the reference:

* ``Gaussian.resample(data)``       internals/weights.py:101, weights.py:161-162,408,413
* ``Gaussian.log_likelihood(x)``    stochastic_block_model.py:288-299, weights.py:325,329,462,468
* ``sample_discrete_from_log(lp)``  stochastic_block_model.py:302, adjacency.py:695, weights.py:488
"""
import numpy as np
import scipy.linalg

LOG2PI = np.log(2.0 * np.pi)


# ----------------------------------------------------------------------------------------------
# pybasicbayes.util.stats.sample_discrete_from_log   (recollection)
# ----------------------------------------------------------------------------------------------
def sample_discrete_from_log(p_log, u=None):
    """cum = cumsum(exp(p_log - max)); r = u * cum[-1]; return sum(r > cum).

    One uniform per call, inverse CDF with strict ``>``.  ``u`` is the injected uniform
    (``np.random.random()`` when None, which is what the reference consumes).
    """
    p_log = np.asarray(p_log, dtype=np.float64)
    cum = np.exp(p_log - p_log.max()).cumsum()
    if u is None:
        u = np.random.random()
    r = u * cum[-1]
    return int(np.sum(r > cum))


# ----------------------------------------------------------------------------------------------
# pybasicbayes.distributions.Gaussian (NIW prior)     (recollection)
# ----------------------------------------------------------------------------------------------
def niw_posterior(mu_0, sigma_0, kappa_0, nu_0, n, sum_x, sum_xxT):
    """Posterior NIW hyper-parameters from sufficient statistics (n, sum x, sum x x^T).

    pybasicbayes ``Gaussian._posterior_hypparams(n, xbar, sumsq)``:
      kappa_n = kappa_0 + n ; nu_n = nu_0 + n
      mu_n    = (kappa_0 mu_0 + n xbar) / kappa_n
      sigma_n = sigma_0 + S + kappa_0 n / kappa_n (xbar - mu_0)(xbar - mu_0)^T,   S = scatter about xbar
    With n == 0 the prior is returned unchanged.
    """
    mu_0 = np.asarray(mu_0, dtype=np.float64)
    sigma_0 = np.asarray(sigma_0, dtype=np.float64)
    n = float(n)
    if n <= 0:
        return mu_0.copy(), sigma_0.copy(), float(kappa_0), float(nu_0)
    sum_x = np.asarray(sum_x, dtype=np.float64)
    sum_xxT = np.asarray(sum_xxT, dtype=np.float64)
    xbar = sum_x / n
    S = sum_xxT - n * np.outer(xbar, xbar)
    kappa_n = kappa_0 + n
    nu_n = nu_0 + n
    mu_n = (kappa_0 * mu_0 + n * xbar) / kappa_n
    d = xbar - mu_0
    sigma_n = sigma_0 + S + kappa_0 * n / kappa_n * np.outer(d, d)
    return mu_n, sigma_n, float(kappa_n), float(nu_n)


def sample_invwishart(S, nu, rng=np.random):
    """pybasicbayes.util.stats.sample_invwishart (recollection): Bartlett/QR construction."""
    n = S.shape[0]
    chol = np.linalg.cholesky(S)
    if (nu <= 81 + n) and (nu == np.round(nu)):
        x = rng.randn(int(nu), n)
    else:
        x = np.diag(np.sqrt(np.atleast_1d(rng.chisquare(nu - np.arange(n)))))
        x[np.triu_indices_from(x, 1)] = rng.randn(n * (n - 1) // 2)
    R = np.linalg.qr(x, 'r')
    T = scipy.linalg.solve_triangular(R.T, chol.T, lower=True).T
    return T.dot(T.T)


def sample_niw(mu, lmbda, kappa, nu, rng=np.random):
    """pybasicbayes.util.stats.sample_niw (recollection): Sigma ~ IW(lmbda, nu); mu ~ N(mu, Sigma/kappa)."""
    sigma = sample_invwishart(lmbda, nu, rng)
    mu = rng.multivariate_normal(mu, sigma / kappa)
    return mu, sigma


def niw_resample(mu_0, sigma_0, kappa_0, nu_0, data=None, stats=None, rng=np.random):
    """``Gaussian.resample(data)``: posterior hyper-parameters then a NIW draw.

    ``data`` is an (n, B) array (the reference passes ``W[mask]``); alternatively pass
    ``stats=(n, sum_x, sum_xxT)``.
    """
    if stats is None:
        if data is None or len(data) == 0:
            stats = (0, None, None)
        else:
            data = np.asarray(data, dtype=np.float64)
            stats = (data.shape[0], data.sum(0), data.T.dot(data))
    mu_n, sigma_n, kappa_n, nu_n = niw_posterior(mu_0, sigma_0, kappa_0, nu_0, *stats)
    return sample_niw(mu_n, sigma_n, kappa_n, nu_n, rng)


def gaussian_log_likelihood(x, mu, sigma):
    """``Gaussian.log_likelihood(x)``: multivariate normal log-pdf for each row of x."""
    x = np.atleast_2d(np.asarray(x, dtype=np.float64))
    mu = np.asarray(mu, dtype=np.float64)
    B = mu.shape[0]
    L = np.linalg.cholesky(sigma)
    xs = scipy.linalg.solve_triangular(L, (x - mu).T, lower=True)
    return -0.5 * np.sum(xs ** 2, axis=0) - 0.5 * B * LOG2PI - np.log(np.diag(L)).sum()


def gaussian_loglik_from_stats(n, sum_x, sum_xxT, mu, sigma):
    """sum_i log N(x_i; mu, sigma) from (n, sum x, sum x x^T)  -- SURVEY.md App. A.2.

    = -1/2 [ n (B log 2pi + log|Sigma|) + tr(Sigma^-1 (Sxx - mu Sx^T - Sx mu^T + n mu mu^T)) ]
    """
    mu = np.asarray(mu, dtype=np.float64)
    B = mu.shape[0]
    Sinv = np.linalg.inv(sigma)
    logdet = np.linalg.slogdet(sigma)[1]
    M = sum_xxT - np.outer(mu, sum_x) - np.outer(sum_x, mu) + n * np.outer(mu, mu)
    return -0.5 * (n * (B * LOG2PI + logdet) + np.sum(Sinv * M))


def sample_niw_injected(mu_n, sigma_n, kappa_n, nu_n, chi2, z_off, z_mu):
    """NIW draw with the random numbers injected (chi-square branch of ``sample_invwishart`` above, i.e. the
    Bartlett factor x = diag(sqrt(chi2)) + strict-upper normals, then the same QR / triangular-solve steps);
    the mean uses the fixed map  mu = mu_n + chol(Sigma / kappa_n) z_mu  instead of numpy's SVD map
    (same distribution; parity on sample paths needs one fixed map -- DESIGN.md "Host draws")."""
    n = sigma_n.shape[0]
    chol = np.linalg.cholesky(sigma_n)
    x = np.diag(np.sqrt(np.atleast_1d(chi2)))
    x[np.triu_indices_from(x, 1)] = z_off
    R = np.linalg.qr(x, 'r')
    T = scipy.linalg.solve_triangular(R.T, chol.T, lower=True).T
    sigma = T.dot(T.T)
    mu = mu_n + np.linalg.cholesky(sigma / kappa_n).dot(z_mu)
    return mu, sigma
