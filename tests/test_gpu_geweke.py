"""Geweke ("getting it right") tests of the eigenmodel Gibbs samplers, ported from the reference's
test/test_logistic_eigenmodel_geweke.py and test/test_probit_eigenmodel_geweke.py (same loop: resample(A); A = rvs();
keep the sample; N = 5, D = 2, p = 0.1).  The reference only draws Q-Q plots; here the criterion is numeric: with the
chain's marginal equal to the prior, the batch-means z statistics of the first two moments of F[0, d], mu_0 and lmbda
must be unremarkable.  Every resample() runs on the GPU through the C ABI (a new network every iteration)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _batch_z(x, mean, nb=25):
    """z statistic of mean(x) against `mean`, standard error from non-overlapping batch means (autocorrelation-robust)."""
    x = np.asarray(x, dtype=float)
    n = len(x) // nb * nb
    bm = x[:n].reshape(nb, -1).mean(1)
    return (bm.mean() - mean) / (bm.std(ddof=1) / np.sqrt(nb))


def _run_chain(model, n_samples, burn):
    A = model.rvs()
    F0, mu0, lam = [], [], []
    for it in range(n_samples + burn):
        model.resample(A)
        A = model.rvs()
        if it >= burn:
            F0.append(model.F[0].copy()); mu0.append(model.mu_0); lam.append(model.lmbda.copy())
    return np.array(F0), np.array(mu0), np.array(lam)


def _check(name, x, mean, var, zmax=4.5):
    z1 = _batch_z(x, mean)
    z2 = _batch_z((x - mean) ** 2, var)
    assert abs(z1) < zmax, "%s: mean %.3f vs prior %.3f (z = %.2f)" % (name, x.mean(), mean, z1)
    assert abs(z2) < zmax, "%s: variance %.3f vs prior %.3f (z = %.2f)" % (name, x.var(), var, z2)


@pytest.mark.parametrize("cls_name", ["LogisticEigenmodel", "ProbitEigenmodel"])
def test_geweke_eigenmodel(cls_name):
    from graphistician_b200 import eigenmodel
    cls = getattr(eigenmodel, cls_name)
    np.random.seed(20261104)
    N, D, p = 5, 2, 0.1
    model = cls(N=N, D=D, p=p, sigma_mu0=1.0, sigma_F=1.0, sigma_lmbda=1.0, dtype="f64")
    assert not model.lmbda_given
    F0, mu0, lam = _run_chain(model, 4000, 200)
    for d in range(D):
        _check("F[0,%d]" % d, F0[:, d], 0.0, model.sigma_F)
        _check("lmbda[%d]" % d, lam[:, d], model.mu_lmbda[d], model.sigma_lmbda)
    _check("mu_0", mu0, model.mu_mu_0, model.sigma_mu0)


def test_geweke_detects_a_broken_sampler():
    """the criterion has power: a chain whose F conditional ignores the prior precision fails it"""
    from graphistician_b200.eigenmodel import LogisticEigenmodel
    np.random.seed(7)
    model = LogisticEigenmodel(N=5, D=2, p=0.1, sigma_mu0=1.0, sigma_F=1.0, sigma_lmbda=1.0, dtype="f64")
    model.sigma_F = 4.0                                  # sampler now targets a different prior than the one checked
    F0, _, _ = _run_chain(model, 1500, 100)
    z2 = _batch_z(F0[:, 0] ** 2, 1.0)
    assert abs(z2) > 4.5
