"""Gibbs sampling of the probit eigenmodel (reference examples/demo_eigenmodel_gibbs.py:17-72: N = 10, D = 2, p = 0.01,
sigma_F = 9, lambda ~ N(1, 0.001), 1000 sweeps)."""
import _common
import numpy as np

from graphistician_b200.eigenmodel import ProbitEigenmodel


def demo(seed=None, N_samples=1000, N=10, D=2, verbose=True):
    _common.seed_everything(seed)
    p, sigma_F, mu_lmbda, sigma_lmbda = 0.01, 3 ** 2, 1.0, 0.001
    true_model = ProbitEigenmodel(N=N, D=D, p=p, sigma_F=sigma_F, mu_lmbda=mu_lmbda, sigma_lmbda=sigma_lmbda)
    A = true_model.rvs()
    test_model = ProbitEigenmodel(N=N, D=D, p=p, sigma_F=sigma_F, mu_lmbda=mu_lmbda, sigma_lmbda=sigma_lmbda)
    lps = []
    for smpl in range(N_samples):
        test_model.resample(A)
        lps.append(test_model.log_probability(A))
        if verbose and smpl % 50 == 0:
            print("Iteration ", smpl, " LP: ", lps[-1])
    return dict(lps=np.array(lps), true_lp=true_model.log_probability(A), lmbda=test_model.lmbda.copy(),
                mu_0=test_model.mu_0)


if __name__ == "__main__":
    _common.cli(demo, 1000)
