"""Polya-gamma Gibbs sampling of the logistic eigenmodel with Gaussian weights (reference
examples/demo_logistic_eigenmodel_gibbs.py:17-90: N = 10, D = 2, B = 2, lambda fixed, 1000 sweeps; BASELINE configs[2] is
the adjacency half of this model at N = 16384, D = 8)."""
import _common
import numpy as np

from graphistician_b200 import GaussianWeightedEigenmodel


def demo(seed=None, N_samples=1000, N=10, D=2, B=2, verbose=True):
    _common.seed_everything(seed)
    eigenmodel_args = {"D": D, "p": 0.01, "sigma_F": 3 ** 2, "lmbda": np.ones(D)}
    true_model = GaussianWeightedEigenmodel(N=N, B=B, **eigenmodel_args)
    network = true_model.rvs()
    test_model = GaussianWeightedEigenmodel(N=N, B=B, **eigenmodel_args)
    lps = []
    for smpl in range(N_samples):
        test_model.resample(network)
        lps.append(test_model.log_probability(network))
        if verbose and smpl % 50 == 0:
            print("Iteration ", smpl, " LP: ", lps[-1])
    return dict(lps=np.array(lps), true_lp=true_model.log_probability(network), mu_w=np.array(test_model.mu_w),
                mu_w_true=np.array(true_model.mu_w), sd_w=np.sqrt(np.diag(test_model.Sigma_w)))


if __name__ == "__main__":
    _common.cli(demo, 1000)
