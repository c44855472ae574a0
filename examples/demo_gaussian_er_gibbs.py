"""Gibbs sampling of the Erdos-Renyi model with Gaussian weights (reference examples/demo_gaussian_er_gibbs.py:17-54;
BASELINE config 1: N = 10, B = 2, 100 sweeps)."""
import _common
import numpy as np

from graphistician_b200 import GaussianErdosRenyiFixedSparsity


def demo(seed=None, N_samples=100, N=10, B=2, verbose=True):
    _common.seed_everything(seed)
    true_model = GaussianErdosRenyiFixedSparsity(N, B)
    network = true_model.rvs()
    test_model = GaussianErdosRenyiFixedSparsity(N, B)
    lps = [test_model.log_probability(network)]
    for smpl in range(N_samples):
        test_model.resample(network)
        lps.append(test_model.log_probability(network))
        if verbose:
            print("Iteration ", smpl, " LP: ", lps[-1])
    return dict(lps=np.array(lps), true_lp=true_model.log_probability(network))


if __name__ == "__main__":
    _common.cli(demo, 100)
