"""Gibbs sampling of the Gaussian stochastic block model (reference examples/demo_gaussian_sbm_gibbs.py:17-70: N = 50,
C = 5, B = 2, 100 sweeps; BASELINE configs[1] is this model at N = 4096, C = 16)."""
import _common
import numpy as np

from graphistician_b200 import GaussianStochasticBlockModel


def demo(seed=None, N_samples=100, N=50, C=5, B=2, verbose=True):
    _common.seed_everything(seed)
    sbm_args = {"C": C}
    true_model = GaussianStochasticBlockModel(N, B, **sbm_args)
    network = true_model.rvs()
    test_model = GaussianStochasticBlockModel(N, B, **sbm_args)
    lps = [test_model.log_probability(network)]
    n_blocks = []
    for smpl in range(N_samples):
        test_model.resample(network)
        lps.append(test_model.log_probability(network))
        n_blocks.append(len(np.unique(test_model.c)))
        if verbose:
            print("Iteration ", smpl, " LP: ", lps[-1])
    same_true = true_model.c[:, None] == true_model.c[None, :]
    same_test = test_model.c[:, None] == test_model.c[None, :]
    return dict(lps=np.array(lps), true_lp=true_model.log_probability(network), n_blocks=np.array(n_blocks),
                pair_agreement=float((same_true == same_test).mean()), c=test_model.c.copy(), c_true=true_model.c.copy())


if __name__ == "__main__":
    _common.cli(demo, 100)
