"""Mean-field inference for the Erdos-Renyi model with Gaussian weights (reference examples/demo_gaussian_er_mf.py:17-56:
meanfieldupdate, then a draw from the variational posterior, 20 times)."""
import _common
import numpy as np

from graphistician_b200 import GaussianErdosRenyiFixedSparsity


def demo(seed=None, N_iters=20, N=10, B=2, verbose=True):
    _common.seed_everything(seed)
    true_model = GaussianErdosRenyiFixedSparsity(N, B)
    network = true_model.rvs()
    test_model = GaussianErdosRenyiFixedSparsity(N, B)
    lps = [test_model.log_probability(network)]
    vlbs = []
    for itr in range(N_iters):
        test_model.meanfieldupdate(network)
        vlbs.append(test_model.get_vlb())
        test_model.resample_from_mf()
        lps.append(test_model.log_probability(network))
        if verbose:
            print("Iteration ", itr, " LP: ", lps[-1])
    return dict(lps=np.array(lps), vlbs=np.array(vlbs))


if __name__ == "__main__":
    _common.cli(demo, 20, iters_name="N_iters")
