"""Mean-field inference for the logistic eigenmodel with Gaussian weights (reference
examples/demo_logistic_eigenmodel_mf.py:17-115: 100 iterations; the bound it plots is get_vlb() + expected_log_likelihood)."""
import _common
import numpy as np

from graphistician_b200 import GaussianWeightedEigenmodel


def demo(seed=None, N_iters=100, N=10, D=2, B=2, verbose=True):
    _common.seed_everything(seed)
    eigenmodel_args = {"D": D, "p": 0.01, "sigma_F": 9.0, "lmbda": np.ones(D)}
    true_model = GaussianWeightedEigenmodel(N=N, B=B, **eigenmodel_args)
    network = true_model.rvs()
    test_model = GaussianWeightedEigenmodel(N=N, B=B, **eigenmodel_args)
    test_model.resample_from_mf()
    lps = [test_model.log_probability(network)]
    vlbs = [test_model.get_vlb() + test_model.expected_log_likelihood(network)]
    for itr in range(N_iters):
        test_model.meanfieldupdate(network)
        vlbs.append(test_model.get_vlb() + test_model.expected_log_likelihood(network))
        test_model.resample_from_mf()
        lps.append(test_model.log_probability(network))
        if verbose and itr % 10 == 0:
            print("Iteration ", itr, " VLB: ", vlbs[-1], " LP: ", lps[-1])
    return dict(lps=np.array(lps), vlbs=np.array(vlbs), mu_w=np.array(test_model.mu_w),
                E_mu_w=np.array(test_model.weight_dist.mf_expected_mu()))


if __name__ == "__main__":
    _common.cli(demo, 100, iters_name="N_iters")
