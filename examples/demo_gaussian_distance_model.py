"""Gaussian weights whose means fall off with latent distance (reference examples/demo_gaussian_distance_model.py:16-89:
complete graph on N = 30 nodes placed on two circles, W_ij ~ N(-||l_i - l_j||^2 + b, Sigma), b = -12, 1000 sweeps of
HMC on L + conjugate b / Sigma / eta updates; seed 1234)."""
import _common
import numpy as np

from graphistician_b200.adjacency import CompleteAdjacencyDistribution
from graphistician_b200.weights import LatentDistanceGaussianWeightDistribution
from graphistician_b200.networks import FactorizedNetworkDistribution

from demo_distance_model_gibbs import circle_locations, distance_error


def demo(seed=1234, N_samples=1000, N=30, D=2, verbose=True):
    _common.seed_everything(seed)
    kw = dict(dim=D, b=-12.)
    true_model = FactorizedNetworkDistribution(N, CompleteAdjacencyDistribution, {}, LatentDistanceGaussianWeightDistribution, kw)
    true_model.weights.L = circle_locations(N)
    A, W = true_model.rvs()
    test_model = FactorizedNetworkDistribution(N, CompleteAdjacencyDistribution, {}, LatentDistanceGaussianWeightDistribution, kw)
    lps = [test_model.log_probability((A, W))]
    lls = [test_model.weights.log_likelihood((A, W))]
    errs = [distance_error(test_model.weights.L, true_model.weights.L)]
    for smpl in range(N_samples):
        test_model.resample((A, W))
        lps.append(test_model.log_probability((A, W)))
        lls.append(test_model.weights.log_likelihood((A, W)))
        errs.append(distance_error(test_model.weights.L, true_model.weights.L))
        if verbose and smpl % 50 == 0:
            print("Iteration ", smpl, " LL: ", lls[-1], " b: ", test_model.weights.b, " eta: ", test_model.weights.eta)
    return dict(lps=np.array(lps), lls=np.array(lls), distance_rmse=np.array(errs), b=test_model.weights.b,
                sigma=float(test_model.weights.cov.sigma[0, 0]), eta=test_model.weights.eta)


if __name__ == "__main__":
    _common.cli(demo, 1000, default_seed=1234)
