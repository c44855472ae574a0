"""Model comparison by held-out predictive likelihood (reference examples/demo_model_comparison.py:22-74: a network of
N + 1 = 31 nodes from the latent-distance x NIW model; the first 30 nodes train 3 x 2 factorised models by Gibbs
sampling, the last node's row and column are scored with approx_predictive_ll after every sweep; seed 1234)."""
import itertools

import _common
import numpy as np

from graphistician_b200.adjacency import LatentDistanceAdjacencyDistribution, SBMAdjacencyDistribution, \
    BetaBernoulliAdjacencyDistribution
from graphistician_b200.weights import NIWGaussianWeightDistribution, SBMGaussianWeightDistribution
from graphistician_b200.networks import FactorizedNetworkDistribution


def demo(seed=1234, N_samples=1000, N=30, verbose=True, M=100):
    _common.seed_everything(seed)
    N_test, B = 1, 1
    true_model = FactorizedNetworkDistribution(N + N_test, LatentDistanceAdjacencyDistribution, {},
                                               NIWGaussianWeightDistribution, {})
    Afull, Wfull = true_model.rvs()
    Afull = np.asarray(Afull, dtype=np.uint8)
    Atrain = np.ascontiguousarray(Afull[:N, :N])
    Wtrain = np.ascontiguousarray(Wfull[:N, :N, :])
    Atest_row = Afull[N:, :].ravel()
    Wtest_row = Wfull[N:, :, :].reshape((N + N_test, 1))
    Atest_col = Afull[:, N:].ravel()
    Wtest_col = Wfull[:, N:, :].reshape((N + N_test, 1))
    adj_models = [LatentDistanceAdjacencyDistribution, SBMAdjacencyDistribution, BetaBernoulliAdjacencyDistribution]
    weight_models = [NIWGaussianWeightDistribution, SBMGaussianWeightDistribution]
    traces = {}
    for adj_model, weight_model in itertools.product(adj_models, weight_models):
        test_model = FactorizedNetworkDistribution(N, adj_model, {}, weight_model, {})
        lps = [test_model.log_probability((Atrain, Wtrain))]
        plls = [test_model.approx_predictive_ll(Atest_row, Atest_col, Wtest_row, Wtest_col, M=M)]
        name = adj_model.__name__.replace("AdjacencyDistribution", "") + "-" + \
            weight_model.__name__.replace("GaussianWeightDistribution", "")
        for smpl in range(N_samples):
            test_model.resample((Atrain, Wtrain))
            lps.append(test_model.log_probability((Atrain, Wtrain)))
            plls.append(test_model.approx_predictive_ll(Atest_row, Atest_col, Wtest_row, Wtest_col, M=M))
        traces[name + "_lps"] = np.array(lps)
        traces[name + "_plls"] = np.array(plls)
        if verbose:
            print("A: ", adj_model.__name__, " W: ", weight_model.__name__, " mean PLL (second half): ",
                  np.mean(plls[len(plls) // 2:]))
    return traces


if __name__ == "__main__":
    _common.cli(demo, 1000, default_seed=1234)
