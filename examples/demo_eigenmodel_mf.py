"""Mean-field variational inference for the probit eigenmodel (reference examples/demo_eigenmodel_mf.py:17-104: N = 10,
D = 2, lambda fixed at ones, 1000 coordinate-ascent iterations from a draw of the variational prior, seed 11223344)."""
import _common
import numpy as np

from graphistician_b200.eigenmodel import ProbitEigenmodel


def demo(seed=11223344, N_iters=1000, N=10, D=2, verbose=True):
    _common.seed_everything(seed)
    p, sigma_F, lmbda = 0.01, 9.0, np.ones(D)
    true_model = ProbitEigenmodel(N=N, D=D, p=p, sigma_F=sigma_F, lmbda=lmbda)
    true_model.lmbda = np.ones((D,))
    A = true_model.rvs()
    test_model = ProbitEigenmodel(N=N, D=D, p=p, sigma_F=sigma_F, lmbda=lmbda)
    test_model.resample_from_mf()
    lps, vlbs = [], []
    for itr in range(N_iters):
        test_model.meanfieldupdate(A)
        vlbs.append(test_model.get_vlb())
        test_model.resample_from_mf()
        lps.append(test_model.log_probability(A))
        if verbose and itr % 50 == 0:
            print("Iteration ", itr, " VLB: ", vlbs[-1])
    return dict(lps=np.array(lps), vlbs=np.array(vlbs), true_lp=true_model.log_probability(A))


if __name__ == "__main__":
    _common.cli(demo, 1000, default_seed=11223344, iters_name="N_iters")
