"""HMC-within-Gibbs for the latent distance model (reference examples/demo_distance_model_gibbs.py:17-84: N = 30 nodes on
two circles, dim = 2, 1000 sweeps, seed 1234; BASELINE configs[3] is this model at N = 16384)."""
import _common
import numpy as np

from graphistician_b200.adjacency import LatentDistanceAdjacencyDistribution


def circle_locations(N):
    """demo_distance_model_gibbs.py:37-42."""
    r = 1.5 + np.arange(N) // (N / 2.)
    th = np.linspace(0, 4 * np.pi, N, endpoint=False)
    return np.hstack(((r * np.cos(th))[:, None], (r * np.sin(th))[:, None]))


def distance_error(L, L_true):
    """RMS error of the pairwise distance matrix (invariant to the rotations / reflections the plot aligns away)."""
    d = lambda X: np.sqrt(((X[:, None, :] - X[None, :, :]) ** 2).sum(2))
    return float(np.sqrt(np.mean((d(L) - d(L_true)) ** 2)))


def demo(seed=1234, N_samples=1000, N=30, D=2, verbose=True):
    _common.seed_everything(seed)
    true_model = LatentDistanceAdjacencyDistribution(N=N, dim=D)
    true_model.L = circle_locations(N)
    A = true_model.rvs()
    test_model = LatentDistanceAdjacencyDistribution(N=N, dim=D)
    lps = [test_model.log_likelihood(A)]
    errs = [distance_error(test_model.L, true_model.L)]
    for smpl in range(N_samples):
        test_model.resample(A)
        lps.append(test_model.log_probability(A))
        errs.append(distance_error(test_model.L, true_model.L))
        if verbose and smpl % 50 == 0:
            print("Iteration ", smpl, " LP: ", lps[-1])
    return dict(lps=np.array(lps), true_lp=true_model.log_probability(A), distance_rmse=np.array(errs),
                L=test_model.L.copy(), step_sz=test_model._L_step_sz)


if __name__ == "__main__":
    _common.cli(demo, 1000, default_seed=1234)
