"""Annealed importance sampling estimates of the marginal likelihood of an Erdos-Renyi and a stochastic block model on
a network drawn from the SBM (reference examples/demo_ais.py:10-127: N = 16, R = 3, 5 particles, 100 temperatures,
2 Gibbs steps per temperature; BASELINE configs[4] is this driver at N = 65536)."""
import _common
import numpy as np

from graphistician_b200.internals.graph_models import ErdosRenyiNetwork, StochasticBlockModel, sample_network, \
    invariant_sbm_order, estimate_marginal_likelihood, estimate_marginal_likelihood_batched


def demo(seed=None, N_samples=5, N=16, R=3, B=100, steps_per_B=2, verbose=True, batched=False):
    """``batched``: the particles run co-resident on the device over the bit-packed matrix (gk_ais_*, configs[4]'s form)
    instead of one after another."""
    _common.seed_everything(seed)
    estimate = estimate_marginal_likelihood_batched if batched else estimate_marginal_likelihood
    er_model = ErdosRenyiNetwork(x=(1.0, 1.0))
    sbm_model = StochasticBlockModel(R=R, b0=1.0, b1=1.0, alpha0=10 * np.ones(R))
    (A_true, f_true, theta_true) = sample_network(sbm_model, N)
    order = invariant_sbm_order(f_true, R)
    A_true = np.ascontiguousarray(np.asarray(A_true)[np.ix_(order, order)])
    unif_marg_lkhd = N ** 2 * np.log(0.5)
    nnz_A = float(A_true.sum())
    N_conns = A_true.size
    eb_rho = nnz_A / N_conns
    eb_er_marg_lkhd = nnz_A * np.log(eb_rho) + (N_conns - nnz_A) * np.log(1 - eb_rho)
    _, f0, theta0 = sample_network(er_model, N)
    est_er, w_er = estimate(A_true, f0, theta0, er_model, N_samples=N_samples, B=B,
                                                steps_per_B=steps_per_B, return_weights=True)
    _, f0, theta0 = sample_network(sbm_model, N)
    est_sbm, w_sbm = estimate(A_true, f0, theta0, sbm_model, N_samples=N_samples, B=B,
                                                  steps_per_B=steps_per_B, return_weights=True)
    if verbose:
        print("Empirical Bayes ER Marg Lkhd: ", eb_er_marg_lkhd)
        print("Estimated Bayesian ER Marg Lkhd: ", est_er)
        print("Estimated SBM Marg Lkhd: ", est_sbm)
    return dict(uniform=unif_marg_lkhd, empirical_bayes_er=eb_er_marg_lkhd, ais_er=est_er, ais_sbm=est_sbm,
                log_weights_er=w_er, log_weights_sbm=w_sbm,
                relative=np.array([eb_er_marg_lkhd, est_er, est_sbm]) - unif_marg_lkhd)


if __name__ == "__main__":
    _common.cli(demo, 5)
