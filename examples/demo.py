"""The Aldous-Hoover graph-model walkthrough (reference examples/demo.py:10-293): draw Erdos-Renyi and SBM networks,
fit the SBM's labels and parameters by Gibbs sampling (``fit_network``), and run the Geweke successive-conditional
test of that sampler (``geweke_test``), comparing the marginals of z_0 and B_00 with their priors.  The reference
steps through the parts with raw_input() and spy plots; here every part returns the numbers its plot showed."""
import _common
import numpy as np
import scipy.special

from graphistician_b200.internals.graph_models import ErdosRenyiNetwork, StochasticBlockModel, sample_network, \
    fit_network, geweke_test, invariant_sbm_order


def sample_er(N=100, a0=0.75, b0=0.75, n_ex=3):
    """demo.py:10-30."""
    model = ErdosRenyiNetwork(rho=None, x=(a0, b0))
    out = []
    for ex in range(n_ex):
        (A, f, theta) = sample_network(model, N)
        out.append((float(theta), float(np.mean(A))))
    return out


def sample_sbm(N=100, R=5, n_ex=3):
    """demo.py:32-75: the block-sorted adjacency is what the reference shows; its block sizes are returned."""
    model = StochasticBlockModel(R, 0.5, 0.5, 0.75 * np.ones(R))
    out = []
    for ex in range(n_ex):
        (A, f, theta) = sample_network(model, N)
        zs = invariant_sbm_order(f, R)
        out.append((np.bincount(np.asarray(f, dtype=int), minlength=R), float(np.mean(np.asarray(A)[np.ix_(zs, zs)]))))
    return out


def gibbs_sbm(N=100, R=5, N_iter=50, verbose=False):
    """demo.py:77-150."""
    model = StochasticBlockModel(R, 0.5, 0.5, np.ones(R))
    (A, f, theta) = sample_network(model, N)
    lp_true = model.logpr(A, f, theta)
    (f_trace, theta_trace, lp_trace) = fit_network(np.asarray(A), model, N_iter=N_iter, verbose=verbose)
    z, zt = np.asarray(f_trace[-1]), np.asarray(f, dtype=int)
    agree = float(((z[:, None] == z[None, :]) == (zt[:, None] == zt[None, :])).mean())
    return dict(lp_true=lp_true, lp_trace=lp_trace, pair_agreement=agree)


def geweke_sbm_test(N=10, R=5, N_iter=500, verbose=False):
    """demo.py:152-245: histogram of z_0 against pi's prior mean and of B_00 against Beta(b1, b0)."""
    b1, b0 = 1.5, 0.5
    model = StochasticBlockModel(R, b0, b1, np.ones(R))
    (f_trace, theta_trace, lp_trace) = geweke_test(N, model, N_iter=N_iter, verbose=verbose)
    Z = np.array(f_trace)
    pZ0 = np.bincount(Z[:, 0].astype(int), minlength=R) / float(len(Z))
    b00s = np.array([th[0][0, 0] for th in theta_trace])
    return dict(pZ0=pZ0, pZ0_prior=np.ones(R) / R, b00_mean=float(b00s.mean()), b00_prior_mean=b1 / (b1 + b0),
                b00_var=float(b00s.var()), b00_prior_var=b1 * b0 / ((b1 + b0) ** 2 * (b1 + b0 + 1)), lp_trace=lp_trace)


def demo(seed=None, N_iter=50, verbose=True, geweke_iter=500):
    _common.seed_everything(seed)
    er = sample_er()
    sbm = sample_sbm()
    fit = gibbs_sbm(N_iter=N_iter)
    gw = geweke_sbm_test(N_iter=geweke_iter)
    if verbose:
        print("ER draws (rho, density): ", er)
        print("SBM draws (block sizes, density): ", sbm)
        print("fit: lp true %.2f, lp first %.2f, lp last %.2f, pair agreement %.3f"
              % (fit["lp_true"], fit["lp_trace"][0], fit["lp_trace"][-1], fit["pair_agreement"]))
        print("Geweke: p(z_0) ", gw["pZ0"], " B_00 mean %.3f (prior %.3f)" % (gw["b00_mean"], gw["b00_prior_mean"]))
    return dict(lp_trace=fit["lp_trace"], lp_true=fit["lp_true"], pair_agreement=fit["pair_agreement"],
                geweke_pZ0=gw["pZ0"], geweke_b00_mean=gw["b00_mean"], geweke_b00_prior_mean=gw["b00_prior_mean"],
                geweke_b00_var=gw["b00_var"], geweke_b00_prior_var=gw["b00_prior_var"])


if __name__ == "__main__":
    _common.cli(demo, 50, iters_name="N_iter")
