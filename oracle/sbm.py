"""
oracle/sbm.py -- TEST INFRASTRUCTURE (see oracle/__init__.py).

CPU restatement of the stochastic-block-model Gibbs hot path:

* block-pair sufficient statistics     adjacency.py:620-648, weights.py:399-413,
                                       stochastic_block_model.py:221-250, internals/weights.py:88-101
* sequential block-assignment sweep    stochastic_block_model.py:252-302 (old API, shared c),
                                       adjacency.py:650-695, weights.py:415-488 (new API),
                                       internals/graph_models.py:261-297 (AIS, beta-tempered)
* Beta / Dirichlet / NIW parameter draws that follow the statistics.

Two tiers:  ``*_loops`` functions are literal loop transcriptions (T0; small N only);
the un-suffixed functions are vectorised NumPy with the same arithmetic (T1; used at scale and as
the timed CPU baseline).  tests/test_oracle_sbm.py holds T1 == T0.

Random draws are *injected*: every stochastic function takes the uniforms it consumes, exactly one
per node for the sweep (pybasicbayes ``sample_discrete_from_log`` draws one ``np.random.random()``).
"""
import numpy as np

from .niw import (sample_discrete_from_log, gaussian_log_likelihood, niw_resample, LOG2PI)

# Diagonal handling of the assignment sweep (which reference function is being restated)
SELF_OLD = "old"          # stochastic_block_model.py:293-299: Bernoulli(p[k,k]) + A11*logN(W11; k,k)
SELF_SKIP = "skip"        # adjacency.py:671-674 / weights.py:460-461 with special_case_self_conns=True
SELF_NEW = "new"          # adjacency.py:675-677 (special_case_self_conns=False); weights.py:_evaluate_lkhd
SELF_AIS = "ais"          # internals/graph_models.py:275-293: diagonal counted in both directions


# ----------------------------------------------------------------------------------------------
# Block-pair sufficient statistics (K1)
# ----------------------------------------------------------------------------------------------
def block_pair_stats_loops(A, W, c, K):
    """Literal mask-per-cell transcription.

    adjacency.py:622-638 (``_get_mask`` + ``A[mask].sum()``), weights.py:308-314,403-408
    (``W[mask]``), with the diagonal split out as adjacency.py:641-648 / weights.py:411-413 do.

    Returns dict: n_edge[K,K], n_pair[K,K] (off-diagonal pairs only), sum_w[K,K,B], sum_ww[K,K,B,B],
    and per-block diagonal stats d_edge[K], d_cnt[K], d_sum_w[K,B], d_sum_ww[K,B,B].
    """
    N = A.shape[0]
    Ab = A.astype(bool)
    B = 0 if W is None else W.shape[2]
    offd = ~np.eye(N, dtype=bool)
    out = dict(n_edge=np.zeros((K, K), np.int64), n_pair=np.zeros((K, K), np.int64),
               sum_w=np.zeros((K, K, B)), sum_ww=np.zeros((K, K, B, B)),
               d_edge=np.zeros(K, np.int64), d_cnt=np.zeros(K, np.int64),
               d_sum_w=np.zeros((K, B)), d_sum_ww=np.zeros((K, B, B)))
    for c1 in range(K):
        for c2 in range(K):
            mask = ((c == c1)[:, None] * (c == c2)[None, :]) & offd
            out["n_edge"][c1, c2] = A[mask].sum()
            out["n_pair"][c1, c2] = mask.sum()
            if B:
                Wm = W[mask & Ab].astype(np.float64)
                out["sum_w"][c1, c2] = Wm.sum(0)
                out["sum_ww"][c1, c2] = Wm.T.dot(Wm)
    for k in range(K):
        idx = np.where(c == k)[0]
        out["d_cnt"][k] = idx.size
        out["d_edge"][k] = A[idx, idx].sum()
        if B:
            Wd = W[idx, idx][Ab[idx, idx]].astype(np.float64)
            out["d_sum_w"][k] = Wd.sum(0)
            out["d_sum_ww"][k] = Wd.T.dot(Wd)
    return out


def block_pair_stats(A, W, c, K, row_chunk=1024):
    """Vectorised K1: same outputs as ``block_pair_stats_loops`` (bincount over a fused cell index)."""
    N = A.shape[0]
    B = 0 if W is None else W.shape[2]
    c = np.asarray(c, dtype=np.int64)
    out = dict(n_edge=np.zeros(K * K, np.int64), n_pair=np.zeros(K * K, np.int64),
               sum_w=np.zeros((K * K, B)), sum_ww=np.zeros((K * K, B, B)),
               d_edge=np.zeros(K, np.int64), d_cnt=np.bincount(c, minlength=K).astype(np.int64),
               d_sum_w=np.zeros((K, B)), d_sum_ww=np.zeros((K, B, B)))
    for r0 in range(0, N, row_chunk):
        r1 = min(N, r0 + row_chunk)
        Ar = (A[r0:r1] != 0)
        cell = (c[r0:r1, None] * K + c[None, :])
        offd = np.ones_like(Ar)
        offd[np.arange(r1 - r0), np.arange(r0, r1)] = False
        e = Ar & offd
        out["n_edge"] += np.bincount(cell[e], minlength=K * K)
        out["n_pair"] += np.bincount(cell[offd], minlength=K * K)
        if B:
            ce = cell[e]
            We = W[r0:r1][e].astype(np.float64)
            for b in range(B):
                out["sum_w"][:, b] += np.bincount(ce, weights=We[:, b], minlength=K * K)
                for b2 in range(b, B):
                    v = np.bincount(ce, weights=We[:, b] * We[:, b2], minlength=K * K)
                    out["sum_ww"][:, b, b2] += v
                    if b2 != b:
                        out["sum_ww"][:, b2, b] += v
    d = np.arange(N)
    Ad = (A[d, d] != 0)
    out["d_edge"] = np.bincount(c[Ad], minlength=K).astype(np.int64)
    if B:
        Wd = W[d, d][Ad].astype(np.float64)
        cd = c[Ad]
        for b in range(B):
            out["d_sum_w"][:, b] = np.bincount(cd, weights=Wd[:, b], minlength=K)
            for b2 in range(B):
                out["d_sum_ww"][:, b, b2] = np.bincount(cd, weights=Wd[:, b] * Wd[:, b2], minlength=K)
    out["n_edge"] = out["n_edge"].reshape(K, K)
    out["n_pair"] = out["n_pair"].reshape(K, K)
    out["sum_w"] = out["sum_w"].reshape(K, K, B)
    out["sum_ww"] = out["sum_ww"].reshape(K, K, B, B)
    return out


def cell_stats_with_diag(st, include_diag):
    """Compose per-cell (n_edge, n_noedge, sum_w, sum_ww) for either API generation.

    Old API (stochastic_block_model.py:223-250) keeps the diagonal inside cell (k,k);
    new API (adjacency.py:626-627, weights.py:311-312) excludes it when special-casing.
    """
    n_edge = st["n_edge"].copy()
    n_pair = st["n_pair"].copy()
    sum_w = st["sum_w"].copy()
    sum_ww = st["sum_ww"].copy()
    if include_diag:
        K = n_edge.shape[0]
        k = np.arange(K)
        n_edge[k, k] += st["d_edge"]
        n_pair[k, k] += st["d_cnt"]
        sum_w[k, k] += st["d_sum_w"]
        sum_ww[k, k] += st["d_sum_ww"]
    return n_edge, n_pair - n_edge, sum_w, sum_ww


# ----------------------------------------------------------------------------------------------
# Parameter draws that follow K1 (host-side in the product as well; O(K^2))
# ----------------------------------------------------------------------------------------------
def resample_p(n_edge, n_noedge, tau1, tau0, beta=1.0, rng=np.random):
    """p[c1,c2] ~ Beta(tau1 + n_conns, tau0 + n_noconns), row-major draw order.

    stochastic_block_model.py:236-250; adjacency.py:630-638; tempered: internals/graph_models.py:361-369.
    """
    K = n_edge.shape[0]
    p = np.zeros((K, K))
    for c1 in range(K):
        for c2 in range(K):
            p[c1, c2] = rng.beta(tau1 + beta * n_edge[c1, c2], tau0 + beta * n_noedge[c1, c2])
    return p


def resample_m(c, pi, rng=np.random):
    """m ~ Dir(pi + bincount(c)).  stochastic_block_model.py:304-309, adjacency.py:697-702, weights.py:493-498."""
    return rng.dirichlet(pi + np.bincount(c, minlength=len(pi)))


def resample_gaussians(n_edge, sum_w, sum_ww, hypers, rng=np.random):
    """Per-cell NIW draws in row-major order.  stochastic_block_model.py:221-226 + internals/weights.py:88-101;
    weights.py:403-408.  ``hypers`` = (mu_0, sigma_0, kappa_0, nu_0)."""
    K = n_edge.shape[0]
    B = sum_w.shape[2]
    mus = np.zeros((K, K, B))
    sigmas = np.zeros((K, K, B, B))
    for c1 in range(K):
        for c2 in range(K):
            mus[c1, c2], sigmas[c1, c2] = niw_resample(
                *hypers, stats=(n_edge[c1, c2], sum_w[c1, c2], sum_ww[c1, c2]), rng=rng)
    return mus, sigmas


# ----------------------------------------------------------------------------------------------
# Sequential assignment sweep (K2): literal transcriptions
# ----------------------------------------------------------------------------------------------
def _bern_logp(p, x):
    """internals/distributions.py:59-67 (``Bernoulli.log_probability`` incl. ``nan_to_num``)."""
    with np.errstate(divide="ignore", invalid="ignore"):
        lp = x * np.log(p) + (1 - x) * np.log(1.0 - p)
    return np.nan_to_num(lp)


def gibbs_sbm_resample_c_loops(A, W, c, m, p, mus, sigmas, u):
    """stochastic_block_model.py:252-302, loop for loop (old API, shared c for A and W).

    ``u[n1]`` is the uniform ``sample_discrete_from_log`` consumes for node n1 (:302).
    Returns (c_new, lps[N,K]) -- lps are the un-normalised conditionals at the moment of each draw.
    """
    N = A.shape[0]
    C = len(m)
    c = c.copy()
    lps = np.zeros((N, C))
    if C == 1:
        return c, lps
    for n1 in range(N):
        lp = np.zeros(C)
        lp += np.log(m)
        for cn1 in range(C):
            for n2 in range(N):
                cn2 = c[n2]
                if n2 != n1:
                    lp[cn1] += _bern_logp(p[cn1, cn2], A[n1, n2])
                    lp[cn1] += _bern_logp(p[cn2, cn1], A[n2, n1])
                    lp[cn1] += A[n1, n2] * gaussian_log_likelihood(W[n1, n2, :], mus[cn1, cn2], sigmas[cn1, cn2]).sum()
                    lp[cn1] += A[n2, n1] * gaussian_log_likelihood(W[n2, n1, :], mus[cn2, cn1], sigmas[cn2, cn1]).sum()
                else:
                    lp[cn1] += _bern_logp(p[cn1, cn1], A[n1, n1])
                    lp[cn1] += A[n1, n1] * gaussian_log_likelihood(W[n1, n1, :], mus[cn1, cn1], sigmas[cn1, cn1]).sum()
        lps[n1] = lp
        c[n1] = sample_discrete_from_log(lp, u[n1])
    return c, lps


def sbm_adjacency_resample_c_loops(A, c, m, p, special_case_self_conns, u):
    """adjacency.py:650-695, loop for loop (new API, adjacency only)."""
    N = A.shape[0]
    C = len(m)
    c = c.copy()
    lps = np.zeros((N, C))
    if C == 1:
        return c, lps
    for n1 in range(N):
        lp = np.zeros(C)
        lp += np.log(m)
        for cn1 in range(C):
            for n2 in range(N):
                cn2 = c[n2]
                if n2 == n1:
                    if special_case_self_conns:
                        continue
                    else:
                        lp[cn1] += A[n1, n1] * np.log(p[cn1, cn1]) + (1 - A[n1, n1]) * np.log(1 - p[cn1, cn1])
                else:
                    lp[cn1] += A[n1, n2] * np.log(p[cn1, cn2]) + (1 - A[n1, n2]) * np.log(1 - p[cn1, cn2])
                    lp[cn1] += A[n2, n1] * np.log(p[cn2, cn1]) + (1 - A[n2, n1]) * np.log(1 - p[cn2, cn1])
        lps[n1] = lp
        c[n1] = sample_discrete_from_log(lp, u[n1])
    return c, lps


def sbm_weights_resample_c_loops(A, W, c, m, mus, sigmas, special_case_self_conns, u):
    """weights.py:415-488 via ``_evaluate_lkhd`` (:451-470), loop for loop (new API, weights only)."""
    N = A.shape[0]
    C = len(m)
    c = c.copy()
    Abool = A.astype(bool)
    lps = np.zeros((N, C))
    if C == 1:
        return c, lps

    def _evaluate_lkhd(n1, cn1):
        chat = c.copy()
        chat[n1] = cn1
        ll = 0
        for c2 in range(C):
            out_mask = (chat == c2) & Abool[n1, :]
            if special_case_self_conns:
                out_mask[n1] = False
            if out_mask.any():
                ll += gaussian_log_likelihood(W[n1, out_mask], mus[cn1, c2], sigmas[cn1, c2]).sum()
            in_mask = (chat == c2) & Abool[:, n1]
            in_mask[n1] = False
            if in_mask.any():
                ll += gaussian_log_likelihood(W[in_mask, n1], mus[c2, cn1], sigmas[c2, cn1]).sum()
        return ll

    for n1 in range(N):
        lp = np.zeros(C)
        lp += np.log(m)
        for cn1 in range(C):
            lp[cn1] += _evaluate_lkhd(n1, cn1)
        lps[n1] = lp
        c[n1] = sample_discrete_from_log(lp, u[n1])
    return c, lps


def ais_naive_sample_f_loops(A, z, B, pi, beta, u):
    """internals/graph_models.py:261-297 applied for n = 0..N-1 in order (examples/demo_ais.py:44-46).

    The diagonal entry A[n,n] is counted in both the outgoing and incoming sums (:275-293).
    hips ``log_sum_exp_sample`` (recollection): normalise, one uniform, first index with cdf >= u.
    """
    N = A.shape[0]
    R = len(pi)
    Ab = A.astype(bool)
    z = np.array(z).astype(int)
    lps = np.zeros((N, R))
    for n in range(N):
        ln_pi_post = np.log(pi).astype(np.float64)
        zother = z.copy()
        for r in range(R):
            zother[n] = r
            o1 = Ab[n, :]
            if np.any(o1):
                ln_pi_post[r] += beta * np.sum(np.log(B[np.ix_([r], zother[o1])]))
            o2 = np.logical_not(Ab[n, :])
            if np.any(o2):
                ln_pi_post[r] += beta * np.sum(np.log(1 - B[np.ix_([r], zother[o2])]))
            i1 = Ab[:, n]
            if np.any(i1):
                ln_pi_post[r] += beta * np.sum(np.log(B[np.ix_(zother[i1], [r])]))
            i2 = np.logical_not(Ab[:, n])
            if np.any(i2):
                ln_pi_post[r] += beta * np.sum(np.log(1 - B[np.ix_(zother[i2], [r])]))
        lps[n] = ln_pi_post
        z[n] = log_sum_exp_sample(ln_pi_post, u[n])
    return z, lps


def ais_collapsed_ln_post_loops(A, z, n, pi, b0, b1):
    """internals/graph_models.py:299-335, the log posterior vector of collapsed_sample_f literally: for every candidate r
    and every (r1, r2), the edges / non-edges of A[np.ix_(zother == r1, zother == r2)]."""
    from scipy.special import betaln
    R = len(pi)
    Ab = A.astype(bool)
    zother = np.array(z).astype(int)
    ln_pi_post = np.log(pi).astype(np.float64)
    for r in range(R):
        zother[n] = r
        for r1 in range(R):
            for r2 in range(R):
                Ar1r2 = Ab[np.ix_(zother == r1, zother == r2)]
                mr1r2 = np.sum(Ar1r2)
                hat_mr1r2 = Ar1r2.size - mr1r2
                ln_pi_post[r] += betaln(mr1r2 + b1, hat_mr1r2 + b0) - betaln(b1, b0)
    return ln_pi_post


def log_sum_exp_sample(lnp, u):
    """hips.inference.log_sum_exp_sample (recollection): p = exp(lnp - logsumexp); first index with cumsum >= u."""
    lnp = np.asarray(lnp, dtype=np.float64)
    mx = lnp.max()
    p = np.exp(lnp - mx)
    p /= p.sum()
    cdf = np.cumsum(p)
    return int(min(np.searchsorted(cdf, u, side="left"), len(p) - 1))


# ----------------------------------------------------------------------------------------------
# Sequential assignment sweep (K2): vectorised count/statistics form
# ----------------------------------------------------------------------------------------------
def gaussian_coefficients(mus, sigmas):
    """Per-cell coefficients so that sum_edges log N(w; mu, Sigma) = n*c0 + c1.sum_w + sum(C2*sum_ww).

    c0 = -1/2 (B log 2pi + log|Sigma| + mu^T Sigma^-1 mu),  c1 = Sigma^-1 mu,  C2 = -1/2 Sigma^-1.
    """
    K = mus.shape[0]
    B = mus.shape[2]
    c0 = np.zeros((K, K))
    c1 = np.zeros((K, K, B))
    C2 = np.zeros((K, K, B, B))
    for a in range(K):
        for b in range(K):
            Sinv = np.linalg.inv(sigmas[a, b])
            logdet = np.linalg.slogdet(sigmas[a, b])[1]
            c1[a, b] = Sinv.dot(mus[a, b])
            c0[a, b] = -0.5 * (B * LOG2PI + logdet + mus[a, b].dot(c1[a, b]))
            C2[a, b] = -0.5 * Sinv
    return c0, c1, C2


def node_conditional(n1, A, W, c, logm, p, gc, self_mode, beta=1.0, At=None, Wt=None):
    """Un-normalised log conditional of c[n1] given everything else (vector over K).

    Count form of stochastic_block_model.py:264-299 / adjacency.py:661-691 / weights.py:451-470:
      lp[k] = log m[k] + beta * sum_k' { cnt_out[k'] log p[k,k'] + (sz[k'] - cnt_out[k']) log(1-p[k,k'])
                                        + cnt_in[k'] log p[k',k] + (sz[k'] - cnt_in[k']) log(1-p[k',k])
                                        + gauss(k,k'; out stats) + gauss(k',k; in stats) } + self term
    ``p`` None -> no adjacency term; ``gc`` None -> no weight term.
    """
    K = len(logm)
    N = A.shape[0]
    lp = np.array(logm, dtype=np.float64).copy()
    a_out = (A[n1, :] != 0)
    a_in = (A[:, n1] != 0) if At is None else (At[n1, :] != 0)
    others = np.ones(N, bool)
    others[n1] = False
    eo = a_out & others
    ei = a_in & others
    if p is not None:
        logp = np.log(p)
        log1mp = np.log1p(-p)
        sz = np.bincount(c[others], minlength=K).astype(np.float64)
        cnt_out = np.bincount(c[eo], minlength=K).astype(np.float64)
        cnt_in = np.bincount(c[ei], minlength=K).astype(np.float64)
        lp += beta * (logp.dot(cnt_out) + log1mp.dot(sz - cnt_out) + logp.T.dot(cnt_in) + log1mp.T.dot(sz - cnt_in))
    if gc is not None:
        c0, c1, C2 = gc
        B = c1.shape[2]
        n_out = np.bincount(c[eo], minlength=K).astype(np.float64)
        n_in = np.bincount(c[ei], minlength=K).astype(np.float64)
        Wo = W[n1, eo].astype(np.float64)
        Wi = (W[ei, n1] if Wt is None else Wt[n1, ei]).astype(np.float64)
        so = np.zeros((K, B)); si = np.zeros((K, B))
        soo = np.zeros((K, B, B)); sii = np.zeros((K, B, B))
        for b in range(B):
            so[:, b] = np.bincount(c[eo], weights=Wo[:, b], minlength=K)
            si[:, b] = np.bincount(c[ei], weights=Wi[:, b], minlength=K)
            for b2 in range(B):
                soo[:, b, b2] = np.bincount(c[eo], weights=Wo[:, b] * Wo[:, b2], minlength=K)
                sii[:, b, b2] = np.bincount(c[ei], weights=Wi[:, b] * Wi[:, b2], minlength=K)
        # out: cell (k, k'), in: cell (k', k)
        lp += beta * (c0.dot(n_out) + np.einsum("kjb,jb->k", c1, so) + np.einsum("kjab,jab->k", C2, soo))
        lp += beta * (c0.T.dot(n_in) + np.einsum("jkb,jb->k", c1, si) + np.einsum("jkab,jab->k", C2, sii))
    # self term
    a11 = float(A[n1, n1] != 0)
    k = np.arange(K)
    if self_mode == SELF_OLD:
        if p is not None:
            lp += beta * (a11 * np.log(p[k, k]) + (1 - a11) * np.log1p(-p[k, k]))
        if gc is not None and a11:
            w = W[n1, n1].astype(np.float64)
            lp += beta * (gc[0][k, k] + gc[1][k, k].dot(w) + np.einsum("kab,a,b->k", gc[2][k, k], w, w))
    elif self_mode == SELF_NEW:
        if p is not None:
            lp += beta * (a11 * np.log(p[k, k]) + (1 - a11) * np.log1p(-p[k, k]))
        if gc is not None and a11:
            w = W[n1, n1].astype(np.float64)
            lp += beta * (gc[0][k, k] + gc[1][k, k].dot(w) + np.einsum("kab,a,b->k", gc[2][k, k], w, w))
    elif self_mode == SELF_AIS:
        if p is not None:
            lp += 2.0 * beta * (a11 * np.log(p[k, k]) + (1 - a11) * np.log1p(-p[k, k]))
    elif self_mode != SELF_SKIP:
        raise ValueError(self_mode)
    return lp


def sbm_assign_sweep(A, W, c, m, p, mus, sigmas, u, self_mode=SELF_OLD, beta=1.0, sampler="cumsum"):
    """Vectorised sequential sweep: for n1 = 0..N-1 in order, conditional -> one injected uniform -> c[n1].

    sampler "cumsum" = pybasicbayes sample_discrete_from_log (strict >); "lse" = hips log_sum_exp_sample.
    Returns (c_new, lps[N,K]).
    """
    N = A.shape[0]
    K = len(m)
    c = np.array(c, dtype=np.int64).copy()
    lps = np.zeros((N, K))
    if K == 1:
        return c, lps
    logm = np.log(m)
    gc = gaussian_coefficients(mus, sigmas) if mus is not None else None
    At = np.ascontiguousarray(A.T)
    Wt = np.ascontiguousarray(W.transpose(1, 0, 2)) if (W is not None and gc is not None) else None
    for n1 in range(N):
        lp = node_conditional(n1, A, W, c, logm, p, gc, self_mode, beta, At, Wt)
        lps[n1] = lp
        c[n1] = sample_discrete_from_log(lp, u[n1]) if sampler == "cumsum" else log_sum_exp_sample(lp, u[n1])
    return c, lps


# ----------------------------------------------------------------------------------------------
# Log-likelihoods (K9) for the SBM classes
# ----------------------------------------------------------------------------------------------
def bernoulli_log_likelihood(A, P):
    """abstractions.py:111-115: sum A log P + (1-A) log(1 - clip(P, 1e-64, 1-1e-64))."""
    Ps = np.clip(P, 1e-64, 1 - 1e-64)
    with np.errstate(divide="ignore", invalid="ignore"):
        return np.sum(A * np.log(P) + (1 - A) * np.log(1 - Ps))


def gaussian_edge_log_likelihood_loops(A, W, Mu, Sigma):
    """abstractions.py:257-274: N^2 loop of multivariate_normal.logpdf over edges."""
    N = A.shape[0]
    ll = 0.0
    for i in range(N):
        for j in range(N):
            if A[i, j]:
                ll += gaussian_log_likelihood(W[i, j], Mu[i, j], Sigma[i, j])[0]
    return ll


def sbm_log_likelihood_from_stats(st, p, mus, sigmas, include_diag, p_self=None, self_gauss=None):
    """Bernoulli + Gaussian-edge log-likelihood of an SBM from K1 statistics.

    Equals abstractions.py:111-115 + 257-274 evaluated with P = p[c,c], Mu = mu[c,c] (old API:
    diagonal inside the cell; new API: diagonal -> p_self / _self_gaussian, weights.py:316-331).
    """
    from .niw import gaussian_loglik_from_stats
    n_edge, n_noedge, sum_w, sum_ww = cell_stats_with_diag(st, include_diag)
    K = n_edge.shape[0]
    ll_a = 0.0
    ll_w = 0.0
    if p is not None:
        ll_a += np.sum(n_edge * np.log(p) + n_noedge * np.log1p(-p))
        if not include_diag and p_self is not None:
            de = st["d_edge"].sum(); dn = st["d_cnt"].sum() - de
            ll_a += de * np.log(p_self) + dn * np.log1p(-p_self)
    if mus is not None:
        for a in range(K):
            for b in range(K):
                ll_w += gaussian_loglik_from_stats(n_edge[a, b], sum_w[a, b], sum_ww[a, b], mus[a, b], sigmas[a, b])
        if not include_diag and self_gauss is not None:
            ll_w += gaussian_loglik_from_stats(st["d_edge"].sum(), st["d_sum_w"].sum(0), st["d_sum_ww"].sum(0), *self_gauss)
    return ll_a, ll_w


# ----------------------------------------------------------------------------------------------
# Full sweeps
# ----------------------------------------------------------------------------------------------
def gaussian_sbm_gibbs_sweep(A, W, state, hypers, u, rng=np.random):
    """One ``GaussianStochasticBlockModel.resample(network)``: stochastic_block_model.py:206-230.

    Order: weight models (row-major cells) -> p -> c -> m.  ``state`` = dict(c, m, p, mus, sigmas);
    ``hypers`` = dict(pi, tau0, tau1, niw=(mu_0, sigma_0, kappa_0, nu_0)).
    Host-side parameter draws use ``rng`` (legacy np.random stream, as the reference does).
    """
    K = len(state["m"])
    st = block_pair_stats(A, W, state["c"], K)
    n_edge, n_noedge, sum_w, sum_ww = cell_stats_with_diag(st, include_diag=True)
    mus, sigmas = resample_gaussians(n_edge, sum_w, sum_ww, hypers["niw"], rng)
    p = resample_p(n_edge, n_noedge, hypers["tau1"], hypers["tau0"], rng=rng)
    c, lps = sbm_assign_sweep(A, W, state["c"], state["m"], p, mus, sigmas, u, SELF_OLD)
    m = resample_m(c, hypers["pi"], rng)
    return dict(c=c, m=m, p=p, mus=mus, sigmas=sigmas), lps
