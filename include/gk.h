/*
 * gk.h -- C ABI of libgk.so, the B200 (sm_100a) Gibbs engine behind graphistician's model classes.
 *
 * The reference (slinderman/graphistician) is pure Python and has NO FFI: its hot path is the body
 * of the model-class methods.  This header is therefore the boundary a maintainer binds with ctypes
 * (see INTEGRATION.md); each entry point names the reference method body it replaces (file:line
 * into the upstream tree).
 *
 * Conventions: extern "C"; plain pointers and sizes; every call returns 0 on success or a negative
 * GK_E* code (message via gk_last_error); one gk_ctx per host thread, not thread-safe; all device
 * work runs on the ctx stream; pointer arguments are HOST pointers unless the name ends in _dev.
 * Matrices are row-major; W is N x N x B with B innermost, as the reference stores it.
 */
#ifndef GK_H
#define GK_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct gk_ctx gk_ctx;
typedef struct gk_net gk_net;

enum { GK_OK = 0, GK_EINVAL = -1, GK_ECUDA = -2, GK_ENOMEM = -3, GK_EUNSUPPORTED = -4 };
enum { GK_F32 = 0, GK_F64 = 1 };

/* diagonal handling of the assignment sweep -- which reference loop is being replaced */
enum {
    GK_SELF_SKIP = 0, /* adjacency.py:671-674, weights.py:460-461,466-467 (special_case_self_conns)   */
    GK_SELF_ONCE = 1, /* stochastic_block_model.py:293-299; adjacency.py:675-677; weights.py:451-470  */
    GK_SELF_TWICE = 2 /* internals/graph_models.py:275-293 (AIS: diagonal in both directions)        */
};
enum {
    GK_SAMPLER_CUMSUM = 0, /* pybasicbayes sample_discrete_from_log: count(u*cum[-1] > cum)          */
    GK_SAMPLER_LSE = 1     /* hips log_sum_exp_sample: first index with normalised cdf >= u          */
};

/* ---- context ------------------------------------------------------------------------------- */
int gk_ctx_create(int device, gk_ctx **out);
void gk_ctx_destroy(gk_ctx *ctx);
const char *gk_last_error(gk_ctx *ctx);
int gk_ctx_set_stream(gk_ctx *ctx, void *cuda_stream);      /* e.g. torch's current stream handle */
int gk_ctx_synchronize(gk_ctx *ctx);
/* engine options.  GK_CTX_OPT_SBM_OVERLAP: how gk_sbm_sweep schedules its strip passes (same results in every mode):
 * 0 each block's strip pass, then its serial kernel; 1 the strip pass of node block b+1 beside the serial kernel of block b when
 * the previous sweep on the network moved <= 256 nodes; 2 the same always; 3 (default) as 1, and a sweep that moved <= 256 nodes
 * leaves the strip passes of EVERY block of the next sweep running in the background from its final labels -- they overlap the
 * caller's host work; the next sweep uses them if it is called with those labels on that network and corrects them for the
 * nodes that move meanwhile; any other use of the engine's scratch buffers or a re-upload cancels them. */
enum { GK_CTX_OPT_SBM_OVERLAP = 1 };
int gk_ctx_set_option(gk_ctx *ctx, int option, int value);
int64_t gk_ctx_launch_count(gk_ctx *ctx);                   /* kernels launched so far by this ctx */
/* per-kernel CUDA-event timing (bench.py roofline; measurement only, no counterpart in the reference); family: 0 = strip statistics kernel,
 * 1 = sequential sampler, 2 = Polya-gamma draw, 3 = eigen slab contraction, 4 = latent pairwise,
 * 5 = eigen sequential feature draw, 6 = mean-field soft strip statistics, 7 = mean-field sequential update */
int gk_ctx_kernel_timing(gk_ctx *ctx, int enable);
int gk_ctx_kernel_timing_read(gk_ctx *ctx, int family, double *ms_total, int64_t *launches, double *bytes_total);

/* ---- network data (A: N x N {0,1} uint8, W: N x N x B) --------------------------------------
 * Replaces the reference's "network" argument ((A, W) tuple, networks.py:69-71; network.A/.W,
 * stochastic_block_model.py:217-218).  A rank holds the row shard [row0, row0+nrows) of A/W and of
 * their transposes.  Device layout: see DESIGN.md "HBM layout". */
int gk_network_create(gk_ctx *ctx, int N, int B, int dev_dtype, int row0, int nrows, gk_net **out);
/* A_rows: nrows x N; W_rows: nrows x N x B (w_dtype GK_F32/GK_F64) or NULL when B == 0.
 * At_rows / Wt_rows: rows [row0,row0+nrows) of A^T / W^T, or NULL to transpose on the device
 * (allowed only when the shard is the whole matrix). */
int gk_network_upload(gk_ctx *ctx, gk_net *net, const uint8_t *A_rows, const void *W_rows, int w_dtype,
                      const uint8_t *At_rows, const void *Wt_rows);
/* diagonal entries A[n,n] (N) and W[n,n,:] (N x B) for ALL N nodes (replicated on every rank);
 * needed by the sweeps' self terms (stochastic_block_model.py:293-299). */
int gk_network_set_diag(gk_ctx *ctx, gk_net *net, const uint8_t *diagA, const double *diagW);
int gk_network_destroy(gk_ctx *ctx, gk_net *net);
size_t gk_network_device_bytes(gk_net *net);

/* ---- K1: block-pair sufficient statistics -----------------------------------------------------
 * Replaces adjacency.py:620-648 (resample_p masks), weights.py:399-413, stochastic_block_model.py:
 * 221-250, internals/weights.py:88-101 (W[A==1] gathers), adjacency.py:153-181 and weights.py:154-162
 * (K == 1).  Off-diagonal pairs only, over this rank's rows; the O(N) diagonal terms are composed by
 * the host class.  n_edge[K*K] exact; sum_w[K*K*B]; sum_ww[K*K*B*B] (symmetric, full). */
int gk_block_stats(gk_ctx *ctx, gk_net *net, const int32_t *c, int K,
                   int64_t *n_edge, double *sum_w, double *sum_ww);

/* ---- K2: sequential block-assignment sweep ----------------------------------------------------
 * Replaces stochastic_block_model.py:252-302, adjacency.py:650-695, weights.py:415-488,
 * internals/graph_models.py:261-297 (beta != 1).  Nodes are resampled in index order with exactly
 * the serial Gibbs semantics; uniforms[n] is the one uniform the reference draws for node n.
 * p (K*K) may be NULL (no adjacency term); mu (K*K*B) / sigma (K*K*B*B) may be NULL (no weight term).
 * out_lp (N*K, optional): un-normalised log conditional of each node at the moment it was drawn.
 * block_nodes: nodes per parallel statistics phase (0 = auto).  Single-rank networks only; the
 * sharded path is gk_sbm_block_partials + gk_sbm_block_sample. */
int gk_sbm_sweep(gk_ctx *ctx, gk_net *net, int32_t *c_inout, int K, const double *log_m,
                 const double *p, const double *mu, const double *sigma, int self_mode, int sampler,
                 double beta, const double *uniforms, int block_nodes, double *out_lp, int64_t *out_nchanged);

/* K1 statistics (off-diagonal pairs, exactly what gk_block_stats returns) of the labels the last unsharded gk_sbm_sweep on
 * `net` ended on.  They are folded from the per-node statistics that sweep computed anyway (sum over the nodes of each
 * destination block of their in-edge statistics, plus a correction over the rows of the nodes that moved), so the next
 * resample() -- stochastic_block_model.py:221-250 run on the labels :252-302 just produced -- needs no separate pass over A / W.
 * *valid = 0 when there is nothing to return (no sweep yet, another network or K, the network re-uploaded since, or a sweep
 * that moved more than 1024 nodes, where a fresh gk_block_stats pass is cheaper). */
int gk_sbm_sweep_stats(gk_ctx *ctx, gk_net *net, int K, int64_t *n_edge, double *sum_w, double *sum_ww, int *valid);

/* Sharded form of the same sweep: per node block, every rank computes the partial
 * per-node statistics of its own rows, the host all-reduces them, every rank samples redundantly. */
int gk_sbm_sweep_begin(gk_ctx *ctx, gk_net *net, const int32_t *c, int K, const double *log_m,
                       const double *p, const double *mu, const double *sigma, int self_mode, int sampler,
                       double beta, const double *uniforms);
/* partial statistics for nodes [n0, n0+nb): stats_dev = device buffer of gk_sbm_partial_len() doubles */
size_t gk_sbm_partial_len(gk_ctx *ctx, int nb);
int gk_sbm_block_partials(gk_ctx *ctx, gk_net *net, int n0, int nb, double *stats_dev);
int gk_sbm_block_sample(gk_ctx *ctx, gk_net *net, int n0, int nb, const double *stats_dev);
/* labels of a node block: read after the owner sampled / overwrite on the other ranks (host arrays) */
int gk_sbm_get_block(gk_ctx *ctx, int n0, int nb, int32_t *c_out);
int gk_sbm_set_block(gk_ctx *ctx, int n0, int nb, const int32_t *c_in);
/* The same exchange through DEVICE buffers, enqueued on the ctx stream without host synchronisation: the owner exports the
 * block's labels, NCCL broadcasts the buffer in-stream, the other ranks import it (labels must lie in [0, K): they come
 * from gk_sbm_block_sample on the owner).  stochastic_block_model.py:252-302 keeps c replicated on every rank this way. */
int gk_sbm_export_block_dev(gk_ctx *ctx, int n0, int nb, int32_t *labels_dev);
int gk_sbm_import_block_dev(gk_ctx *ctx, int n0, int nb, const int32_t *labels_dev);
int gk_sbm_sweep_end(gk_ctx *ctx, gk_net *net, int32_t *c_out, double *out_lp, int64_t *out_nchanged);

/* ---- eigenmodel (logistic / probit), K3-K6 ---------------------------------------------------------
 * State of one eigenmodel over one network (A only; W is ignored): S = Omega + Omega^T (row shard),
 * F (N x D), lmbda (D), mu_0.  1 <= D <= 8. */
typedef struct gk_eig gk_eig;
int gk_eig_create(gk_ctx *ctx, gk_net *net, int D, gk_eig **out);
int gk_eig_destroy(gk_ctx *ctx, gk_eig *eig);
size_t gk_eig_device_bytes(gk_eig *eig);
/* options: GK_EIG_OPT_CHOL_F64 = factorise the D x D conditionals in fp64 even with fp32 storage
 * (default with fp32 storage: fp32 factorisation; the conditionals themselves are always fp64) */
enum { GK_EIG_OPT_CHOL_F64 = 0,
       /* 1 (default): unsharded gk_eig_resample_F keeps a running contraction G for all nodes (one full pass + per-block
        * pushes); 0: re-contract a slab of S per node block (the form the row-sharded block API uses) */
       GK_EIG_OPT_FAST_SWEEP = 1,
       /* 1 (default): within the running-G sweep, the push of block b's change to all strips except the next block's
        * runs on a second stream under the serial kernel of block b+1; 0: everything on the context stream */
       GK_EIG_OPT_OVERLAP_PUSH = 2,
       /* 1 (default): with fp32 storage and fp32 factorisation the serial kernel is k_eig_seq5 (the solver warp keeps the
        * previous node's features in registers, helper warps prepare the next conditional's base row); 0: k_eig_seq4 */
       GK_EIG_OPT_SEQ5 = 3,
       /* 1 (default): with fp32 storage, D = 8 and the logistic link the once-per-sweep contraction G = S . Phi runs on the
        * tcgen05 tensor cores fed by TMA (tf32 operands, fp32 accumulation in tensor memory); 0: the FMA kernel k_eig_gfull */
       GK_EIG_OPT_TC = 4 };
/* scheduling / kernel choices of the engine, for A/B measurements and the parity tests; every setting draws from the same
 * conditionals (eigenmodel.py:871-943) -- the reference has no such switch */
int gk_eig_set_option(gk_ctx *ctx, gk_eig *eig, int option, int value);
/* upload the model attributes F (N x D), mu_0, lmbda (D)  (eigenmodel.py:41-73 state) */
int gk_eig_set_params(gk_ctx *ctx, gk_eig *eig, const double *F, double mu0, const double *lmbda);
/* K3: replaces eigenmodel.py:859-874 (ppg.pgdrawvpar over Mu.ravel()) fused with Mu of :75-83.  Draws
 * Omega[n,m] ~ PG(1, Mu[n,m]) for all N^2 entries with a counter-based generator keyed by (seed, sweep, pair)
 * and stores S = Omega + Omega^T (zero diagonal).  sum_omega (optional) = sum of ALL N^2 draws, diagonal
 * included (the precision term O.sum() of eigenmodel.py:919). */
int gk_eig_pg_draw(gk_ctx *ctx, gk_eig *eig, uint64_t seed, double *sum_omega);
/* injection path of the parity tests: S rows (nrows x N doubles) computed by the caller from a given Omega */
int gk_eig_set_S(gk_ctx *ctx, gk_eig *eig, const double *S_rows);
int gk_eig_get_S(gk_ctx *ctx, gk_eig *eig, double *S_rows);
/* K7 (probit link): replaces eigenmodel.py:268-287 (_resample_Z) with internals/utils.py:53-76 (sample_truncnorm, its
 * 1e-15 clip included) fused with Mu of :75-83.  Draws Z[n,m] for all n != m with a counter-based generator keyed by
 * (seed, sweep, n, m), stores Z + Z^T (zero diagonal) in the S buffer and switches the state to the probit link: the
 * feature sweep then runs eigenmodel.py:289-321 (S == 2, h from Z - mu_0) and `stats` feed :323-379.
 * sum_z (optional) = sum of Z[n,m] over n != m (eigenmodel.py:336-342). */
int gk_eig_tn_draw(gk_ctx *ctx, gk_eig *eig, uint64_t seed, double *sum_z);
/* injection path of the probit parity tests: rows of Z + Z^T computed by the caller (nrows x N doubles) */
int gk_eig_set_Zsym(gk_ctx *ctx, gk_eig *eig, const double *Zsym_rows);
/* the truncated-normal map by itself, uniforms supplied: out[i] = sample_truncnorm(mu[i], lb/ub from edge[i]; u[i]) */
int gk_truncnorm_draw(gk_ctx *ctx, const double *mu, const uint8_t *edge, const double *u, double *out, size_t n);
/* No counterpart in the reference.  The same map as gk_truncnorm_draw for
 * callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops): the staging scratch is a file-scope static
 * device array, so the call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine).  n is limited to
 * gk_truncnorm_static_max() = 65536 entries (1.6 MB held by every GPU that loads the library, until exit); larger n is
 * refused with GK_EINVAL and a message.  All calls on one device share the array, which keeps the previous call's
 * contents (nothing in it needs clearing: out[0, n) is overwritten before it is read): calls that overlap on the GPU race. */
int gk_truncnorm_draw_static(gk_ctx *ctx, const double *mu, const uint8_t *edge, const double *u, double *out, size_t n);
size_t gk_truncnorm_static_max(void);
/* K4+K5: replaces eigenmodel.py:876-908 (_resample_F), same serial semantics: node n is drawn from
 * N(prec_n^-1 h_n, prec_n^-1) given the NEW features of m < n and the OLD ones of m > n, with the fixed map
 * f = L^-T (L^-1 h + z), prec = L L^T, z = normals[n] (N x D standard normals supplied by the caller).
 * Outputs (all optional): F_out N x D; out_prec N x D x D and out_h N x D = the conditionals at the moment
 * each node was drawn; stats = Pq[D*D] | hS[D] | hK[D], the pair sums over n1 < n2 with the NEW F that the
 * mu_0 and lmbda conditionals need (eigenmodel.py:910-965):
 *   Pq[a,b] = sum S f1a f1b f2a f2b,  hS[a] = sum S f1a f2a,  hK[a] = sum (A12 + A21 - 1) f1a f2a. */
int gk_eig_resample_F(gk_ctx *ctx, gk_eig *eig, double sigma_F, const double *normals, double *F_out,
                      double *out_prec, double *out_h, double *stats);
/* Sharded form of the same sweep: begin; per node block of gk_eig_block_nodes() nodes
 * { block_partials on every rank -> all-reduce(acc_dev, gk_eig_acc_len doubles) -> block_sample on the rank that owns
 * the block's rows (writes the new features, nb x D doubles, to fblk_dev) -> broadcast(fblk_dev) -> set_block on the
 * other ranks }; end.  The pair statistics accumulate on the rank that draws a block: sum `stats` over the ranks. */
int gk_eig_sweep_begin(gk_ctx *ctx, gk_eig *eig, double sigma_F, const double *normals);
size_t gk_eig_acc_len(gk_eig *eig);
int gk_eig_block_nodes(gk_eig *eig);
int gk_eig_block_partials(gk_ctx *ctx, gk_eig *eig, int n0, int nb, double *acc_dev);
int gk_eig_block_sample(gk_ctx *ctx, gk_eig *eig, int n0, int nb, double *acc_dev, double *fblk_dev);
int gk_eig_set_block(gk_ctx *ctx, gk_eig *eig, int n0, int nb, const double *fblk_dev);
int gk_eig_sweep_end(gk_ctx *ctx, gk_eig *eig, double *F_out, double *stats);
/* K9: replaces eigenmodel.py:110-129 with P of :803-811 (link 0, logistic) or :227-236 (link 1, probit) */
int gk_eig_log_likelihood(gk_ctx *ctx, gk_eig *eig, int link, double *ll);

/* ---- latent-distance adjacency model, K8 -----------------------------------------------------------
 * State of one LatentDistanceAdjacencyDistribution (adjacency.py:189-207) over one network (A only):
 * L (N x dim, 1 <= dim <= 3), mu_0, mu_self, sigma.  Arithmetic in fp32 (network dtype GK_F32) with fp64
 * accumulation, or fp64 (GK_F64, parity mode). */
typedef struct gk_latent gk_latent;
int gk_latent_create(gk_ctx *ctx, gk_net *net, int dim, gk_latent **out);
int gk_latent_destroy(gk_ctx *ctx, gk_latent *lat);
/* L may be NULL (keep the device copy) */
int gk_latent_set_params(gk_ctx *ctx, gk_latent *lat, const double *L, double mu0, double mu_self, double sigma);
int gk_latent_get_params(gk_ctx *ctx, gk_latent *lat, double *L, double *mu0);
/* doubles in one accumulator of a pair pass (N*dim pair gradient sums + 4 scalars): the buffer the
 * ranks all-reduce (SUM) in the row-sharded form */
size_t gk_latent_acc_len(gk_latent *lat);
/* Replaces adjacency.py:244-275 (_hmc_log_probability) and its autograd gradient (:375-379, 398, 415) at the
 * current parameters: ll = the Bernoulli log likelihood (abstractions.py:111-115 with P of :216-219),
 * lp = ll + log priors as :266-275 write them, gradL (N x dim) = d lp / d L, gmu0, gmuself.  All outputs
 * optional.  Unsharded networks; the sharded form is eval_pass -> all-reduce(acc_dev) -> eval_finish
 * (acc_dev = device buffer of gk_latent_acc_len doubles, NULL = the state's own buffer). */
int gk_latent_lp_grad(gk_ctx *ctx, gk_latent *lat, double *ll, double *lp, double *gradL, double *gmu0, double *gmuself);
int gk_latent_eval_pass(gk_ctx *ctx, gk_latent *lat, double *acc_dev);
int gk_latent_eval_finish(gk_ctx *ctx, gk_latent *lat, const double *acc_dev, double *ll, double *lp, double *gradL,
                          double *gmu0, double *gmuself);
/* Replaces hips.inference.hmc.hmc as called by _resample_L (adjacency.py:370-385, target 0, p0 = N x dim
 * standard normals) and _resample_mu_0 (:389-404, target 1, p0 = one standard normal): leapfrog with nsteps
 * steps = nsteps + 1 pair passes; accepted iff log_u < U_curr - U_prop + K_curr - K_prop; on accept the
 * device state moves to the proposal.  log_ratio (optional) returns that right-hand side. */
int gk_latent_hmc(gk_ctx *ctx, gk_latent *lat, int target, double step, int nsteps, const double *p0, double log_u,
                  int *accepted, double *log_ratio);
/* the same trajectory in pieces, for row shards: begin; (nsteps + 1) x { pass -> all-reduce(acc_dev) -> advance }; end */
int gk_latent_hmc_begin(gk_ctx *ctx, gk_latent *lat, int target, double step, int nsteps, const double *p0);
int gk_latent_hmc_pass(gk_ctx *ctx, gk_latent *lat, double *acc_dev);
int gk_latent_hmc_advance(gk_ctx *ctx, gk_latent *lat, const double *acc_dev);
int gk_latent_hmc_end(gk_ctx *ctx, gk_latent *lat, double log_u, int *accepted, double *log_ratio);
/* Gaussian-weight sibling of the same state: LatentDistanceGaussianWeightDistribution (weights.py:613-901, B = 1),
 *   W_ij | A_ij = 1 ~ N(-||l_i - l_j||^2 + b, 1 / lam),  i != j.
 * on = 1 switches every pass of this state (lp_grad, eval_*, hmc*) to the target of _hmc_log_probability
 * (weights.py:697-731): lp = -lam/2 sum A_ij r_ij^2 - sum L^2 / (2 eta) - b^2 / 2 with r = W + d^2 - b; gk_latent_set_params'
 * mu0 is then b, sigma is eta, mu_self is unused.  Replaces the autograd gradient inside _resample_L (:765-780). */
int gk_latent_set_weight_mode(gk_ctx *ctx, gk_latent *lat, int on, double lam);
/* Statistics of the conjugate updates _resample_b (:838-860) and _resample_cov (:863-867) at the current (L, b), one
 * row pass over (A, W): out3 = { n = sum A_ij, sum A_ij r_ij, sum A_ij r_ij^2 } over i != j.  The sharded form is
 * weight_stats_pass -> all-reduce(acc_dev) -> read acc_dev[N*dim + {2, 0, 1}]. */
int gk_latent_weight_stats(gk_ctx *ctx, gk_latent *lat, double *out3);
int gk_latent_weight_stats_pass(gk_ctx *ctx, gk_latent *lat, double *acc_dev);

/* ---- K11: mean-field logistic eigenmodel ------------------------------------------------
 * Unsharded networks.  q(f_n) = N(mf_mu_F[n], mf_Sigma_F[n]) (N x D, N x D x D row-major), q(lmbda) through E_lmbda (D) and
 * E_llT (D x D), q(mu_0) = N(mf_mu_mu0, mf_sigma_mu0).
 * gk_eig_mf_omega replaces eigenmodel.py:1055-1064 (_meanfieldupdate_Omega) fused with mf_expected_mu / mf_expected_musq /
 * mf_sample_mus (:1177-1248): per directed entry n_samples activations z ~ N(E[mu], Var[mu]) from a counter-based generator
 * keyed by (seed, tile, lane), E_O = 1/2 mean tanh(z/2)/z; stores S = E_O + E_O^T (zero diagonal) in the state's S buffer.
 * sum_EO = sum of ALL N^2 entries of E_O (the precision term of :1076); sum_EO_FdotF = sum_{n != m} E_O[n,m] E_FdotF[n,m]
 * (:1079-1083).  Both optional. */
int gk_eig_mf_omega(gk_ctx *ctx, gk_eig *eig, const double *mf_mu_F, const double *mf_Sigma_F, const double *E_llT,
                    const double *E_lmbda, double mf_mu_mu0, double mf_sigma_mu0, int n_samples, uint64_t seed,
                    double *sum_EO, double *sum_EO_FdotF);
/* gk_eig_mf_update_F replaces eigenmodel.py:1090-1124 (_meanfieldupdate_F), same serial semantics: node n's q is recomputed
 * from the CURRENT q of every other node, in index order:  prec_n = I/sigma_F + E_llT o sum_m S[n,m] E[f_m f_m^T],
 * h_n = E_lmbda o sum_m (A[n,m] + A[m,n] - 1 - E_mu0 S[n,m]) lin_m,  Sigma_n <- (1-stepsize) Sigma_n + stepsize prec_n^-1,
 * mu_n <- (1-stepsize) mu_n + stepsize Sigma_n h_n.  lin = F_lin (N x D; the reference uses self.F, the Gibbs sample,
 * :1113-1114) or, when F_lin == NULL, mf_mu_F itself as it is updated.  out_prec (N x D x D) / out_h (N x D) optional:
 * the conditionals at the moment each node was updated.  S comes from gk_eig_mf_omega or gk_eig_set_S. */
int gk_eig_mf_update_F(gk_ctx *ctx, gk_eig *eig, double *mf_mu_F, double *mf_Sigma_F, const double *F_lin, const double *E_llT,
                       const double *E_lmbda, double E_mu0, double sigma_F, double stepsize, double *out_prec, double *out_h);
/* pair sums over n1 < n2 of eigenmodel.py:1126-1157 (_meanfieldupdate_lmbda): stats = Pq[D*D] | hS[D] | hK[D] with
 * Pq[a,b] = sum S E[f1 f1^T][a,b] E[f2 f2^T][a,b],  hS[a] = sum S mu1a mu2a,  hK[a] = sum (A12 + A21 - 1) mu1a mu2a. */
int gk_eig_mf_pair_stats(gk_ctx *ctx, gk_eig *eig, const double *mf_mu_F, const double *mf_Sigma_F, double *stats);

/* Mean-field PROBIT eigenmodel (eigenmodel.py:382-770, _MeanFieldProbitEigenModel).  gk_eig_mf_probit_Z replaces
 * _meanfieldupdate_Z (:570-585: mf_mu_Z = E[mu_0] + E[f_n]^T diag(E[lmbda]) E[f_m], diagonal = E[mu_0]) fused with mf_expected_Z
 * (:519-534) and internals/utils.py:78-101 (expected_truncnorm) for a fixed network (mf_A = A): E[Z] of every directed entry is
 * computed per tile, E[Z] + E[Z]^T (zero diagonal) is stored in the state's S buffer and the state switches to the probit link, so
 * that gk_eig_mf_update_F then runs :610-641 (S == 2, h from E[Z] - E[mu_0]) and gk_eig_mf_pair_stats feeds :643-674.
 * exact_normalizer = 0 keeps the reference's normaliser (utils.py:93 takes normal_cdf of the RAW bounds, i.e. 1/2 whatever mu:
 * E = mu +- 2 phi(mu)); 1 = the truncated-normal mean mu +- phi(mu)/Phi(+-mu).  sum_EZ (optional) = sum of E[Z] over n != m (:598-604). */
int gk_eig_mf_probit_Z(gk_ctx *ctx, gk_eig *eig, const double *mf_mu_F, const double *E_lmbda, double E_mu0,
                       int exact_normalizer, double *sum_EZ);

/* ---- K10: mean-field SBM ---------------------------------------------------------
 * Fixed networks only (E_A = A, E_W = W, E_WWT = W W^T: the reference's FixedGaussianNetwork); unsharded.
 * gk_mf_sbm_soft_stats replaces the C x C loops of stochastic_block_model.py:366-389 (mf_update_p) and :340-344 with
 * internals/weights.py:158-171 (weighted NIW statistics): n_soft[c1,c2] = sum_{i != j} mf_m[i,c1] mf_m[j,c2] A[i,j],
 * sum_w / sum_ww the same sums of A*W_b / A*W_b*W_b'.  Off-diagonal pairs only; the O(N) diagonal terms are composed
 * by the host class.  mf_m: N x K row-major. */
int gk_mf_sbm_soft_stats(gk_ctx *ctx, gk_net *net, const double *mf_m, int K, double *n_soft, double *sum_w,
                         double *sum_ww);
/* gk_mf_sbm_update_c replaces stochastic_block_model.py:392-445 (mf_update_c), same serial semantics: node n1's
 * class probabilities are recomputed from the CURRENT mf_m of all other nodes, in index order.  Per-partner terms of
 * lp[k] for a partner of class k' (tables K x K [x S], candidate class k first, S = 1 + B + B(B+1)/2 packed
 * statistics (a, a*w_b, a*w_b*w_b' for b <= b')):
 *   forward  n1 -> n2 (n2 == n1 included):  base_out[k,k'] + sum_s x_s(n1,n2) coef_out[k,k',s]
 *   reverse  n2 -> n1 (n2 != n1):           base_in [k,k'] + sum_s x_s(n2,n1) coef_in [k,k',s]
 * mf_m[n1,:] <- (1-stepsize) mf_m[n1,:] + stepsize softmax(lp).  out_lp (N x K, optional) = lp of each node at the
 * moment it was updated.  block_nodes: nodes per parallel statistics phase (0 = auto, <= 256). */
int gk_mf_sbm_update_c(gk_ctx *ctx, gk_net *net, double *mf_m_inout, int K, const double *E_log_m,
                       const double *base_out, const double *base_in, const double *coef_out, const double *coef_in,
                       double stepsize, int block_nodes, double *out_lp);

/* ---- AIS over a bit-packed adjacency matrix, batched over particles -------------------
 * Replaces the body of the driver loop of examples/demo_ais.py:10-62 for P co-resident particles.  A and A^T are held
 * bit-packed (N^2/8 bytes each); each particle keeps its labels as R class bit-masks, so the class counts of node n are
 * popc(row_n & mask) and one streamed row of A serves every particle.  R <= 8 (template instantiations). */
typedef struct gk_ais gk_ais;
enum { GK_AIS_ER = 0,   /* internals/graph_models.py:105-182 (ErdosRenyiNetwork)     */
       GK_AIS_SBM = 1   /* internals/graph_models.py:185-371 (StochasticBlockModel)  */ };
int gk_ais_create(gk_ctx *ctx, int N, int P, int R, gk_ais **out);
int gk_ais_destroy(gk_ctx *ctx, gk_ais *ais);
size_t gk_ais_device_bytes(gk_ais *ais);
/* GK_AIS_OPT_INCREMENTAL: keep the class counts of every node on the device and move them along with the labels (no counting
 * pass in the sweep; same results) -- 0 never, 1 (default) after a sweep in which no particle moved more than 256 nodes,
 * 2 always.  Costs P x N x 2 (R - 1) x 4 bytes. */
enum { GK_AIS_OPT_INCREMENTAL = 1 };
int gk_ais_set_option(gk_ctx *ctx, gk_ais *ais, int option, int value);
/* rows [row0, row0 + nrows) of the {0,1} uint8 matrix A (nrows x N, row0 a multiple of 32; the `A` of
 * estimate_marginal_likelihood, demo_ais.py:10) -> the bit-packed rows of A and bit columns of A^T.  src_on_device != 0:
 * A_rows is a device pointer. */
int gk_ais_upload_rows(gk_ctx *ctx, gk_ais *ais, const uint8_t *A_rows, int row0, int nrows, int src_on_device);
/* the same from the uint8 copy an unsharded gk_net already holds on the device */
int gk_ais_upload_network(gk_ctx *ctx, gk_ais *ais, gk_net *net);
/* sum(A) over all N^2 entries: nnz_A of ErdosRenyiNetwork.sample_theta (internals/graph_models.py:178) */
int gk_ais_nnz(gk_ctx *ctx, gk_ais *ais, int64_t *nnz);
/* labels z (P x N, or N when broadcast != 0: `f, theta = copy.deepcopy((f0, theta0))`, demo_ais.py:22); builds the class
 * masks, class sizes and the R x R edge-count table of every particle */
int gk_ais_set_labels(gk_ctx *ctx, gk_ais *ais, const int32_t *z, int broadcast);
int gk_ais_get_labels(gk_ctx *ctx, gk_ais *ais, int32_t *z);
/* n_edge[P x R x R] = sum(A[np.ix_(z == r1, z == r2)]) (diagonal included, internals/graph_models.py:364-367 and the
 * N^2 loop of log_lkhd :42-48), sizes[P x R] = sum(z == r) (:347-349).  recompute = 0: the tables the sweeps keep
 * current (exact integer updates per move); 1: a fresh pass over A. */
int gk_ais_counts(gk_ctx *ctx, gk_ais *ais, int recompute, int64_t *n_edge, int64_t *sizes);
/* `for n in arange(N): f[n] = model.sample_f(theta, (n, A, f), beta)` (demo_ais.py:44-46 -> naive_sample_f,
 * internals/graph_models.py:261-297) for every particle: exact serial semantics per particle, particles in parallel.
 * log_pi[P x R], B[P x R x R] (theta of each particle), beta[P]; uniforms[P x N] = the uniform log_sum_exp_sample draws
 * for node n, or NULL: a counter-based stream keyed by (seed, sweep number, particle, node).  out_lp (P x N x R,
 * optional): ln_pi_post of each node at the moment it was drawn.  out_nchanged (P, optional). */
int gk_ais_sweep(gk_ctx *ctx, gk_ais *ais, const double *log_pi, const double *B, const double *beta,
                 const double *uniforms, uint64_t seed, double *out_lp, int64_t *out_nchanged);
/* The whole of estimate_marginal_likelihood (demo_ais.py:10-62) for the P particles of this state, all starting at f0
 * (N labels; unused for GK_AIS_ER): n_temps temperatures linspace(0, 1, n_temps), steps_per_temp x (sample_theta, feature
 * sweep) per temperature, log_weights[p] = sum_b (beta_b - beta_{b-1}) log_lkhd.  Beta(b1 + beta edges, b0 + beta non-edges)
 * (ER: x = (b1, b0)), Dirichlet(alpha0[R] + sizes); the O(R^2) conjugate draws run on the host from a xoshiro256** stream
 * seeded by `seed` (the Python shell draws the seed from np.random); the caller forms log Z = -log(P) + logsumexp. */
int gk_ais_run(gk_ctx *ctx, gk_ais *ais, int model, double b0, double b1, const double *alpha0, int n_temps,
               int steps_per_temp, const int32_t *f0, uint64_t seed, double *log_weights);

/* ---- the engine's own communicator (no counterpart in the reference, which is single-process) ---------
 * One process per GPU.  Rank 0 asks for an id (128 bytes, NCCL's ncclUniqueId), the host passes it to the other ranks by
 * any channel it has (torch.distributed, MPI, a file), every rank calls gk_comm_init.  From then on the calls that take a
 * ROW-SHARDED network -- gk_sbm_sweep, gk_eig_resample_F, gk_latent_hmc -- run their collectives themselves, in-stream
 * between their kernels: per node block the partial statistics are all-reduced, the owner of the block's rows draws it and
 * its labels / features are broadcast (rows of rank r: whole blocks of 128 rows, spread evenly: [128*(nblk*r/world),
 * 128*(nblk*(r+1)/world)) with nblk = ceil(N/128)).  libnccl.so.2 is opened at run time (dlopen). */
int gk_comm_unique_id(gk_ctx *ctx, uint8_t *id128);
int gk_comm_init(gk_ctx *ctx, const uint8_t *id128, int rank, int world);
int gk_comm_destroy(gk_ctx *ctx);
int gk_comm_rank(gk_ctx *ctx);
int gk_comm_world(gk_ctx *ctx);
/* in-stream SUM over the ranks of n doubles at a device address (for the host-driven block forms) */
int gk_comm_allreduce(gk_ctx *ctx, double *buf_dev, size_t n);

#ifdef __cplusplus
}
#endif
#endif /* GK_H */
