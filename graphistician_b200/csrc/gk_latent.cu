// gk_latent.cu -- host side of the latent-distance path (K8): state, HMC trajectory pieces, C-ABI entry points.
//
// hips.inference.hmc.hmc (called at adjacency.py:381-385, 402-404) is a leapfrog trajectory of n steps that needs
// n + 1 gradient evaluations and 2 value evaluations; here every evaluation is ONE pair pass (k_latent_pair), the
// first and the last one also producing the value.  The O(N dim) kick / drift updates and the accept test run in
// single-CTA kernels so a whole trajectory is enqueued without a host round trip; in the row-sharded form the
// caller all-reduces the accumulator between gk_latent_hmc_pass and gk_latent_hmc_advance.
#include <math.h>
#include <algorithm>
#include <vector>
#include "gk_latent_kernels.cuh"

enum { SC_MU0 = 0, SC_MU0_Q, SC_P_MU0, SC_LP_CURR, SC_LP_PROP, SC_K_CURR, SC_K_PROP, SC_ACCEPT, SC_RATIO, SC_LL, SC_GMU0,
       SC_GMUSELF, SC_LP, SC_LEN = 16 };

struct gk_latent {
    gk_net *net = nullptr;
    int N = 0, dim = 0; bool f64 = false;
    double *L = nullptr, *q = nullptr, *p = nullptr;   // N x dim each: current state, proposal, momentum
    void *LT = nullptr, *qT = nullptr;                 // compute-type shadows of L and q
    double *sc = nullptr, *acc = nullptr, *grad = nullptr;
    double sigma = 1.0, mu_self = 0.0;
    bool wmode = false; double lam = 1.0;              // Gaussian-weight target (weights.py:613-901): mu_0 plays b, sigma plays eta
    double scale() const { return (f64 || wmode) ? 1.0 : GK_LAT_SCALE_F32; }   // pre-scaling of the compute-type shadows
    long long nd = 0;                                  // number of self edges (sum of diag A)
    // open trajectory
    int target = -1, nsteps = 0, stage = 0; double step = 0.0;
};

static inline size_t lat_acc_len(const gk_latent *l) { return (size_t)l->N * l->dim + GK_LAT_ACC_EXTRA; }

// compute-type shadow of a coordinate array, pre-scaled for the pair kernel (gk_lat_math<T>::scale())
template <typename T>
__global__ void k_lat_shadow(const double *__restrict__ src, T *__restrict__ dst, size_t n, double scale) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = (T)(src[i] * scale);
}

__device__ __forceinline__ double lat_block_sum(double v, double *red) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
    return t;
}
__device__ __forceinline__ double lat_softplus(double x) { return fmax(x, 0.0) + log1p(exp(-fabs(x))); }
__device__ __forceinline__ double lat_sigmoid(double x) { return 1.0 / (1.0 + exp(-x)); }

struct gk_lat_fin {
    const double *acc; double *sc; int N, dim; double sigma, mu_self, nd; int wmode; double lam;
};
// value of adjacency.py:244-275 from the pair sum: diagonal (eta_ii = mu_0 + mu_self) and priors in closed form
__device__ __forceinline__ double lat_value(const gk_lat_fin &f, double mu0, double sumsq, double *ll_out) {
    if (f.wmode) {                                             // weights.py:697-731: Gaussian residuals, priors on L and b
        const double ll = -0.5 * f.lam * f.acc[(size_t)f.N * f.dim + GK_LAT_ACC_SV];
        if (ll_out) *ll_out = ll;
        return ll - 0.5 * sumsq / f.sigma - 0.5 * mu0 * mu0;
    }
    const double ed = mu0 + f.mu_self;
    const double ll = f.acc[(size_t)f.N * f.dim + GK_LAT_ACC_SV] + f.nd * ed - (double)f.N * lat_softplus(ed);
    if (ll_out) *ll_out = ll;
    return ll - 0.5 * sumsq / f.sigma - 0.5 * mu0 * mu0 - 0.5 * f.mu_self * f.mu_self;
}
__device__ __forceinline__ double lat_gmu0(const gk_lat_fin &f, double mu0) {
    if (f.wmode) return f.lam * f.acc[(size_t)f.N * f.dim + GK_LAT_ACC_SW] - mu0;      // d lp / d b = Lambda sum A r - b
    return f.acc[(size_t)f.N * f.dim + GK_LAT_ACC_SW] + f.nd - (double)f.N * lat_sigmoid(mu0 + f.mu_self) - mu0;
}

// one-shot evaluation at the current parameters: gradients and values to sc / grad
__global__ void __launch_bounds__(1024) k_lat_eval_finish(gk_lat_fin f, const double *__restrict__ L, double *__restrict__ grad) {
    __shared__ double red[32];
    const size_t n = (size_t)f.N * f.dim;
    double ss = 0.0;
    for (size_t i = threadIdx.x; i < n; i += blockDim.x) {
        const double l = L[i];
        ss += l * l;
        grad[i] = -2.0 * f.acc[i] - l / f.sigma;
    }
    ss = lat_block_sum(ss, red);
    if (threadIdx.x == 0) {
        const double mu0 = f.sc[SC_MU0];
        double ll;
        f.sc[SC_LP] = lat_value(f, mu0, ss, &ll);
        f.sc[SC_LL] = ll;
        f.sc[SC_GMU0] = lat_gmu0(f, mu0);
        f.sc[SC_GMUSELF] = f.wmode ? 0.0 : f.nd - (double)f.N * lat_sigmoid(mu0 + f.mu_self) - f.mu_self;
    }
}

// leapfrog piece for the L trajectory.  kind 0: first (record U_curr, K_curr; half kick; drift), 1: middle (kick; drift),
// 2: last (half kick; record U_prop, K_prop).  One-step trajectories (nsteps == 1) use kinds 0 then 2.
template <typename T>
__global__ void __launch_bounds__(1024) k_lat_advance_L(gk_lat_fin f, double *__restrict__ q, double *__restrict__ p,
                                                        T *__restrict__ qT, double eps, int kind, double scale) {
    __shared__ double red[32];
    const size_t n = (size_t)f.N * f.dim;
    const double kick = kind == 1 ? eps : 0.5 * eps;
    double ssq = 0.0, ssp = 0.0;
    for (size_t i = threadIdx.x; i < n; i += blockDim.x) {
        double qi = q[i], pi = p[i];
        if (kind == 0) ssp += pi * pi;
        if (kind != 1) ssq += qi * qi;
        const double dlp = -2.0 * f.acc[i] - qi / f.sigma;      // d log p / d q ; U = -log p
        pi += kick * dlp;
        if (kind == 2) ssp += pi * pi;
        if (kind != 2) { qi += eps * pi; q[i] = qi; qT[i] = (T)(qi * scale); }
        p[i] = pi;
    }
    if (kind == 1) return;
    ssq = lat_block_sum(ssq, red);
    ssp = lat_block_sum(ssp, red);
    if (threadIdx.x == 0) {
        const double v = lat_value(f, f.sc[SC_MU0], ssq, nullptr);
        if (kind == 0) { f.sc[SC_LP_CURR] = v; f.sc[SC_K_CURR] = 0.5 * ssp; }
        else { f.sc[SC_LP_PROP] = v; f.sc[SC_K_PROP] = 0.5 * ssp; }
    }
}

// the same for the scalar mu_0 trajectory (the L prior term is the same constant at both ends and is left out)
__global__ void k_lat_advance_mu0(gk_lat_fin f, double eps, int kind) {
    if (threadIdx.x || blockIdx.x) return;
    double q = f.sc[SC_MU0_Q], p = f.sc[SC_P_MU0];
    const double kick = kind == 1 ? eps : 0.5 * eps;
    if (kind == 0) { f.sc[SC_LP_CURR] = lat_value(f, q, 0.0, nullptr); f.sc[SC_K_CURR] = 0.5 * p * p; }
    p += kick * lat_gmu0(f, q);
    if (kind == 2) { f.sc[SC_LP_PROP] = lat_value(f, q, 0.0, nullptr); f.sc[SC_K_PROP] = 0.5 * p * p; }
    else { q += eps * p; f.sc[SC_MU0_Q] = q; }
    f.sc[SC_P_MU0] = p;
}

// accept test of hips hmc: log u < U_curr - U_prop + K_curr - K_prop  (U = -lp); on accept the proposal becomes the state
template <typename T>
__global__ void __launch_bounds__(1024) k_lat_end(double *sc, double log_u, int target, size_t n, const double *__restrict__ q,
                                                  const T *__restrict__ qT, double *__restrict__ L, T *__restrict__ LT) {
    const double ratio = sc[SC_LP_PROP] - sc[SC_LP_CURR] + sc[SC_K_CURR] - sc[SC_K_PROP];
    const bool accept = log_u < ratio;                         // NaN (diverged trajectory) rejects
    if (accept && target == 0)
        for (size_t i = threadIdx.x; i < n; i += blockDim.x) { L[i] = q[i]; LT[i] = qT[i]; }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (accept && target == 1) sc[SC_MU0] = sc[SC_MU0_Q];
        sc[SC_ACCEPT] = accept ? 1.0 : 0.0; sc[SC_RATIO] = ratio;
    }
}

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
extern "C" int gk_latent_create(gk_ctx *ctx, gk_net *net, int dim, gk_latent **out) {
    if (!ctx || !net || !out) return GK_EINVAL;
    *out = nullptr;
    if (dim < 1 || dim > GK_LAT_MAXDIM) return gk_fail(ctx, GK_EUNSUPPORTED, "dim=%d: this build supports 1 <= dim <= %d", dim, GK_LAT_MAXDIM);
    if (net->row0 % GK_LAT_TILE || (net->row0 + net->nrows != net->N && net->nrows % GK_LAT_TILE))
        return gk_fail(ctx, GK_EINVAL, "row shards must start and end on multiples of %d (or at N)", GK_LAT_TILE);
    if (!net->have_diag) return gk_fail(ctx, GK_EINVAL, "gk_network_set_diag must be called first (self-edge count)");
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = gk_net_ensure_asym(ctx, net);
    if (rc) return rc;
    gk_latent *l = new gk_latent();
    l->net = net; l->N = net->N; l->dim = dim; l->f64 = net->dtype == GK_F64;
    const size_t n = (size_t)net->N * dim, ts = l->f64 ? 8 : 4;
    const size_t npad = (size_t)((net->N + GK_LAT_TILE - 1) / GK_LAT_TILE) * GK_LAT_TILE * dim;   // shadows are zero padded to whole tiles
    cudaError_t e = cudaMalloc(&l->L, n * 8);
    if (e == cudaSuccess) e = cudaMalloc(&l->q, n * 8);
    if (e == cudaSuccess) e = cudaMalloc(&l->p, n * 8);
    if (e == cudaSuccess) e = cudaMalloc(&l->grad, n * 8);
    if (e == cudaSuccess) e = cudaMalloc(&l->LT, npad * ts);
    if (e == cudaSuccess) e = cudaMalloc(&l->qT, npad * ts);
    if (e == cudaSuccess) e = cudaMalloc(&l->sc, SC_LEN * 8);
    if (e == cudaSuccess) e = cudaMalloc(&l->acc, lat_acc_len(l) * 8);
    if (e != cudaSuccess) {
        cudaFree(l->L); cudaFree(l->q); cudaFree(l->p); cudaFree(l->grad); cudaFree(l->LT); cudaFree(l->qT); cudaFree(l->sc); cudaFree(l->acc);
        delete l;
        return gk_fail(ctx, GK_ENOMEM, "latent state alloc: %s", cudaGetErrorString(e));
    }
    GK_CUDA(ctx, cudaMemsetAsync(l->L, 0, n * 8, ctx->stream));
    GK_CUDA(ctx, cudaMemsetAsync(l->LT, 0, npad * ts, ctx->stream));
    GK_CUDA(ctx, cudaMemsetAsync(l->qT, 0, npad * ts, ctx->stream));
    GK_CUDA(ctx, cudaMemsetAsync(l->sc, 0, SC_LEN * 8, ctx->stream));
    std::vector<uint8_t> d(net->N);
    GK_CUDA(ctx, cudaMemcpyAsync(d.data(), net->diagA, net->N, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < net->N; ++i) l->nd += d[i] != 0;
    *out = l;
    return GK_OK;
}

extern "C" int gk_latent_destroy(gk_ctx *ctx, gk_latent *l) {
    if (!l) return GK_OK;
    if (ctx) { cudaSetDevice(ctx->device); cudaStreamSynchronize(ctx->stream); }
    cudaFree(l->L); cudaFree(l->q); cudaFree(l->p); cudaFree(l->grad); cudaFree(l->LT); cudaFree(l->qT); cudaFree(l->sc); cudaFree(l->acc);
    delete l;
    return GK_OK;
}

extern "C" size_t gk_latent_acc_len(gk_latent *l) { return l ? lat_acc_len(l) : 0; }

static int lat_shadow(gk_ctx *ctx, gk_latent *l, const double *src, void *dst) {
    const size_t n = (size_t)l->N * l->dim;
    const int grid = (int)std::min<size_t>((n + 255) / 256, (size_t)ctx->num_sms * 4);
    if (l->f64) k_lat_shadow<double><<<grid, 256, 0, ctx->stream>>>(src, (double *)dst, n, 1.0);
    else k_lat_shadow<float><<<grid, 256, 0, ctx->stream>>>(src, (float *)dst, n, l->scale());
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

extern "C" int gk_latent_set_params(gk_ctx *ctx, gk_latent *l, const double *L, double mu0, double mu_self, double sigma) {
    if (!ctx || !l) return GK_EINVAL;
    if (!(sigma > 0)) return gk_fail(ctx, GK_EINVAL, "sigma must be positive");
    if (l->target >= 0) return gk_fail(ctx, GK_EINVAL, "a trajectory is open");
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    l->sigma = sigma; l->mu_self = mu_self;
    GK_CUDA(ctx, cudaMemcpyAsync(l->sc + SC_MU0, &mu0, 8, cudaMemcpyHostToDevice, ctx->stream));
    if (L) {
        GK_CUDA(ctx, cudaMemcpyAsync(l->L, L, (size_t)l->N * l->dim * 8, cudaMemcpyHostToDevice, ctx->stream));
        int rc = lat_shadow(ctx, l, l->L, l->LT);
        if (rc) return rc;
    }
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GK_OK;
}

extern "C" int gk_latent_get_params(gk_ctx *ctx, gk_latent *l, double *L, double *mu0) {
    if (!ctx || !l) return GK_EINVAL;
    if (L) GK_CUDA(ctx, cudaMemcpyAsync(L, l->L, (size_t)l->N * l->dim * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (mu0) GK_CUDA(ctx, cudaMemcpyAsync(mu0, l->sc + SC_MU0, 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GK_OK;
}

template <typename T, int DIM>
static void lat_launch(gk_ctx *ctx, const gk_lat_pass_params &P, bool value, bool grad, int grid) {
    if (value && grad) k_latent_pair<T, DIM, true, true><<<grid, GK_LAT_THREADS, 0, ctx->stream>>>(P);
    else if (value) k_latent_pair<T, DIM, true, false><<<grid, GK_LAT_THREADS, 0, ctx->stream>>>(P);
    else if (grad) k_latent_pair<T, DIM, false, true><<<grid, GK_LAT_THREADS, 0, ctx->stream>>>(P);
    else k_latent_pair<T, DIM, false, false><<<grid, GK_LAT_THREADS, 0, ctx->stream>>>(P);
}

template <typename T, int DIM>
static void latw_launch(gk_ctx *ctx, const gk_latw_params &P, bool grad, int grid) {
    if (grad) k_latw_rows<T, DIM, true><<<grid, GK_LATW_THREADS, 0, ctx->stream>>>(P);
    else k_latw_rows<T, DIM, false><<<grid, GK_LATW_THREADS, 0, ctx->stream>>>(P);
}

// Gaussian-weight target: one row pass over this rank's rows of (A, W) and (A^T, W^T) at (Lq, *b) into acc_dev
static int latw_pass(gk_ctx *ctx, gk_latent *l, const void *Lq, const double *b_dev, bool grad, double *acc_dev) {
    gk_net *net = l->net;
    gk_latw_params P;
    P.A = net->A; P.At = net->At; P.W = net->W; P.Wt = net->Wt; P.ld = net->ld; P.row0 = net->row0; P.nrows = net->nrows; P.N = net->N;
    P.Lq = Lq; P.b = b_dev; P.lam = l->lam; P.acc = acc_dev;
    const int groups = (net->nrows + GK_LATW_ROWS - 1) / GK_LATW_ROWS;
    const int grid = std::max(1, std::min(groups, 8 * ctx->num_sms));
    gk_time_begin(ctx, 4, (double)net->nrows * net->N * (grad ? 2.0 : 1.0) * (1.0 + (l->f64 ? 8.0 : 4.0)));
    if (l->f64) {
        if (l->dim == 1) latw_launch<double, 1>(ctx, P, grad, grid);
        else if (l->dim == 2) latw_launch<double, 2>(ctx, P, grad, grid);
        else latw_launch<double, 3>(ctx, P, grad, grid);
    } else {
        if (l->dim == 1) latw_launch<float, 1>(ctx, P, grad, grid);
        else if (l->dim == 2) latw_launch<float, 2>(ctx, P, grad, grid);
        else latw_launch<float, 3>(ctx, P, grad, grid);
    }
    gk_time_end(ctx, 4);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

// one pair pass over this rank's rows at (Lq, *mu0) into acc_dev (zeroed here)
static int lat_pass(gk_ctx *ctx, gk_latent *l, const void *Lq, const double *mu0_dev, bool value, bool grad, double *acc_dev) {
    gk_net *net = l->net;
    GK_CUDA(ctx, cudaMemsetAsync(acc_dev, 0, lat_acc_len(l) * 8, ctx->stream));
    if (l->wmode) return latw_pass(ctx, l, Lq, mu0_dev, grad, acc_dev);
    gk_lat_pass_params P;
    P.Asym = net->Asym; P.ld = net->ld; P.row0 = net->row0; P.nrows = net->nrows; P.N = net->N;
    P.Lq = Lq; P.mu0 = mu0_dev; P.acc = acc_dev; P.magic = 0x4B000000u;
    P.NT = (net->N + GK_LAT_TILE - 1) / GK_LAT_TILE;
    P.nkc = (P.NT / 2 + 1 + GK_LAT_KCH - 1) / GK_LAT_KCH;
    P.nunits = ((net->nrows + GK_LAT_TILE - 1) / GK_LAT_TILE) * P.nkc;
    const int grid = std::max(1, std::min(P.nunits, 2 * ctx->num_sms));
    gk_time_begin(ctx, 4, (double)net->nrows * net->N);
    if (l->f64) {
        if (l->dim == 1) lat_launch<double, 1>(ctx, P, value, grad, grid);
        else if (l->dim == 2) lat_launch<double, 2>(ctx, P, value, grad, grid);
        else lat_launch<double, 3>(ctx, P, value, grad, grid);
    } else {
        if (l->dim == 1) lat_launch<float, 1>(ctx, P, value, grad, grid);
        else if (l->dim == 2) lat_launch<float, 2>(ctx, P, value, grad, grid);
        else lat_launch<float, 3>(ctx, P, value, grad, grid);
    }
    gk_time_end(ctx, 4);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

static gk_lat_fin lat_fin(gk_latent *l, const double *acc) {
    gk_lat_fin f; f.acc = acc; f.sc = l->sc; f.N = l->N; f.dim = l->dim; f.sigma = l->sigma; f.mu_self = l->mu_self; f.nd = (double)l->nd;
    f.wmode = l->wmode; f.lam = l->lam;
    return f;
}

extern "C" int gk_latent_eval_pass(gk_ctx *ctx, gk_latent *l, double *acc_dev) {
    if (!ctx || !l) return GK_EINVAL;
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    return lat_pass(ctx, l, l->LT, l->sc + SC_MU0, true, true, acc_dev ? acc_dev : l->acc);
}

extern "C" int gk_latent_eval_finish(gk_ctx *ctx, gk_latent *l, const double *acc_dev, double *ll, double *lp, double *gradL,
                                     double *gmu0, double *gmuself) {
    if (!ctx || !l) return GK_EINVAL;
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    k_lat_eval_finish<<<1, 1024, 0, ctx->stream>>>(lat_fin(l, acc_dev ? acc_dev : l->acc), l->L, l->grad);
    GK_CHECK_LAUNCH(ctx);
    double sc[SC_LEN];
    GK_CUDA(ctx, cudaMemcpyAsync(sc, l->sc, sizeof sc, cudaMemcpyDeviceToHost, ctx->stream));
    if (gradL) GK_CUDA(ctx, cudaMemcpyAsync(gradL, l->grad, (size_t)l->N * l->dim * 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ll) *ll = sc[SC_LL];
    if (lp) *lp = sc[SC_LP];
    if (gmu0) *gmu0 = sc[SC_GMU0];
    if (gmuself) *gmuself = sc[SC_GMUSELF];
    return GK_OK;
}

extern "C" int gk_latent_lp_grad(gk_ctx *ctx, gk_latent *l, double *ll, double *lp, double *gradL, double *gmu0, double *gmuself) {
    if (!ctx || !l) return GK_EINVAL;
    if (l->net->row0 != 0 || l->net->nrows != l->N) return gk_fail(ctx, GK_EINVAL, "gk_latent_lp_grad needs an unsharded network; use eval_pass / all-reduce / eval_finish");
    int rc = gk_latent_eval_pass(ctx, l, nullptr);
    if (rc) return rc;
    return gk_latent_eval_finish(ctx, l, nullptr, ll, lp, gradL, gmu0, gmuself);
}

// ---- Gaussian-weight target (weights.py:613-901) -------------------------------------------------------
extern "C" int gk_latent_set_weight_mode(gk_ctx *ctx, gk_latent *l, int on, double lam) {
    if (!ctx || !l) return GK_EINVAL;
    if (l->target >= 0) return gk_fail(ctx, GK_EINVAL, "a trajectory is open");
    if (on && l->net->B != 1) return gk_fail(ctx, GK_EUNSUPPORTED, "the distance model of the weights is one-dimensional (B = 1, weights.py:705); network has B=%d", l->net->B);
    if (on && !(lam > 0)) return gk_fail(ctx, GK_EINVAL, "the precision of the weights must be positive");
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    const bool rescale = (on != 0) != l->wmode;
    l->wmode = on != 0; if (on) l->lam = lam;
    if (rescale) { int rc = lat_shadow(ctx, l, l->L, l->LT); if (rc) return rc; }
    return GK_OK;
}

extern "C" int gk_latent_weight_stats_pass(gk_ctx *ctx, gk_latent *l, double *acc_dev) {
    if (!ctx || !l || !l->wmode) return GK_EINVAL;
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    return lat_pass(ctx, l, l->LT, l->sc + SC_MU0, true, false, acc_dev ? acc_dev : l->acc);
}

extern "C" int gk_latent_weight_stats(gk_ctx *ctx, gk_latent *l, double *out3) {
    if (!ctx || !l || !out3) return GK_EINVAL;
    if (!l->wmode) return gk_fail(ctx, GK_EINVAL, "gk_latent_set_weight_mode first");
    if (l->net->row0 != 0 || l->net->nrows != l->N) return gk_fail(ctx, GK_EINVAL, "gk_latent_weight_stats needs an unsharded network; use gk_latent_weight_stats_pass / all-reduce");
    int rc = gk_latent_weight_stats_pass(ctx, l, nullptr);
    if (rc) return rc;
    double ex[GK_LAT_ACC_EXTRA];
    GK_CUDA(ctx, cudaMemcpyAsync(ex, l->acc + (size_t)l->N * l->dim, sizeof ex, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    out3[0] = ex[GK_LAT_ACC_N]; out3[1] = ex[GK_LAT_ACC_SW]; out3[2] = ex[GK_LAT_ACC_SV];
    return GK_OK;
}

// ---- HMC trajectory ---------------------------------------------------------------------------------
extern "C" int gk_latent_hmc_begin(gk_ctx *ctx, gk_latent *l, int target, double step, int nsteps, const double *p0) {
    if (!ctx || !l || !p0) return GK_EINVAL;
    if (target < 0 || target > 1 || nsteps < 1 || !(step > 0)) return gk_fail(ctx, GK_EINVAL, "bad HMC arguments");
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)l->N * l->dim;
    if (target == 0) {
        GK_CUDA(ctx, cudaMemcpyAsync(l->p, p0, n * 8, cudaMemcpyHostToDevice, ctx->stream));
        GK_CUDA(ctx, cudaMemcpyAsync(l->q, l->L, n * 8, cudaMemcpyDeviceToDevice, ctx->stream));
        GK_CUDA(ctx, cudaMemcpyAsync(l->qT, l->LT, n * (l->f64 ? 8 : 4), cudaMemcpyDeviceToDevice, ctx->stream));
    } else {
        GK_CUDA(ctx, cudaMemcpyAsync(l->sc + SC_P_MU0, p0, 8, cudaMemcpyHostToDevice, ctx->stream));
        GK_CUDA(ctx, cudaMemcpyAsync(l->sc + SC_MU0_Q, l->sc + SC_MU0, 8, cudaMemcpyDeviceToDevice, ctx->stream));
    }
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));      // p0 is a host temporary
    l->target = target; l->step = step; l->nsteps = nsteps; l->stage = 0;
    return GK_OK;
}

extern "C" int gk_latent_hmc_pass(gk_ctx *ctx, gk_latent *l, double *acc_dev) {
    if (!ctx || !l || l->target < 0 || l->stage > l->nsteps) return GK_EINVAL;
    const bool value = l->stage == 0 || l->stage == l->nsteps;
    if (l->target == 0) return lat_pass(ctx, l, l->qT, l->sc + SC_MU0, value, true, acc_dev ? acc_dev : l->acc);
    return lat_pass(ctx, l, l->LT, l->sc + SC_MU0_Q, value, false, acc_dev ? acc_dev : l->acc);
}

extern "C" int gk_latent_hmc_advance(gk_ctx *ctx, gk_latent *l, const double *acc_dev) {
    if (!ctx || !l || l->target < 0 || l->stage > l->nsteps) return GK_EINVAL;
    const int kind = l->stage == 0 ? 0 : (l->stage == l->nsteps ? 2 : 1);
    gk_lat_fin f = lat_fin(l, acc_dev ? acc_dev : l->acc);
    if (l->target == 0) {
        if (l->f64) k_lat_advance_L<double><<<1, 1024, 0, ctx->stream>>>(f, l->q, l->p, (double *)l->qT, l->step, kind, 1.0);
        else k_lat_advance_L<float><<<1, 1024, 0, ctx->stream>>>(f, l->q, l->p, (float *)l->qT, l->step, kind, l->scale());
    } else {
        k_lat_advance_mu0<<<1, 32, 0, ctx->stream>>>(f, l->step, kind);
    }
    GK_CHECK_LAUNCH(ctx);
    l->stage++;
    return GK_OK;
}

extern "C" int gk_latent_hmc_end(gk_ctx *ctx, gk_latent *l, double log_u, int *accepted, double *log_ratio) {
    if (!ctx || !l || l->target < 0) return GK_EINVAL;
    if (l->stage != l->nsteps + 1) return gk_fail(ctx, GK_EINVAL, "trajectory not finished: stage %d of %d", l->stage, l->nsteps + 1);
    const size_t n = (size_t)l->N * l->dim;
    if (l->f64) k_lat_end<double><<<1, 1024, 0, ctx->stream>>>(l->sc, log_u, l->target, n, l->q, (const double *)l->qT, l->L, (double *)l->LT);
    else k_lat_end<float><<<1, 1024, 0, ctx->stream>>>(l->sc, log_u, l->target, n, l->q, (const float *)l->qT, l->L, (float *)l->LT);
    GK_CHECK_LAUNCH(ctx);
    double sc[SC_LEN];
    GK_CUDA(ctx, cudaMemcpyAsync(sc, l->sc, sizeof sc, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (accepted) *accepted = sc[SC_ACCEPT] != 0.0;
    if (log_ratio) *log_ratio = sc[SC_RATIO];
    l->target = -1;
    return GK_OK;
}

extern "C" int gk_latent_hmc(gk_ctx *ctx, gk_latent *l, int target, double step, int nsteps, const double *p0, double log_u,
                             int *accepted, double *log_ratio) {
    if (!ctx || !l) return GK_EINVAL;
    const bool sharded = l->net->row0 != 0 || l->net->nrows != l->N;
    if (sharded && !ctx->comm)
        return gk_fail(ctx, GK_EINVAL, "gk_latent_hmc on a row-sharded network needs gk_comm_init (or use the begin/pass/advance/end form)");
    int rc = gk_latent_hmc_begin(ctx, l, target, step, nsteps, p0);
    if (rc) return rc;
    for (int s = 0; s <= nsteps; ++s) {
        // row shards: the accumulator of this rank's pair pass is summed over the ranks in-stream (NCCL) before the update
        if ((rc = gk_latent_hmc_pass(ctx, l, nullptr)) || (sharded && (rc = gk_comm_allreduce_f64(ctx, l->acc, lat_acc_len(l)))) ||
            (rc = gk_latent_hmc_advance(ctx, l, nullptr))) { l->target = -1; return rc; }
    }
    return gk_latent_hmc_end(ctx, l, log_u, accepted, log_ratio);
}
