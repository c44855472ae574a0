// gk_eig.cu -- host side of the eigenmodel path (K3-K6): state, launch plans, C-ABI entry points.
#include <math.h>
#include <string.h>
#include <stdlib.h>
#include <algorithm>
#include "gk_eig_mf.cuh"

struct gk_eig {
    gk_net *net = nullptr;
    int N = 0, D = 0, dtype = GK_F32;
    void *S = nullptr;                 // nrows x ld of dtype
    double *F = nullptr, *lmbda = nullptr, *z = nullptr;   // N x D, D, N x D
    double *acc = nullptr;             // [2][GK_EIG_NBMAX][PA_max]
    double *glob = nullptr;            // GK_EIG_GLOB_LEN
    double *part = nullptr; size_t npart = 0;              // per-tile partial sums of K3
    double *scal = nullptr;            // small device scalars (sum of omega, ...)
    double *out_prec = nullptr, *out_h = nullptr;
    cudaStream_t s2 = nullptr; cudaEvent_t evE[2] = {nullptr, nullptr}, evR[2] = {nullptr, nullptr};   // push overlap (sweep_F_fast)
    void *dphi2 = nullptr; int overlap_push = 1;
    double *G = nullptr, *spart = nullptr; void *dphi = nullptr;   // running-G sweep (gk_eig_fast.cuh): [ld][PA], [NB][PA], [64][PA]
    double *mfMu = nullptr, *mfSig = nullptr, *mfLin = nullptr, *mfPhi = nullptr, *mfSmall = nullptr, *mfPart = nullptr;   // mean-field q (K11)
    long long *prof = nullptr;         // debug cycle counters of k_eig_seq3 (env GK_EIG_PROF=1)
    int use_fast = 1;                  // unsharded sweeps use the running-G form (GK_EIG_OPT_FAST_SWEEP)
    int use_tc = 1;                    // fp32, D = 8, logistic: the full contraction runs on tcgen05 + TMA (GK_EIG_OPT_TC)
    int use_seq5 = 1;                  // fp32 path: register-resident solver k_eig_seq5 (GK_EIG_OPT_SEQ5); 0 = k_eig_seq4
    double mu0 = 0.0;
    bool have_S = false;
    bool chol_f64 = false;             // f32 storage: factorise the D x D conditionals in fp32 unless set
    double inv_sigma_F = 1.0;          // of the open block-form sweep
    int probit = 0;                    // 1 after gk_eig_tn_draw / gk_eig_set_Zsym: the S buffer holds Z + Z^T (probit link)
    uint32_t sweep = 0;
};

#define GK_EIG_PA_MAX (GK_EIG_MAXD * (GK_EIG_MAXD + 1) / 2 + 2 * GK_EIG_MAXD)

int gk_net_refresh_asym(gk_ctx *ctx, gk_net *net) {
    k_eig_asym<<<ctx->num_sms * 8, 256, 0, ctx->stream>>>(net->A, net->At, net->Asym, net->nrows, net->N, net->ld, net->row0);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

int gk_net_ensure_asym(gk_ctx *ctx, gk_net *net) {
    if (net->Asym) return GK_OK;
    size_t bytes = (size_t)net->nrows * net->ld;
    GK_CUDA(ctx, cudaMalloc(&net->Asym, bytes));
    net->bytes += bytes;
    return gk_net_refresh_asym(ctx, net);
}

extern "C" int gk_eig_create(gk_ctx *ctx, gk_net *net, int D, gk_eig **out) {
    if (!ctx || !net || !out) return GK_EINVAL;
    *out = nullptr;
    if (D < 1 || D > GK_EIG_MAXD) return gk_fail(ctx, GK_EUNSUPPORTED, "D=%d: this build supports 1 <= D <= %d", D, GK_EIG_MAXD);
    if (net->row0 % GK_EIG_NBMAX) return gk_fail(ctx, GK_EINVAL, "row shards must start on a multiple of %d", GK_EIG_NBMAX);
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    gk_eig *e = new gk_eig();
    e->net = net; e->N = net->N; e->D = D; e->dtype = net->dtype;
    const size_t N = net->N;
    const int tiles_y = (net->nrows + GK_PG_TILE - 1) / GK_PG_TILE, tiles_x = (net->N + GK_PG_TILE - 1) / GK_PG_TILE;
    e->npart = (size_t)tiles_x * tiles_y;
    cudaError_t err = cudaMalloc(&e->S, (size_t)net->nrows * net->ld * gk_esize(e->dtype));
    if (err == cudaSuccess) err = cudaMalloc(&e->F, N * D * 8);
    if (err == cudaSuccess) err = cudaMalloc(&e->z, N * D * 8);
    if (err == cudaSuccess) err = cudaMalloc(&e->lmbda, GK_EIG_MAXD * 8);
    if (err == cudaSuccess) err = cudaMalloc(&e->acc, (size_t)2 * GK_EIG_NBMAX * GK_EIG_PA_MAX * 8);
    if (err == cudaSuccess) err = cudaMalloc(&e->glob, GK_EIG_GLOB_LEN * 8);
    if (err == cudaSuccess) err = cudaMalloc(&e->part, e->npart * 8);
    if (err == cudaSuccess) err = cudaMalloc(&e->scal, 64);
    if (err != cudaSuccess) {
        cudaFree(e->S); cudaFree(e->F); cudaFree(e->z); cudaFree(e->lmbda); cudaFree(e->acc); cudaFree(e->glob); cudaFree(e->part); cudaFree(e->scal);
        delete e;
        return gk_fail(ctx, GK_ENOMEM, "eigenmodel state alloc: %s", cudaGetErrorString(err));
    }
    GK_CUDA(ctx, cudaMemsetAsync(e->S, 0, (size_t)net->nrows * net->ld * gk_esize(e->dtype), ctx->stream));
    GK_CUDA(ctx, cudaMemsetAsync(e->acc, 0, (size_t)2 * GK_EIG_NBMAX * GK_EIG_PA_MAX * 8, ctx->stream));
    int rc = gk_net_ensure_asym(ctx, net);
    if (rc) return rc;
    *out = e;
    return GK_OK;
}

extern "C" int gk_eig_destroy(gk_ctx *ctx, gk_eig *e) {
    if (!e) return GK_OK;
    if (ctx) { cudaSetDevice(ctx->device); cudaStreamSynchronize(ctx->stream); }
    cudaFree(e->S); cudaFree(e->F); cudaFree(e->z); cudaFree(e->lmbda); cudaFree(e->acc); cudaFree(e->glob);
    cudaFree(e->part); cudaFree(e->scal); cudaFree(e->out_prec); cudaFree(e->out_h);
    cudaFree(e->G); cudaFree(e->dphi); cudaFree(e->dphi2); cudaFree(e->spart); cudaFree(e->prof);
    if (e->s2) cudaStreamDestroy(e->s2);
    for (int q = 0; q < 2; ++q) { if (e->evE[q]) cudaEventDestroy(e->evE[q]); if (e->evR[q]) cudaEventDestroy(e->evR[q]); }
    cudaFree(e->mfMu); cudaFree(e->mfSig); cudaFree(e->mfLin); cudaFree(e->mfPhi); cudaFree(e->mfSmall); cudaFree(e->mfPart);
    delete e;
    return GK_OK;
}

extern "C" size_t gk_eig_device_bytes(gk_eig *e) {
    return e ? (size_t)e->net->nrows * e->net->ld * gk_esize(e->dtype) + (size_t)e->N * e->D * 16 : 0;
}

extern "C" int gk_eig_set_option(gk_ctx *ctx, gk_eig *e, int option, int value) {
    if (!ctx || !e) return GK_EINVAL;
    if (option == GK_EIG_OPT_CHOL_F64) { e->chol_f64 = value != 0; return GK_OK; }
    if (option == GK_EIG_OPT_FAST_SWEEP) { e->use_fast = value != 0; return GK_OK; }
    if (option == GK_EIG_OPT_OVERLAP_PUSH) { e->overlap_push = value != 0; return GK_OK; }
    if (option == GK_EIG_OPT_SEQ5) { e->use_seq5 = value != 0; return GK_OK; }
    if (option == GK_EIG_OPT_TC) { e->use_tc = value != 0; return GK_OK; }
    return gk_fail(ctx, GK_EINVAL, "unknown eigenmodel option %d", option);
}

extern "C" int gk_eig_set_params(gk_ctx *ctx, gk_eig *e, const double *F, double mu0, const double *lmbda) {
    if (!ctx || !e || !F || !lmbda) return GK_EINVAL;
    GK_CUDA(ctx, cudaMemcpyAsync(e->F, F, (size_t)e->N * e->D * 8, cudaMemcpyHostToDevice, ctx->stream));
    GK_CUDA(ctx, cudaMemcpyAsync(e->lmbda, lmbda, (size_t)e->D * 8, cudaMemcpyHostToDevice, ctx->stream));
    e->mu0 = mu0;
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GK_OK;
}

// parity / injection path: S rows (nrows x N, host doubles, S = Omega + Omega^T with any diagonal: it is zeroed)
template <typename T>
__global__ void k_eig_store_S(const double *__restrict__ src, T *__restrict__ S, int nrows, int N, int ld, int row0) {
    size_t total = (size_t)nrows * ld;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int r = (int)(i / ld), j = (int)(i - (size_t)r * ld);
        S[i] = (j < N && j != r + row0) ? (T)src[(size_t)r * N + j] : T(0);
    }
}

extern "C" int gk_eig_set_S(gk_ctx *ctx, gk_eig *e, const double *S_rows) {
    if (!ctx || !e || !S_rows) return GK_EINVAL;
    gk_net *net = e->net;
    size_t bytes = (size_t)net->nrows * net->N * 8;
    double *tmp = (double *)gk_scratch(ctx, 0, bytes);
    if (!tmp) return GK_ENOMEM;
    GK_CUDA(ctx, cudaMemcpyAsync(tmp, S_rows, bytes, cudaMemcpyHostToDevice, ctx->stream));
    if (e->dtype == GK_F32) k_eig_store_S<float><<<ctx->num_sms * 8, 256, 0, ctx->stream>>>(tmp, (float *)e->S, net->nrows, net->N, net->ld, net->row0);
    else k_eig_store_S<double><<<ctx->num_sms * 8, 256, 0, ctx->stream>>>(tmp, (double *)e->S, net->nrows, net->N, net->ld, net->row0);
    GK_CHECK_LAUNCH(ctx);
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    e->have_S = true; e->probit = 0;
    return GK_OK;
}

// probit injection path: rows of Z + Z^T (any diagonal: it is zeroed); switches the state to the probit link
extern "C" int gk_eig_set_Zsym(gk_ctx *ctx, gk_eig *e, const double *Zsym_rows) {
    int rc = gk_eig_set_S(ctx, e, Zsym_rows);
    if (rc == GK_OK) e->probit = 1;
    return rc;
}

template <typename T>
__global__ void k_eig_fetch_S(const T *__restrict__ S, double *__restrict__ dst, int nrows, int N, int ld) {
    size_t total = (size_t)nrows * N;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int r = (int)(i / N), j = (int)(i - (size_t)r * N);
        dst[i] = (double)S[(size_t)r * ld + j];
    }
}

extern "C" int gk_eig_get_S(gk_ctx *ctx, gk_eig *e, double *S_rows) {
    if (!ctx || !e || !S_rows) return GK_EINVAL;
    gk_net *net = e->net;
    size_t bytes = (size_t)net->nrows * net->N * 8;
    double *tmp = (double *)gk_scratch(ctx, 0, bytes);
    if (!tmp) return GK_ENOMEM;
    if (e->dtype == GK_F32) k_eig_fetch_S<float><<<ctx->num_sms * 8, 256, 0, ctx->stream>>>((const float *)e->S, tmp, net->nrows, net->N, net->ld);
    else k_eig_fetch_S<double><<<ctx->num_sms * 8, 256, 0, ctx->stream>>>((const double *)e->S, tmp, net->nrows, net->N, net->ld);
    GK_CHECK_LAUNCH(ctx);
    GK_CUDA(ctx, cudaMemcpyAsync(S_rows, tmp, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GK_OK;
}

// ---------------------------------------------------------------------------------------------
// K3
// ---------------------------------------------------------------------------------------------
extern "C" int gk_eig_pg_draw(gk_ctx *ctx, gk_eig *e, uint64_t seed, double *sum_omega) {
    if (!ctx || !e) return GK_EINVAL;
    gk_net *net = e->net;
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    dim3 grid((net->N + GK_PG_TILE - 1) / GK_PG_TILE, (net->nrows + GK_PG_TILE - 1) / GK_PG_TILE);
    const double bytes = (double)net->nrows * net->N * gk_esize(e->dtype);
    const size_t pg_dyn = 8 * 2 * (GK_PG_TILE * GK_PG_TILE / 256) * 32 * sizeof(unsigned short);   // per-warp draw lists
    GK_CUDA(ctx, cudaFuncSetAttribute(k_eig_pg_tile<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pg_dyn));
    GK_CUDA(ctx, cudaFuncSetAttribute(k_eig_pg_tile<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pg_dyn));
    gk_time_begin(ctx, 2, bytes);
    if (e->dtype == GK_F32)
        k_eig_pg_tile<float><<<grid, 256, pg_dyn, ctx->stream>>>((float *)e->S, net->ld, net->row0, net->nrows, net->N, e->D, e->F, e->lmbda, e->mu0, seed, e->sweep, e->part);
    else
        k_eig_pg_tile<double><<<grid, 256, pg_dyn, ctx->stream>>>((double *)e->S, net->ld, net->row0, net->nrows, net->N, e->D, e->F, e->lmbda, e->mu0, seed, e->sweep, e->part);
    gk_time_end(ctx, 2);
    GK_CHECK_LAUNCH(ctx);
    e->sweep++;
    k_sum_partials<<<1, 1024, 0, ctx->stream>>>(e->part, (int)e->npart, e->scal);
    GK_CHECK_LAUNCH(ctx);
    e->have_S = true; e->probit = 0;
    if (sum_omega) {
        GK_CUDA(ctx, cudaMemcpyAsync(sum_omega, e->scal, 8, cudaMemcpyDeviceToHost, ctx->stream));
        GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return GK_OK;
}

// ---------------------------------------------------------------------------------------------
// K7 (probit): Z[n,m] ~ N(Mu[n,m], 1) truncated to [0, inf) if A[n,m] else (-inf, 0]  (eigenmodel.py:268-287 with
// internals/utils.py:53-76: inverse-CDF map  z = mu - sqrt(2) erfcinv(2 clip(Phi_lb + u (Phi_ub - Phi_lb), 1e-15, 1-1e-15)),
// the 1e-15 clip included), fused with Mu of :75-83.  The uniform of the directed entry (src, dst) is a counter-based hash
// of (seed, sweep, src, dst), so entry (n,m) and its mirror (m,n) -- possibly on another rank -- see the same two draws.
// Stores Z + Z^T (zero diagonal) in the S buffer; part[tile] = sum of this tile's Z[n,m], n != m.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double gk_hash_uniform(uint64_t seed, uint32_t sweep, uint32_t src, uint32_t dst) {
    // the (seed, sweep) key is mixed BEFORE the entry index enters: consecutive seeds must not alias neighbouring entries
    // (seed + 1, dst) == (seed, dst + 1) would replay one iteration's uniforms in the next -- caught by the Geweke test)
    uint64_t x = seed + 0x9E3779B97F4A7C15ull * ((uint64_t)sweep + 1);
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 27; x *= 0x94D049BB133111EBull; x ^= x >> 31;
    x ^= ((uint64_t)src << 32) | (uint64_t)dst;
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 27; x *= 0x94D049BB133111EBull; x ^= x >> 31;
    x += 0x9E3779B97F4A7C15ull;
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 27; x *= 0x94D049BB133111EBull; x ^= x >> 31;
    return ((double)(x >> 11) + 0.5) * (1.0 / 9007199254740992.0);
}
// E[z] of z ~ N(mu, 1) truncated to [0, inf) (edge) or (-inf, 0] (internals/utils.py:78-101, expected_truncnorm).  The reference's
// normaliser is normal_cdf(ub) - normal_cdf(lb) of the RAW bounds (:93), i.e. 1/2 whatever mu: literal = that arithmetic,
// E = mu +- 2 phi(mu); otherwise the truncated-normal mean E = mu +- phi(mu) / Phi(+-mu).
__device__ __forceinline__ double gk_expected_truncnorm(double mu, bool edge, bool literal) {
    const double phi = 0.39894228040143267794 * exp(-0.5 * mu * mu);
    if (literal) return edge ? mu + 2.0 * phi : mu - 2.0 * phi;
    const double cdf = 0.5 * erfc((edge ? -mu : mu) * 0.70710678118654752440);      // Phi(mu) or Phi(-mu)
    const double r = phi / fmax(cdf, 1e-300);
    return edge ? mu + r : mu - r;
}
__device__ __forceinline__ double gk_truncnorm(double mu, bool edge, double u) {
    const double c0 = 0.5 * erfc(mu * 0.70710678118654752440);          // Phi(0 - mu)
    const double cl = edge ? c0 : 0.0, cu = edge ? 1.0 : c0;
    double cdf = cl + u * (cu - cl);
    cdf = fmin(fmax(cdf, 1e-15), 1.0 - 1e-15);
    return mu - 1.41421356237309504880 * erfcinv(2.0 * cdf);
}

// MODE 0: draws (Gibbs, K7); 1 / 2: expectations (mean field, eigenmodel.py:519-534 mf_expected_Z; 1 = the reference's literal normaliser)
template <typename T, int MODE>
__global__ void __launch_bounds__(256) k_eig_tn_tile(T *__restrict__ S, const uint8_t *__restrict__ A, const uint8_t *__restrict__ At,
                                                     int ld, int row0, int nrows, int N, int D, const double *__restrict__ F,
                                                     const double *__restrict__ lmbda, double mu0, uint64_t seed, uint32_t sweep,
                                                     double *__restrict__ part) {
    constexpr int TS = GK_PG_TILE;
    __shared__ double fi[TS][GK_EIG_MAXD], fj[TS][GK_EIG_MAXD];
    __shared__ double red[8];
    const int tid = threadIdx.x, I = row0 + blockIdx.y * TS, J = blockIdx.x * TS;
    for (int t = tid; t < TS * GK_EIG_MAXD; t += 256) {
        int r = t / GK_EIG_MAXD, d = t % GK_EIG_MAXD;
        fi[r][d] = (d < D && I + r < N) ? F[(size_t)(I + r) * D + d] * lmbda[d] : 0.0;
        fj[r][d] = (d < D && J + r < N) ? F[(size_t)(J + r) * D + d] : 0.0;
    }
    __syncthreads();
    const int j = tid & (TS - 1);
    double local = 0.0;
    for (int i = tid / TS; i < TS; i += 256 / TS) {
        const int n = I + i, m = J + j;
        if (n >= N || m >= N || n - row0 >= nrows) continue;
        const size_t off = (size_t)(n - row0) * ld + m;
        if (n == m) { S[off] = T(0); continue; }
        double mu = mu0;
#pragma unroll
        for (int d = 0; d < GK_EIG_MAXD; ++d) mu += fi[i][d] * fj[j][d];
        double z1, z2;
        if constexpr (MODE == 0) {
            z1 = gk_truncnorm(mu, A[off] != 0, gk_hash_uniform(seed, sweep, (uint32_t)n, (uint32_t)m));
            z2 = gk_truncnorm(mu, At[off] != 0, gk_hash_uniform(seed, sweep, (uint32_t)m, (uint32_t)n));
        } else {
            z1 = gk_expected_truncnorm(mu, A[off] != 0, MODE == 1);
            z2 = gk_expected_truncnorm(mu, At[off] != 0, MODE == 1);
        }
        S[off] = (T)(z1 + z2);
        local += z1;
    }
    for (int o = 16; o; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((tid & 31) == 0) red[tid >> 5] = local;
    __syncthreads();
    if (tid == 0) { double s = 0.0; for (int w = 0; w < 8; ++w) s += red[w]; part[blockIdx.y * gridDim.x + blockIdx.x] = s; }
}

extern "C" int gk_eig_tn_draw(gk_ctx *ctx, gk_eig *e, uint64_t seed, double *sum_z) {
    if (!ctx || !e) return GK_EINVAL;
    gk_net *net = e->net;
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    dim3 grid((net->N + GK_PG_TILE - 1) / GK_PG_TILE, (net->nrows + GK_PG_TILE - 1) / GK_PG_TILE);
    gk_time_begin(ctx, 2, (double)net->nrows * net->N * (gk_esize(e->dtype) + 2.0));
    if (e->dtype == GK_F32)
        k_eig_tn_tile<float, 0><<<grid, 256, 0, ctx->stream>>>((float *)e->S, net->A, net->At, net->ld, net->row0, net->nrows, net->N, e->D, e->F, e->lmbda, e->mu0, seed, e->sweep, e->part);
    else
        k_eig_tn_tile<double, 0><<<grid, 256, 0, ctx->stream>>>((double *)e->S, net->A, net->At, net->ld, net->row0, net->nrows, net->N, e->D, e->F, e->lmbda, e->mu0, seed, e->sweep, e->part);
    gk_time_end(ctx, 2);
    GK_CHECK_LAUNCH(ctx);
    e->sweep++;
    k_sum_partials<<<1, 1024, 0, ctx->stream>>>(e->part, (int)e->npart, e->scal);
    GK_CHECK_LAUNCH(ctx);
    e->have_S = true; e->probit = 1;
    if (sum_z) {
        GK_CUDA(ctx, cudaMemcpyAsync(sum_z, e->scal, 8, cudaMemcpyDeviceToHost, ctx->stream));
        GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return GK_OK;
}

// Mean-field probit (eigenmodel.py:570-585 _meanfieldupdate_Z fused with :519-534 mf_expected_Z): E[Z] per directed entry
// from mf_mu_Z = E[mu_0] + E[f_n]^T diag(E[lmbda]) E[f_m]; stores E[Z] + E[Z]^T in the S buffer (probit link).
extern "C" int gk_eig_mf_probit_Z(gk_ctx *ctx, gk_eig *e, const double *mf_mu_F, const double *E_lmbda, double E_mu0,
                                  int exact_normalizer, double *sum_EZ) {
    if (!ctx || !e || !mf_mu_F || !E_lmbda) return GK_EINVAL;
    int rc = gk_eig_set_params(ctx, e, mf_mu_F, E_mu0, E_lmbda);
    if (rc) return rc;
    gk_net *net = e->net;
    dim3 grid((net->N + GK_PG_TILE - 1) / GK_PG_TILE, (net->nrows + GK_PG_TILE - 1) / GK_PG_TILE);
    gk_time_begin(ctx, 2, (double)net->nrows * net->N * (gk_esize(e->dtype) + 2.0));
#define GK_TN_E(T, M) k_eig_tn_tile<T, M><<<grid, 256, 0, ctx->stream>>>((T *)e->S, net->A, net->At, net->ld, net->row0, net->nrows, net->N, e->D, e->F, e->lmbda, e->mu0, 0ull, 0u, e->part)
    if (e->dtype == GK_F32) { if (exact_normalizer) GK_TN_E(float, 2); else GK_TN_E(float, 1); }
    else { if (exact_normalizer) GK_TN_E(double, 2); else GK_TN_E(double, 1); }
    gk_time_end(ctx, 2);
    GK_CHECK_LAUNCH(ctx);
    k_sum_partials<<<1, 1024, 0, ctx->stream>>>(e->part, (int)e->npart, e->scal);
    GK_CHECK_LAUNCH(ctx);
    e->have_S = true; e->probit = 1;
    if (sum_EZ) {
        GK_CUDA(ctx, cudaMemcpyAsync(sum_EZ, e->scal, 8, cudaMemcpyDeviceToHost, ctx->stream));
        GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return GK_OK;
}

// the same inverse-CDF map elementwise, with the uniforms supplied (known-answer tests against the reference's own
// sample_truncnorm, internals/utils.py:53-76; also the injection path of the parity tests)
__global__ void k_truncnorm(const double *__restrict__ mu, const uint8_t *__restrict__ a, const double *__restrict__ u,
                            double *__restrict__ out, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = gk_truncnorm(mu[i], a[i] != 0, u[i]);
}
extern "C" int gk_truncnorm_draw(gk_ctx *ctx, const double *mu, const uint8_t *edge, const double *u, double *out, size_t n) {
    if (!ctx || !mu || !edge || !u || !out) return GK_EINVAL;
    if (n == 0) return GK_OK;
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    char *buf = (char *)gk_scratch(ctx, 0, n * 25 + 1024);
    if (!buf) return GK_ENOMEM;
    double *dmu = (double *)buf, *du = dmu + n, *dout = du + n; uint8_t *da = (uint8_t *)(dout + n);
    GK_CUDA(ctx, cudaMemcpyAsync(dmu, mu, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    GK_CUDA(ctx, cudaMemcpyAsync(du, u, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    GK_CUDA(ctx, cudaMemcpyAsync(da, edge, n, cudaMemcpyHostToDevice, ctx->stream));
    k_truncnorm<<<(int)std::min<size_t>((n + 255) / 256, (size_t)ctx->num_sms * 8), 256, 0, ctx->stream>>>(dmu, da, du, dout, n);
    GK_CHECK_LAUNCH(ctx);
    GK_CUDA(ctx, cudaMemcpyAsync(out, dout, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GK_OK;
}

// The same map with NO allocation on the call path (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA-graph capture, tight serving loops).  The staging scratch that
// gk_truncnorm_draw takes from the context's cudaMalloc'd slots lives in a file-scope device array instead; every GPU that
// loads the library holds these GK_TN_STATIC_BYTES (1.6 MB) until exit.  Layout: mu | u | out (GK_TN_STATIC_MAX doubles each) | edge.
// In HOST code `gk_tn_static` is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress, once per
// device, with that device current.  All calls on a device share the array and it keeps the previous call's contents; nothing
// in it has to start at zero (out[0, n) is fully overwritten before it is read back) -- but two calls that overlap on the
// GPU (two contexts / streams of one device) race on it.
#define GK_TN_STATIC_MAX 65536
#define GK_TN_STATIC_BYTES ((size_t)GK_TN_STATIC_MAX * 25)
static __device__ __align__(256) unsigned char gk_tn_static[GK_TN_STATIC_BYTES];
#define GK_TN_MAX_DEVICES 64
static void *gk_tn_static_ptr[GK_TN_MAX_DEVICES];      // per device: address of that device's copy of gk_tn_static

extern "C" size_t gk_truncnorm_static_max(void) { return GK_TN_STATIC_MAX; }

extern "C" int gk_truncnorm_draw_static(gk_ctx *ctx, const double *mu, const uint8_t *edge, const double *u, double *out, size_t n) {
    if (!ctx || !mu || !edge || !u || !out) return GK_EINVAL;
    if (n > GK_TN_STATIC_MAX)
        return gk_fail(ctx, GK_EINVAL, "gk_truncnorm_draw_static: n=%zu exceeds the static scratch of %d entries (use gk_truncnorm_draw, which allocates)",
                       n, GK_TN_STATIC_MAX);
    if (n == 0) return GK_OK;
    if (ctx->device < 0 || ctx->device >= GK_TN_MAX_DEVICES) return gk_fail(ctx, GK_EUNSUPPORTED, "device %d: static scratch is tracked for %d devices", ctx->device, GK_TN_MAX_DEVICES);
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!gk_tn_static_ptr[ctx->device]) {                      // set-up, once per device, with that device current
        void *p = nullptr;
        GK_CUDA(ctx, cudaGetSymbolAddress(&p, gk_tn_static));
        if (!p) return gk_fail(ctx, GK_ECUDA, "cudaGetSymbolAddress(gk_tn_static) returned a null address");
        gk_tn_static_ptr[ctx->device] = p;
    }
    char *buf = (char *)gk_tn_static_ptr[ctx->device];
    double *dmu = (double *)buf, *du = dmu + GK_TN_STATIC_MAX, *dout = du + GK_TN_STATIC_MAX; uint8_t *da = (uint8_t *)(dout + GK_TN_STATIC_MAX);
    GK_CUDA(ctx, cudaMemcpyAsync(dmu, mu, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    GK_CUDA(ctx, cudaMemcpyAsync(du, u, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    GK_CUDA(ctx, cudaMemcpyAsync(da, edge, n, cudaMemcpyHostToDevice, ctx->stream));
    k_truncnorm<<<(int)std::min<size_t>((n + 255) / 256, (size_t)ctx->num_sms * 8), 256, 0, ctx->stream>>>(dmu, da, du, dout, n);
    GK_CHECK_LAUNCH(ctx);
    GK_CUDA(ctx, cudaMemcpyAsync(out, dout, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GK_OK;
}

// ---------------------------------------------------------------------------------------------
// K4 + K5 (+ K6 statistics): the sequential feature sweep
// ---------------------------------------------------------------------------------------------
template <typename T, int D>
static int launch_slab(gk_ctx *ctx, gk_eig *e, int n0, int nb, double *acc) {
    typedef typename gk_eacc<T>::type ACC;
    gk_net *net = e->net;
    // rows per chunk: a divisor of the block stride, small enough to fill the SMs
    const int stride = gk_eig_nb<T>::NB;
    int mc = stride;
    while (mc > 16 && (net->nrows + mc - 1) / mc < ctx->num_sms) mc >>= 1;
    const int nchunks = (net->nrows + mc - 1) / mc;
    const size_t smem = (size_t)((mc + 3) / 4 * 4) * gk_edim<D>::PA * sizeof(ACC);
    const double bytes = (double)net->nrows * nb * (sizeof(T) + 1.0);
    gk_time_begin(ctx, 3, bytes);
    k_eig_slab<T, D><<<nchunks, GK_EIG_SLAB_THREADS, smem, ctx->stream>>>((const T *)e->S, net->Asym, net->ld, net->row0, net->nrows,
                                                                         e->F, n0, nb, mc, acc, e->probit);
    gk_time_end(ctx, 3);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

template <typename T, typename CH, int D>
static int launch_seq_ch(gk_ctx *ctx, gk_eig *e, int n0, int nb, const void *Sblk, int ldS, const uint8_t *Ablk, int ldA,
                         double inv_sigma_F, double *acc) {
    const size_t smem = gk_eig_seq_smem<T, D>();
    if (gk_attr_needed(ctx, (const void *)k_eig_seq<T, CH, D>)) {
        GK_CUDA(ctx, cudaFuncSetAttribute(k_eig_seq<T, CH, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    gk_eig_seq_params P;
    P.n0 = n0; P.nb = nb; P.N = e->N; P.Sblk = Sblk; P.ldS = ldS; P.Ablk = Ablk; P.ldA = ldA;
    P.F = e->F; P.z = e->z; P.lmbda = e->lmbda; P.mu0 = e->mu0; P.inv_sigma_F = inv_sigma_F;
    P.acc = acc; P.glob = e->glob; P.out_prec = e->out_prec; P.out_h = e->out_h; P.probit = e->probit;
    gk_time_begin(ctx, 5, 0.0);
    k_eig_seq<T, CH, D><<<1, GK_EIG_SEQ_THREADS, smem, ctx->stream>>>(P);
    gk_time_end(ctx, 5);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

template <typename T, int D>
static int launch_seq(gk_ctx *ctx, gk_eig *e, int n0, int nb, const void *Sblk, int ldS, const uint8_t *Ablk, int ldA,
                      double inv_sigma_F, double *acc) {
    if (sizeof(T) == 4 && !e->chol_f64) return launch_seq_ch<T, float, D>(ctx, e, n0, nb, Sblk, ldS, Ablk, ldA, inv_sigma_F, acc);
    return launch_seq_ch<T, double, D>(ctx, e, n0, nb, Sblk, ldS, Ablk, ldA, inv_sigma_F, acc);
}

// block form (row shards): slab partials of one node block / the owner's sequential draw of it
template <typename T, int D>
static int block_partials_T(gk_ctx *ctx, gk_eig *e, int n0, int nb, double *acc) {
    if (n0 % gk_eig_nb<T>::NB || nb > gk_eig_nb<T>::NB) return gk_fail(ctx, GK_EINVAL, "node blocks are %d-aligned", gk_eig_nb<T>::NB);
    return launch_slab<T, D>(ctx, e, n0, nb, acc);
}
template <typename T, int D>
static int block_sample_T(gk_ctx *ctx, gk_eig *e, int n0, int nb, double *acc) {
    gk_net *net = e->net;
    if (n0 < net->row0 || n0 + nb > net->row0 + net->nrows) return gk_fail(ctx, GK_EINVAL, "the sampling rank must own rows [%d,%d)", n0, n0 + nb);
    const T *Sblk = (const T *)e->S + (size_t)(n0 - net->row0) * net->ld + n0;
    const uint8_t *Ablk = net->Asym + (size_t)(n0 - net->row0) * net->ld + n0;
    return launch_seq<T, D>(ctx, e, n0, nb, Sblk, net->ld, Ablk, net->ld, e->inv_sigma_F, acc);
}

template <typename T, int D>
static int sweep_F(gk_ctx *ctx, gk_eig *e, double inv_sigma_F) {
    gk_net *net = e->net;
    const int NB = gk_eig_nb<T>::NB;
    for (int n0 = 0; n0 < e->N; n0 += NB) {
        const int nb = std::min(NB, e->N - n0);
        int rc = launch_slab<T, D>(ctx, e, n0, nb, e->acc);
        if (rc) return rc;
        const T *Sblk = (const T *)e->S + (size_t)(n0 - net->row0) * net->ld + n0;
        const uint8_t *Ablk = net->Asym + (size_t)(n0 - net->row0) * net->ld + n0;
        if ((rc = launch_seq<T, D>(ctx, e, n0, nb, Sblk, net->ld, Ablk, net->ld, inv_sigma_F, e->acc))) return rc;
    }
    return GK_OK;
}

// running-G form of the sweep (gk_eig_fast.cuh): one full contraction, then per node block { serial draws ; push }
template <typename T, typename CH, int D>
static int launch_seq4_ch(gk_ctx *ctx, gk_eig *e, int n0, int nb, double inv_sigma_F) {
    const size_t smem = gk_eig_seq4_smem<T, CH, D>();
    if (gk_attr_needed(ctx, (const void *)k_eig_seq4<T, CH, D>)) {
        GK_CUDA(ctx, cudaFuncSetAttribute(k_eig_seq4<T, CH, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    gk_net *net = e->net;
    gk_eig_seq4_params P;
    P.n0 = n0; P.nb = nb; P.N = e->N;
    P.Sblk = (const T *)e->S + (size_t)n0 * net->ld + n0; P.ldS = net->ld;
    P.Ablk = net->Asym + (size_t)n0 * net->ld + n0; P.ldA = net->ld;
    P.F = e->F; P.z = e->z; P.lmbda = e->lmbda; P.mu0 = e->mu0; P.inv_sigma_F = inv_sigma_F;
    P.G = e->G; P.dphi = e->dphi; P.glob = e->glob; P.out_prec = e->out_prec; P.out_h = e->out_h; P.probit = e->probit;
    P.prof = e->prof;
    gk_time_begin(ctx, 5, 0.0);
    if (e->overlap_push) {
        // programmatic dependent launch: the prologue (tile, feature and normal loads) may run while the preceding
        // k_eig_push_cols is still in flight; the kernel waits (griddepcontrol.wait) before it reads G / the column sums
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(1); cfg.blockDim = dim3(GK_EIG_SEQ_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = ctx->stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        GK_CUDA(ctx, cudaLaunchKernelEx(&cfg, k_eig_seq4<T, CH, D>, P));
    } else {
        k_eig_seq4<T, CH, D><<<1, GK_EIG_SEQ_THREADS, smem, ctx->stream>>>(P);
    }
    gk_time_end(ctx, 5);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

// fp32 storage + fp32 factorisation: the register-resident solver of gk_eig_seq5.cuh (GK_EIG_OPT_SEQ5, default on)
template <int D, bool PROF>
static int launch_seq5_p(gk_ctx *ctx, gk_eig *e, int n0, int nb, double inv_sigma_F) {
    const size_t smem = gk_eig_seq5_smem<D>();
    { static int dbg_set = 0; if (!dbg_set) { dbg_set = 1; const char *ev = getenv("GK_SEQ5_DEBUG"); if (ev) { int v = atoi(ev); cudaMemcpyToSymbol(gk_seq5_debug, &v, sizeof v); } } }
    GK_CUDA(ctx, cudaFuncSetAttribute(k_eig_seq5<D, PROF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    gk_net *net = e->net;
    gk_eig_seq4_params P;
    P.n0 = n0; P.nb = nb; P.N = e->N;
    P.Sblk = (const float *)e->S + (size_t)n0 * net->ld + n0; P.ldS = net->ld;
    P.Ablk = net->Asym + (size_t)n0 * net->ld + n0; P.ldA = net->ld;
    P.F = e->F; P.z = e->z; P.lmbda = e->lmbda; P.mu0 = e->mu0; P.inv_sigma_F = inv_sigma_F;
    P.G = e->G; P.dphi = e->dphi; P.glob = e->glob; P.out_prec = e->out_prec; P.out_h = e->out_h; P.probit = e->probit;
    P.prof = e->prof;
    gk_time_begin(ctx, 5, 0.0);
    if (e->overlap_push) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(1); cfg.blockDim = dim3(GK_SEQ5_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = ctx->stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        GK_CUDA(ctx, cudaLaunchKernelEx(&cfg, k_eig_seq5<D, PROF>, P));
    } else {
        k_eig_seq5<D, PROF><<<1, GK_SEQ5_THREADS, smem, ctx->stream>>>(P);
    }
    gk_time_end(ctx, 5);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}
template <int D>
static int launch_seq5(gk_ctx *ctx, gk_eig *e, int n0, int nb, double inv_sigma_F) {
    return e->prof ? launch_seq5_p<D, true>(ctx, e, n0, nb, inv_sigma_F) : launch_seq5_p<D, false>(ctx, e, n0, nb, inv_sigma_F);
}

// one full contraction G[n] = sum_m X[m,n] Phi_m over the whole network (Phi from e->F, or given rows phi_given)
template <typename T, int D>
static int launch_gfull(gk_ctx *ctx, gk_eig *e, const double *phi_given, size_t *smem_push_out) {
    typedef typename gk_eacc<T>::type ACC;
    constexpr int PA = gk_edim<D>::PA, PH = gk_edim<D>::PH, NB = gk_eig_nb<T>::NB;
    gk_net *net = e->net;
    const int N = e->N, npad = net->ld, strips = npad / GK_EF_COLS;
    if (!e->G) {
        GK_CUDA(ctx, cudaMalloc(&e->G, (size_t)npad * GK_EIG_PA_MAX * 8));
        GK_CUDA(ctx, cudaMalloc(&e->dphi, (size_t)GK_EIG_NBMAX * GK_EIG_PA_MAX * 8));       // ACC [NB][PA]
        GK_CUDA(ctx, cudaMalloc(&e->dphi2, (size_t)GK_EIG_NBMAX * GK_EIG_PA_MAX * 8));      // second buffer: blocks alternate
        GK_CUDA(ctx, cudaStreamCreateWithFlags(&e->s2, cudaStreamNonBlocking));
        for (int q = 0; q < 2; ++q) {
            GK_CUDA(ctx, cudaEventCreateWithFlags(&e->evE[q], cudaEventDisableTiming));
            GK_CUDA(ctx, cudaEventCreateWithFlags(&e->evR[q], cudaEventDisableTiming));
        }
        GK_CUDA(ctx, cudaMalloc(&e->spart, (size_t)64 * GK_EIG_PA_MAX * 8));
        if (getenv("GK_EIG_PROF")) GK_CUDA(ctx, cudaMalloc(&e->prof, 16 * 8));
    }
    if (e->prof) GK_CUDA(ctx, cudaMemsetAsync(e->prof, 0, 16 * 8, ctx->stream));
    // row chunks so that the grid covers >= 2 waves of 2 CTAs/SM
    int nchunks = std::max(1, std::min(16, (4 * ctx->num_sms + strips - 1) / strips));
    constexpr int PR = gk_ef_piece_rows<T>::R;
    int chunk_rows = ((N + nchunks - 1) / nchunks + PR - 1) / PR * PR;
    nchunks = (N + chunk_rows - 1) / chunk_rows;
    ACC *partial = (ACC *)gk_scratch(ctx, 2, (size_t)nchunks * npad * PA * sizeof(ACC));
    if (!partial) return GK_ENOMEM;
    const size_t fold_elems = (size_t)128 * (2 * PH + 1);
    const size_t smem_full = std::max((size_t)2 * PR * GK_EF_COLS * (sizeof(T) + 1) + (size_t)2 * PR * PA * sizeof(ACC), fold_elems * sizeof(ACC));
    const size_t smem_push = std::max((size_t)NB * GK_EF_COLS * (sizeof(T) + 1) + (size_t)NB * PA * sizeof(ACC), fold_elems * sizeof(ACC));
    if (gk_attr_needed(ctx, (const void *)k_eig_gfull<T, D>)) {
        GK_CUDA(ctx, cudaFuncSetAttribute(k_eig_gfull<T, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_full));
        GK_CUDA(ctx, cudaFuncSetAttribute(k_eig_push<T, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_push));
    }
    bool on_tc = false;
    if constexpr (sizeof(T) == 4 && D == 8) {
        // tensor-core form (gk_eig_tc.cu): fp32 storage, logistic link, features given by F; needs the symmetry of an unsharded S
        if (e->use_tc && !phi_given && !e->probit && net->row0 == 0 && net->nrows == N) {
            // partner chunks: the grid (strips x chunks, 2 CTAs per SM) should fill whole waves
            int tc_chunks = 1; double best = 1e30;
            for (int c = 1; c <= 16; ++c) {
                const double waves = (double)strips * c / (2.0 * ctx->num_sms), cost = ceil(waves - 1e-9) / waves * (1.0 + 0.004 * c);
                if ((N + c - 1) / c >= 256 && cost < best - 1e-12) { best = cost; tc_chunks = c; }
            }
            if (const char *ev = getenv("GK_TC_CHUNKS")) tc_chunks = std::max(1, std::min(32, atoi(ev)));     // tuning experiments
            int tc_rows = ((N + tc_chunks - 1) / tc_chunks + 31) / 32 * 32;
            const int tc_n = (N + tc_rows - 1) / tc_rows;
            void *work = gk_scratch(ctx, 9, gk_tc_work_bytes(npad));
            partial = (ACC *)gk_scratch(ctx, 2, (size_t)tc_n * npad * PA * sizeof(ACC));
            if (!work || !partial) return GK_ENOMEM;
            gk_time_begin(ctx, 3, (double)N * N * (sizeof(T) + 1.0));
            int rc = gk_tc_gfull(ctx, (const float *)e->S, net->Asym, net->ld, N, e->F, work, (float *)partial, tc_n, tc_rows);
            gk_time_end(ctx, 3);
            if (rc) return rc;
            nchunks = tc_n;
            on_tc = true;
        }
    }
    if (!on_tc) {
    gk_time_begin(ctx, 3, (double)N * N * (sizeof(T) + 1.0));
    k_eig_gfull<T, D><<<dim3(strips, nchunks), GK_EF_THREADS, smem_full, ctx->stream>>>(
        (const T *)e->S, net->Asym, net->ld, N, e->F, phi_given, chunk_rows, npad, partial, e->probit);
    gk_time_end(ctx, 3);
    GK_CHECK_LAUNCH(ctx);
    }
    const size_t per_chunk = (size_t)npad * PA;
    k_eig_greduce<ACC><<<(int)std::min<size_t>((per_chunk + 255) / 256, (size_t)ctx->num_sms * 8), 256, 0, ctx->stream>>>(
        partial, nchunks, per_chunk, e->G);
    GK_CHECK_LAUNCH(ctx);
    if (smem_push_out) *smem_push_out = smem_push;
    return GK_OK;
}

template <typename T, int D>
static int launch_push(gk_ctx *ctx, gk_eig *e, int n0, int nb, size_t smem_push, cudaStream_t st = nullptr, int skip_cb = -1) {
    typedef typename gk_eacc<T>::type ACC;
    gk_net *net = e->net;
    cudaStream_t keep = ctx->stream;
    if (st) ctx->stream = st;                                  // the timing events must sit on the launching stream
    gk_time_begin(ctx, 3, (double)nb * (e->N - (skip_cb >= 0 ? GK_EF_COLS : 0)) * (sizeof(T) + 1.0));
    k_eig_push<T, D><<<net->ld / GK_EF_COLS, GK_EF_THREADS, smem_push, ctx->stream>>>((const T *)e->S, net->Asym, net->ld, e->N,
                                                                                    (const ACC *)e->dphi, n0, nb, e->G, e->probit, skip_cb);
    gk_time_end(ctx, 3);
    ctx->stream = keep;
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

// the strip of the next node block only (see k_eig_push_cols)
template <typename T, int D>
static int launch_push_cols(gk_ctx *ctx, gk_eig *e, int n0, int nb, int c0, int ncols) {
    typedef typename gk_eacc<T>::type ACC;
    gk_net *net = e->net;
    gk_time_begin(ctx, 3, (double)nb * ncols * (sizeof(T) + 1.0));
    k_eig_push_cols<T, D><<<(ncols + GK_EPC_COLS - 1) / GK_EPC_COLS, 32 * GK_EPC_COLS, 0, ctx->stream>>>(
        (const T *)e->S, net->Asym, net->ld, e->N, (const ACC *)e->dphi, n0, nb, e->G, e->probit, c0, ncols);
    gk_time_end(ctx, 3);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

template <typename T, int D>
static int sweep_F_fast(gk_ctx *ctx, gk_eig *e, double inv_sigma_F) {
    constexpr int NB = gk_eig_nb<T>::NB;
    const int N = e->N;
    size_t smem_push = 0;
    int rc0 = launch_gfull<T, D>(ctx, e, nullptr, &smem_push);
    if (rc0) return rc0;
    int blk = 0;
    for (int n0 = 0; n0 < N; n0 += NB) {
        const int nb = std::min(NB, N - n0);
        int rc;
        if (sizeof(T) == 4 && !e->chol_f64 && e->use_seq5) rc = launch_seq5<D>(ctx, e, n0, nb, inv_sigma_F);
        else rc = (sizeof(T) == 4 && !e->chol_f64) ? launch_seq4_ch<T, float, D>(ctx, e, n0, nb, inv_sigma_F)
                                                   : launch_seq4_ch<T, double, D>(ctx, e, n0, nb, inv_sigma_F);
        if (rc) return rc;
        // Push of this block's change.  Only the strip holding the NEXT block's columns is needed before the next
        // k_eig_seq4; every other strip goes to a second stream and runs under it.  Ordering: rest(b) after seq4(b)
        // (evE); strip pushes and seq4 of later blocks after rest(b-1) (evR) -- so no two kernels ever touch the same
        // rows of G at once, and the dphi buffer of block b is not rewritten (by seq4(b+2)) before rest(b) is done.
        const int next0 = n0 + nb, q = blk & 1;
        if (blk > 0 && e->overlap_push) GK_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, e->evR[q ^ 1], 0));
        if (!e->overlap_push || next0 >= N) {
            if ((rc = launch_push<T, D>(ctx, e, n0, nb, smem_push))) return rc;
        } else {
            const int cbu = next0 / GK_EF_COLS;
            GK_CUDA(ctx, cudaEventRecord(e->evE[q], ctx->stream));
            GK_CUDA(ctx, cudaStreamWaitEvent(e->s2, e->evE[q], 0));
            if ((rc = launch_push<T, D>(ctx, e, n0, nb, smem_push, e->s2, cbu))) return rc;
            GK_CUDA(ctx, cudaEventRecord(e->evR[q], e->s2));
            if ((rc = launch_push_cols<T, D>(ctx, e, n0, nb, cbu * GK_EF_COLS, std::min(GK_EF_COLS, e->net->ld - cbu * GK_EF_COLS)))) return rc;
            std::swap(e->dphi, e->dphi2);
        }
        ++blk;
    }
    k_eig_gstats<D><<<64, 512, 0, ctx->stream>>>(e->G, e->F, N, e->glob, e->probit, e->spart);
    GK_CHECK_LAUNCH(ctx);
    k_eig_gstats_final<D><<<1, 64, 0, ctx->stream>>>(e->spart, 64, e->glob);
    GK_CHECK_LAUNCH(ctx);
    if (e->prof) {
        long long h[16];
        GK_CUDA(ctx, cudaMemcpyAsync(h, e->prof, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
        GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (e->use_seq5 && e->dtype == GK_F32 && !e->chol_f64)
            fprintf(stderr, "[gk_eig seq5 prof] cycles/node: solver loop %.0f = build+factor %.0f, wait base %.0f, load+backsub+store %.0f, fence %.0f, arrive %.0f | finisher: wait f %.0f, phase %.0f (->dphi %.0f, ->tot %.0f, bar4 %.0f) | summer pre-sum %.0f\n",
                    (double)h[0] / N, (double)h[2] / N, (double)h[1] / N, (double)h[3] / N, (double)h[4] / N, (double)h[5] / N, (double)h[8] / N, (double)h[9] / N,
                    (double)h[11] / N, (double)h[12] / N, (double)h[13] / N, (double)h[10] / N);
        else
        fprintf(stderr, "[gk_eig prof] cycles/node: solver busy %.0f barrier wait %.0f | updater(63) busy %.0f barrier wait %.0f\n",
                (double)h[0] / N, (double)h[7] / N, (double)h[8] / N, (double)h[8 + 7] / N);
        fprintf(stderr, "[gk_eig prof] cycles/node: loop %.0f setup %.0f flush %.0f whole kernel %.0f\n", (double)h[3] / N, (double)h[4] / N, (double)h[5] / N, (double)h[6] / N);
    }
    return GK_OK;
}

#define GK_EIG_SWITCH_D(T, D, FN, ...)                                              \
    ((D) == 1 ? FN<T, 1>(__VA_ARGS__) : (D) == 2 ? FN<T, 2>(__VA_ARGS__) : (D) == 3 ? FN<T, 3>(__VA_ARGS__)      \
     : (D) == 4 ? FN<T, 4>(__VA_ARGS__) : (D) == 5 ? FN<T, 5>(__VA_ARGS__) : (D) == 6 ? FN<T, 6>(__VA_ARGS__)    \
     : (D) == 7 ? FN<T, 7>(__VA_ARGS__) : FN<T, 8>(__VA_ARGS__))

extern "C" int gk_eig_resample_F(gk_ctx *ctx, gk_eig *e, double sigma_F, const double *normals, double *F_out,
                                 double *out_prec, double *out_h, double *stats) {
    if (!ctx || !e || !normals) return GK_EINVAL;
    if (!e->have_S) return gk_fail(ctx, GK_EINVAL, "gk_eig_resample_F: call gk_eig_pg_draw or gk_eig_set_S first");
    if (!(sigma_F > 0)) return gk_fail(ctx, GK_EINVAL, "sigma_F must be positive");
    gk_net *net = e->net;
    if (net->row0 != 0 || net->nrows != net->N) {
        // Row shards with the engine's own communicator: per node block, slab partials on every rank -> NCCL all-reduce ->
        // the owner of the block's rows draws it -> NCCL broadcast of the block's features; the pair statistics, which
        // accumulate on whichever rank drew a block, are summed at the end.  All in-stream, no host round trips.
        if (!ctx->comm) return gk_fail(ctx, GK_EINVAL, "gk_eig_resample_F on a row-sharded network needs gk_comm_init (or use the block form)");
        if (out_prec || out_h) return gk_fail(ctx, GK_EUNSUPPORTED, "per-node conditionals are available on single-rank networks only");
        const int world = gk_comm_world(ctx), rank = gk_comm_rank(ctx);
        int rc = gk_eig_sweep_begin(ctx, e, sigma_F, normals);
        if (rc) return rc;
        const int bn = gk_eig_block_nodes(e);
        double *acc = (double *)gk_scratch(ctx, 10, gk_eig_acc_len(e) * 8);
        double *fblk = (double *)gk_scratch(ctx, 11, (size_t)bn * e->D * 8);
        if (!acc || !fblk) return GK_ENOMEM;
        for (int rk = 0; rk < world; ++rk) {
            int s0, s1;
            gk_comm_row_range(e->N, rk, world, &s0, &s1);
            if (rk == rank && (s0 != net->row0 || s1 != net->row0 + net->nrows))
                return gk_fail(ctx, GK_EINVAL, "rank %d of %d must hold rows [%d, %d)", rank, world, s0, s1);
            for (int n0 = s0; n0 < s1; n0 += bn) {
                const int nb = std::min(bn, s1 - n0);
                if ((rc = gk_eig_block_partials(ctx, e, n0, nb, acc))) return rc;
                if ((rc = gk_comm_allreduce_f64(ctx, acc, gk_eig_acc_len(e)))) return rc;
                if (rank == rk && (rc = gk_eig_block_sample(ctx, e, n0, nb, acc, fblk))) return rc;
                if ((rc = gk_comm_broadcast(ctx, fblk, (size_t)nb * e->D * 8, rk))) return rc;
                if (rank != rk && (rc = gk_eig_set_block(ctx, e, n0, nb, fblk))) return rc;
            }
        }
        if ((rc = gk_comm_allreduce_f64(ctx, e->glob + GK_EIG_GLOB_PQ, GK_EIG_GLOB_LEN - GK_EIG_GLOB_PQ))) return rc;
        return gk_eig_sweep_end(ctx, e, F_out, stats);
    }
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t N = e->N; const int D = e->D;
    GK_CUDA(ctx, cudaMemcpyAsync(e->z, normals, N * D * 8, cudaMemcpyHostToDevice, ctx->stream));
    if (out_prec && !e->out_prec) GK_CUDA(ctx, cudaMalloc(&e->out_prec, N * D * D * 8));
    if (out_h && !e->out_h) GK_CUDA(ctx, cudaMalloc(&e->out_h, N * D * 8));
    double *keep_prec = e->out_prec, *keep_h = e->out_h;
    if (!out_prec) e->out_prec = nullptr;
    if (!out_h) e->out_h = nullptr;
    GK_CUDA(ctx, cudaMemsetAsync(e->glob, 0, GK_EIG_GLOB_LEN * 8, ctx->stream));
    k_eig_colsum<<<1, 1024, 0, ctx->stream>>>(e->F, e->N, D, e->glob + GK_EIG_GLOB_SUMF);
    GK_CHECK_LAUNCH(ctx);
    int rc;
    if (e->use_fast) rc = e->dtype == GK_F32 ? GK_EIG_SWITCH_D(float, D, sweep_F_fast, ctx, e, 1.0 / sigma_F)
                                             : GK_EIG_SWITCH_D(double, D, sweep_F_fast, ctx, e, 1.0 / sigma_F);
    else rc = e->dtype == GK_F32 ? GK_EIG_SWITCH_D(float, D, sweep_F, ctx, e, 1.0 / sigma_F)
                                 : GK_EIG_SWITCH_D(double, D, sweep_F, ctx, e, 1.0 / sigma_F);
    e->out_prec = keep_prec; e->out_h = keep_h;
    if (rc) return rc;
    if (F_out) GK_CUDA(ctx, cudaMemcpyAsync(F_out, e->F, N * D * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_prec) GK_CUDA(ctx, cudaMemcpyAsync(out_prec, keep_prec, N * D * D * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_h) GK_CUDA(ctx, cudaMemcpyAsync(out_h, keep_h, N * D * 8, cudaMemcpyDeviceToHost, ctx->stream));
    double g[GK_EIG_GLOB_LEN];
    GK_CUDA(ctx, cudaMemcpyAsync(g, e->glob, sizeof g, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (stats) {
        // stats: Pq[D*D] (full symmetric) | hS[D] | hK[D]
        for (int a = 0; a < D; ++a)
            for (int b = 0; b < D; ++b) stats[a * D + b] = g[GK_EIG_GLOB_PQ + gk_qidx(D, std::min(a, b), std::max(a, b))];
        for (int a = 0; a < D; ++a) { stats[D * D + a] = g[GK_EIG_GLOB_HS + a]; stats[D * D + D + a] = g[GK_EIG_GLOB_HK + a]; }
    }
    return GK_OK;
}

// ---------------------------------------------------------------------------------------------
// block form of the same sweep, for row shards: begin; per node block { partials on every rank ->
// all-reduce(acc_dev) -> sample on the owner -> broadcast(fblk_dev) -> set_block on the others }; end.
// The pair statistics accumulate on whichever rank draws a block, so the caller sums `stats` over the ranks.
// ---------------------------------------------------------------------------------------------
__global__ void k_eig_set_block(double *__restrict__ F, const double *__restrict__ fblk, int n0, int nb, int D, double *__restrict__ glob) {
    __shared__ double dsum[GK_EIG_MAXD], dnew[GK_EIG_MAXD];
    if (threadIdx.x < GK_EIG_MAXD) { dsum[threadIdx.x] = 0.0; dnew[threadIdx.x] = 0.0; }
    __syncthreads();
    for (int t = threadIdx.x; t < nb * D; t += blockDim.x) {
        const int d = t % D;
        const double o = F[(size_t)n0 * D + t], v = fblk[t];
        F[(size_t)n0 * D + t] = v;
        atomicAdd(&dsum[d], v - o); atomicAdd(&dnew[d], v);
    }
    __syncthreads();
    if (threadIdx.x < D) { glob[GK_EIG_GLOB_SUMF + threadIdx.x] += dsum[threadIdx.x]; glob[GK_EIG_GLOB_PREF + threadIdx.x] += dnew[threadIdx.x]; }
}

extern "C" int gk_eig_sweep_begin(gk_ctx *ctx, gk_eig *e, double sigma_F, const double *normals) {
    if (!ctx || !e || !normals) return GK_EINVAL;
    if (!e->have_S) return gk_fail(ctx, GK_EINVAL, "gk_eig_sweep_begin: call gk_eig_pg_draw or gk_eig_set_S first");
    if (!(sigma_F > 0)) return gk_fail(ctx, GK_EINVAL, "sigma_F must be positive");
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    GK_CUDA(ctx, cudaMemcpyAsync(e->z, normals, (size_t)e->N * e->D * 8, cudaMemcpyHostToDevice, ctx->stream));
    GK_CUDA(ctx, cudaMemsetAsync(e->glob, 0, GK_EIG_GLOB_LEN * 8, ctx->stream));
    k_eig_colsum<<<1, 1024, 0, ctx->stream>>>(e->F, e->N, e->D, e->glob + GK_EIG_GLOB_SUMF);
    GK_CHECK_LAUNCH(ctx);
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));       // normals is a host temporary
    e->inv_sigma_F = 1.0 / sigma_F;
    return GK_OK;
}
extern "C" size_t gk_eig_acc_len(gk_eig *e) { return e ? (size_t)2 * GK_EIG_NBMAX * GK_EIG_PA_MAX : 0; }
extern "C" int gk_eig_block_nodes(gk_eig *e) { return !e ? 0 : (e->dtype == GK_F32 ? gk_eig_nb<float>::NB : gk_eig_nb<double>::NB); }

extern "C" int gk_eig_block_partials(gk_ctx *ctx, gk_eig *e, int n0, int nb, double *acc_dev) {
    if (!ctx || !e || !acc_dev || n0 < 0 || nb <= 0 || n0 + nb > e->N) return GK_EINVAL;
    GK_CUDA(ctx, cudaMemsetAsync(acc_dev, 0, gk_eig_acc_len(e) * 8, ctx->stream));
    return e->dtype == GK_F32 ? GK_EIG_SWITCH_D(float, e->D, block_partials_T, ctx, e, n0, nb, acc_dev)
                              : GK_EIG_SWITCH_D(double, e->D, block_partials_T, ctx, e, n0, nb, acc_dev);
}
extern "C" int gk_eig_block_sample(gk_ctx *ctx, gk_eig *e, int n0, int nb, double *acc_dev, double *fblk_dev) {
    if (!ctx || !e || !acc_dev || n0 < 0 || nb <= 0 || n0 + nb > e->N) return GK_EINVAL;
    double *kp = e->out_prec, *kh = e->out_h;
    e->out_prec = nullptr; e->out_h = nullptr;
    int rc = e->dtype == GK_F32 ? GK_EIG_SWITCH_D(float, e->D, block_sample_T, ctx, e, n0, nb, acc_dev)
                                : GK_EIG_SWITCH_D(double, e->D, block_sample_T, ctx, e, n0, nb, acc_dev);
    e->out_prec = kp; e->out_h = kh;
    if (rc) return rc;
    if (fblk_dev) GK_CUDA(ctx, cudaMemcpyAsync(fblk_dev, e->F + (size_t)n0 * e->D, (size_t)nb * e->D * 8, cudaMemcpyDeviceToDevice, ctx->stream));
    return GK_OK;
}
extern "C" int gk_eig_set_block(gk_ctx *ctx, gk_eig *e, int n0, int nb, const double *fblk_dev) {
    if (!ctx || !e || !fblk_dev || n0 < 0 || nb <= 0 || n0 + nb > e->N) return GK_EINVAL;
    k_eig_set_block<<<1, 256, 0, ctx->stream>>>(e->F, fblk_dev, n0, nb, e->D, e->glob);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}
extern "C" int gk_eig_sweep_end(gk_ctx *ctx, gk_eig *e, double *F_out, double *stats) {
    if (!ctx || !e) return GK_EINVAL;
    const int D = e->D;
    if (F_out) GK_CUDA(ctx, cudaMemcpyAsync(F_out, e->F, (size_t)e->N * D * 8, cudaMemcpyDeviceToHost, ctx->stream));
    double g[GK_EIG_GLOB_LEN];
    GK_CUDA(ctx, cudaMemcpyAsync(g, e->glob, sizeof g, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (stats) {
        for (int a = 0; a < D; ++a)
            for (int b = 0; b < D; ++b) stats[a * D + b] = g[GK_EIG_GLOB_PQ + gk_qidx(D, std::min(a, b), std::max(a, b))];
        for (int a = 0; a < D; ++a) { stats[D * D + a] = g[GK_EIG_GLOB_HS + a]; stats[D * D + D + a] = g[GK_EIG_GLOB_HK + a]; }
    }
    return GK_OK;
}

// ---------------------------------------------------------------------------------------------
// K9 (eigenmodel form): log likelihood  sum A log P + (1-A) log(1-P),  P = clip(logistic(Mu), 1e-16, 1-1e-16)
// (eigenmodel.py:110-129 with P of :803-811; link = 1 -> probit P of :227-236).  fp64 math, fixed order.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_eig_loglik(const uint8_t *__restrict__ A, int ld, int row0, int nrows, int N, int D,
                                                    const double *__restrict__ F, const double *__restrict__ lmbda, double mu0,
                                                    int link, double *__restrict__ part) {
    constexpr int TS = 64;
    __shared__ double fi[TS][GK_EIG_MAXD], fj[TS][GK_EIG_MAXD];
    __shared__ double red[8];
    const int tid = threadIdx.x, I = row0 + blockIdx.y * TS, J = blockIdx.x * TS;
    for (int t = tid; t < TS * GK_EIG_MAXD; t += 256) {
        int r = t / GK_EIG_MAXD, d = t % GK_EIG_MAXD;
        fi[r][d] = (d < D && I + r < N) ? F[(size_t)(I + r) * D + d] * lmbda[d] : 0.0;
        fj[r][d] = (d < D && J + r < N) ? F[(size_t)(J + r) * D + d] : 0.0;
    }
    __syncthreads();
    const int j = tid & (TS - 1);
    double local = 0.0;
    for (int i = tid / TS; i < TS; i += 256 / TS) {
        const int n = I + i, m = J + j;
        if (n >= N || m >= N || n - row0 >= nrows) continue;
        double z = mu0;
        if (n != m) for (int d = 0; d < GK_EIG_MAXD; ++d) z += fi[i][d] * fj[j][d];
        double p = link == 0 ? 1.0 / (1.0 + exp(-z)) : 0.5 * erfc(-z * 0.70710678118654752440);
        p = fmin(fmax(p, 1e-16), 1.0 - 1e-16);
        local += A[(size_t)(n - row0) * ld + m] ? log(p) : log(1.0 - p);
    }
    for (int o = 16; o; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((tid & 31) == 0) red[tid >> 5] = local;
    __syncthreads();
    if (tid == 0) { double s = 0.0; for (int w = 0; w < 8; ++w) s += red[w]; part[blockIdx.y * gridDim.x + blockIdx.x] = s; }
}

extern "C" int gk_eig_log_likelihood(gk_ctx *ctx, gk_eig *e, int link, double *ll) {
    if (!ctx || !e || !ll || link < 0 || link > 1) return GK_EINVAL;
    gk_net *net = e->net;
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    dim3 grid((net->N + 63) / 64, (net->nrows + 63) / 64);
    k_eig_loglik<<<grid, 256, 0, ctx->stream>>>(net->A, net->ld, net->row0, net->nrows, net->N, e->D, e->F, e->lmbda, e->mu0, link, e->part);
    GK_CHECK_LAUNCH(ctx);
    k_sum_partials<<<1, 1024, 0, ctx->stream>>>(e->part, (int)e->npart, e->scal + 1);
    GK_CHECK_LAUNCH(ctx);
    GK_CUDA(ctx, cudaMemcpyAsync(ll, e->scal + 1, 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GK_OK;
}

// ---------------------------------------------------------------------------------------------
// K11: mean-field logistic eigenmodel (eigenmodel.py:998-1157).  Unsharded networks.
// Device state of q: e->mfMu (N x D), e->mfSig (N x D x D), e->mfPhi (N x PA = E[ff^T] packed | lin | lin).
// ---------------------------------------------------------------------------------------------
static int mf_ensure(gk_ctx *ctx, gk_eig *e) {
    if (e->net->row0 != 0 || e->net->nrows != e->net->N) return gk_fail(ctx, GK_EINVAL, "the mean-field eigenmodel path needs an unsharded network");
    if (!e->mfMu) {
        const size_t N = e->N, D = e->D;
        GK_CUDA(ctx, cudaMalloc(&e->mfMu, N * D * 8));
        GK_CUDA(ctx, cudaMalloc(&e->mfSig, N * D * D * 8));
        GK_CUDA(ctx, cudaMalloc(&e->mfLin, N * D * 8));
        GK_CUDA(ctx, cudaMalloc(&e->mfPhi, N * GK_EIG_PA_MAX * 8));
        GK_CUDA(ctx, cudaMalloc(&e->mfSmall, 256 * 8));
        const size_t nt = (N + GK_MFE_TILE - 1) / GK_MFE_TILE;
        GK_CUDA(ctx, cudaMalloc(&e->mfPart, 2 * nt * nt * 8));
    }
    return GK_OK;
}

template <typename T, int D>
static int mf_build_phi(gk_ctx *ctx, gk_eig *e, const double *mu, const double *Sig, const double *Flin) {
    const size_t N = e->N;
    GK_CUDA(ctx, cudaMemcpyAsync(e->mfMu, mu, N * D * 8, cudaMemcpyHostToDevice, ctx->stream));
    GK_CUDA(ctx, cudaMemcpyAsync(e->mfSig, Sig, N * D * D * 8, cudaMemcpyHostToDevice, ctx->stream));
    if (Flin) GK_CUDA(ctx, cudaMemcpyAsync(e->mfLin, Flin, N * D * 8, cudaMemcpyHostToDevice, ctx->stream));
    k_eig_mf_phi<D><<<ctx->num_sms * 4, 256, 0, ctx->stream>>>(e->mfMu, e->mfSig, Flin ? e->mfLin : nullptr, e->N, e->mfPhi);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

template <typename T, int D>
static int mf_omega_T(gk_ctx *ctx, gk_eig *e, const double *mu, const double *Sig, const double *wq, const double *lam, double mu0,
                      double sig_mu0, int n_samples, uint64_t seed, double *sums) {
    constexpr int Q = gk_edim<D>::Q;
    int rc = mf_build_phi<T, D>(ctx, e, mu, Sig, nullptr);
    if (rc) return rc;
    GK_CUDA(ctx, cudaMemcpyAsync(e->mfSmall, wq, Q * 8, cudaMemcpyHostToDevice, ctx->stream));
    GK_CUDA(ctx, cudaMemcpyAsync(e->mfSmall + 64, lam, D * 8, cudaMemcpyHostToDevice, ctx->stream));
    const int nt = (e->N + GK_MFE_TILE - 1) / GK_MFE_TILE;
    gk_time_begin(ctx, 2, (double)e->N * e->N * sizeof(T));
    k_eig_mf_omega<T, D><<<dim3(nt, nt), 256, 0, ctx->stream>>>((T *)e->S, e->net->ld, e->N, e->mfPhi, e->mfSmall, e->mfSmall + 64, mu0,
                                                                 sig_mu0, n_samples, seed, e->mfPart);
    gk_time_end(ctx, 2);
    GK_CHECK_LAUNCH(ctx);
    k_sum_partials<<<1, 1024, 0, ctx->stream>>>(e->mfPart, nt * nt, e->mfSmall + 128);
    GK_CHECK_LAUNCH(ctx);
    k_sum_partials<<<1, 1024, 0, ctx->stream>>>(e->mfPart + (size_t)nt * nt, nt * nt, e->mfSmall + 129);
    GK_CHECK_LAUNCH(ctx);
    GK_CUDA(ctx, cudaMemcpyAsync(sums, e->mfSmall + 128, 16, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    e->have_S = true; e->probit = 0;
    return GK_OK;
}

extern "C" int gk_eig_mf_omega(gk_ctx *ctx, gk_eig *e, const double *mf_mu_F, const double *mf_Sigma_F, const double *E_llT,
                               const double *E_lmbda, double mf_mu_mu0, double mf_sigma_mu0, int n_samples, uint64_t seed,
                               double *sum_EO, double *sum_EO_FdotF) {
    if (!ctx || !e || !mf_mu_F || !mf_Sigma_F || !E_llT || !E_lmbda) return GK_EINVAL;
    if (n_samples < 1) return gk_fail(ctx, GK_EINVAL, "n_samples must be positive");
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = mf_ensure(ctx, e);
    if (rc) return rc;
    rc = gk_net_ensure_asym(ctx, e->net);
    if (rc) return rc;
    const int D = e->D;
    double wq[GK_EIG_MAXD * (GK_EIG_MAXD + 1) / 2], sums[2];
    for (int a = 0; a < D; ++a)
        for (int b = a; b < D; ++b) wq[gk_qidx(D, a, b)] = (a == b ? 1.0 : 2.0) * 0.5 * (E_llT[a * D + b] + E_llT[b * D + a]);
    rc = e->dtype == GK_F32 ? GK_EIG_SWITCH_D(float, D, mf_omega_T, ctx, e, mf_mu_F, mf_Sigma_F, wq, E_lmbda, mf_mu_mu0, mf_sigma_mu0, n_samples, seed, sums)
                            : GK_EIG_SWITCH_D(double, D, mf_omega_T, ctx, e, mf_mu_F, mf_Sigma_F, wq, E_lmbda, mf_mu_mu0, mf_sigma_mu0, n_samples, seed, sums);
    if (rc) return rc;
    if (sum_EO) *sum_EO = sums[0];
    if (sum_EO_FdotF) *sum_EO_FdotF = sums[1];
    return GK_OK;
}

template <typename T, int D>
static int mf_update_F_T(gk_ctx *ctx, gk_eig *e, double *mu, double *Sig, const double *Flin, const double *llT, const double *lam,
                         double mu0, double sigma_F, double stepsize, double *out_prec, double *out_h) {
    constexpr int NB = gk_eig_nb<T>::NB;
    const size_t N = e->N;
    gk_net *net = e->net;
    int rc = mf_build_phi<T, D>(ctx, e, mu, Sig, Flin);
    if (rc) return rc;
    GK_CUDA(ctx, cudaMemcpyAsync(e->mfSmall, llT, D * D * 8, cudaMemcpyHostToDevice, ctx->stream));
    GK_CUDA(ctx, cudaMemcpyAsync(e->mfSmall + 64, lam, D * 8, cudaMemcpyHostToDevice, ctx->stream));
    if (out_prec && !e->out_prec) GK_CUDA(ctx, cudaMalloc(&e->out_prec, N * D * D * 8));
    if (out_h && !e->out_h) GK_CUDA(ctx, cudaMalloc(&e->out_h, N * D * 8));
    // running column sums of the linear features (the "- sum_m f_m" of kappa + kappa^T = Asym - 1)
    k_eig_colsum<<<1, 1024, 0, ctx->stream>>>(Flin ? e->mfLin : e->mfMu, e->N, D, e->glob + GK_EIG_GLOB_SUMF);
    GK_CHECK_LAUNCH(ctx);
    size_t smem_push = 0;
    if ((rc = launch_gfull<T, D>(ctx, e, e->mfPhi, &smem_push))) return rc;
    const size_t smem = gk_eig_mfseq_smem<T, D>();
    if (gk_attr_needed(ctx, (const void *)k_eig_mf_seq<T, D>)) {
        GK_CUDA(ctx, cudaFuncSetAttribute(k_eig_mf_seq<T, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    for (int n0 = 0; n0 < e->N; n0 += NB) {
        const int nb = std::min(NB, e->N - n0);
        gk_eig_mfseq_params P;
        P.n0 = n0; P.nb = nb; P.N = e->N;
        P.Sblk = (const T *)e->S + (size_t)n0 * net->ld + n0; P.ldS = net->ld;
        P.Ablk = net->Asym + (size_t)n0 * net->ld + n0; P.ldA = net->ld;
        P.G = e->G; P.Phi = e->mfPhi; P.mu = e->mfMu; P.Sig = e->mfSig; P.llT = e->mfSmall; P.lam = e->mfSmall + 64;
        P.mu0 = mu0; P.inv_sigma_F = 1.0 / sigma_F; P.stepsize = stepsize; P.track = Flin ? 0 : 1;
        P.dphi = e->dphi; P.glob = e->glob; P.out_prec = out_prec ? e->out_prec : nullptr; P.out_h = out_h ? e->out_h : nullptr;
        P.probit = e->probit;
        gk_time_begin(ctx, 5, 0.0);
        k_eig_mf_seq<T, D><<<1, 256, smem, ctx->stream>>>(P);
        gk_time_end(ctx, 5);
        GK_CHECK_LAUNCH(ctx);
        if ((rc = launch_push<T, D>(ctx, e, n0, nb, smem_push))) return rc;
    }
    GK_CUDA(ctx, cudaMemcpyAsync(mu, e->mfMu, N * D * 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaMemcpyAsync(Sig, e->mfSig, N * D * D * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_prec) GK_CUDA(ctx, cudaMemcpyAsync(out_prec, e->out_prec, N * D * D * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_h) GK_CUDA(ctx, cudaMemcpyAsync(out_h, e->out_h, N * D * 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GK_OK;
}

extern "C" int gk_eig_mf_update_F(gk_ctx *ctx, gk_eig *e, double *mf_mu_F, double *mf_Sigma_F, const double *F_lin,
                                  const double *E_llT, const double *E_lmbda, double E_mu0, double sigma_F, double stepsize,
                                  double *out_prec, double *out_h) {
    if (!ctx || !e || !mf_mu_F || !mf_Sigma_F || !E_llT || !E_lmbda) return GK_EINVAL;
    if (!e->have_S) return gk_fail(ctx, GK_EINVAL, "gk_eig_mf_update_F: call gk_eig_mf_omega, gk_eig_mf_probit_Z or gk_eig_set_S first");
    if (!(sigma_F > 0) || !(stepsize >= 0.0 && stepsize <= 1.0)) return gk_fail(ctx, GK_EINVAL, "bad sigma_F / stepsize");
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = mf_ensure(ctx, e);
    if (rc) return rc;
    const int D = e->D;
    return e->dtype == GK_F32 ? GK_EIG_SWITCH_D(float, D, mf_update_F_T, ctx, e, mf_mu_F, mf_Sigma_F, F_lin, E_llT, E_lmbda, E_mu0, sigma_F, stepsize, out_prec, out_h)
                              : GK_EIG_SWITCH_D(double, D, mf_update_F_T, ctx, e, mf_mu_F, mf_Sigma_F, F_lin, E_llT, E_lmbda, E_mu0, sigma_F, stepsize, out_prec, out_h);
}

template <typename T, int D>
static int mf_pair_stats_T(gk_ctx *ctx, gk_eig *e, const double *mu, const double *Sig, double *stats) {
    int rc = mf_build_phi<T, D>(ctx, e, mu, Sig, nullptr);
    if (rc) return rc;
    k_eig_colsum<<<1, 1024, 0, ctx->stream>>>(e->mfMu, e->N, D, e->glob + GK_EIG_GLOB_SUMF);
    GK_CHECK_LAUNCH(ctx);
    if ((rc = launch_gfull<T, D>(ctx, e, e->mfPhi, nullptr))) return rc;
    k_eig_gstats_phi<D><<<64, 512, 0, ctx->stream>>>(e->G, e->mfPhi, e->N, e->glob, e->spart, e->probit);
    GK_CHECK_LAUNCH(ctx);
    k_eig_gstats_final<D><<<1, 64, 0, ctx->stream>>>(e->spart, 64, e->glob);
    GK_CHECK_LAUNCH(ctx);
    double g[GK_EIG_GLOB_LEN];
    GK_CUDA(ctx, cudaMemcpyAsync(g, e->glob, sizeof g, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int a = 0; a < D; ++a)
        for (int b = 0; b < D; ++b) stats[a * D + b] = g[GK_EIG_GLOB_PQ + gk_qidx(D, std::min(a, b), std::max(a, b))];
    for (int a = 0; a < D; ++a) { stats[D * D + a] = g[GK_EIG_GLOB_HS + a]; stats[D * D + D + a] = g[GK_EIG_GLOB_HK + a]; }
    return GK_OK;
}

extern "C" int gk_eig_mf_pair_stats(gk_ctx *ctx, gk_eig *e, const double *mf_mu_F, const double *mf_Sigma_F, double *stats) {
    if (!ctx || !e || !mf_mu_F || !mf_Sigma_F || !stats) return GK_EINVAL;
    if (!e->have_S) return gk_fail(ctx, GK_EINVAL, "gk_eig_mf_pair_stats: call gk_eig_mf_omega, gk_eig_mf_probit_Z or gk_eig_set_S first");
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = mf_ensure(ctx, e);
    if (rc) return rc;
    const int D = e->D;
    return e->dtype == GK_F32 ? GK_EIG_SWITCH_D(float, D, mf_pair_stats_T, ctx, e, mf_mu_F, mf_Sigma_F, stats)
                              : GK_EIG_SWITCH_D(double, D, mf_pair_stats_T, ctx, e, mf_mu_F, mf_Sigma_F, stats);
}
