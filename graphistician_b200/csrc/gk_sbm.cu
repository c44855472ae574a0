// gk_sbm.cu -- host side of the SBM path: coefficient tables, launch plans, C-ABI entry points.
#include <math.h>
#include <string.h>
#include <algorithm>
#include "gk_sbm_seq.cuh"

struct gk_sweep {
    int K = 0, S = 0, N = 0, sampler = 0;
    // device pointers (carved from ctx scratch)
    double *tables = nullptr; gk_tables T{};
    int32_t *c = nullptr; uint8_t *c8 = nullptr; int *sizes = nullptr; unsigned long long *nchanged = nullptr;
    double *u = nullptr, *out_lp = nullptr, *lp = nullptr, *stats = nullptr;
    bool want_lp = false;
};

#define GK_LOG_ZERO (-1e10)
#define GK_CELL_MOVER_CAP 1024            // sweeps that moved more nodes than this leave no K1 table behind (a full pass is cheaper)
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// sum_edges log N(w; mu, Sigma) = n*c0 + c1.sum_w + sum_{b<=b'} C2s[b,b'] sum_ww[b,b']   (Cholesky of the symmetrised Sigma, B <= 4)
static bool gauss_coef(int B, const double *mu, const double *sg, double *c0, double *c1, double *C2s) {
    const double LOG2PI = 1.8378770664093453;
    double L[16] = {0}, Li[16] = {0}, P[16];
    for (int i = 0; i < B; ++i)
        for (int j = 0; j <= i; ++j) {
            double v = 0.5 * (sg[i * B + j] + sg[j * B + i]);
            for (int k = 0; k < j; ++k) v -= L[i * B + k] * L[j * B + k];
            if (i == j) { if (!(v > 0)) return false; L[i * B + i] = sqrt(v); }
            else L[i * B + j] = v / L[j * B + j];
        }
    double logdet = 0.0;
    for (int i = 0; i < B; ++i) logdet += 2.0 * log(L[i * B + i]);
    for (int i = 0; i < B; ++i) {                       // Li = L^-1 (lower)
        Li[i * B + i] = 1.0 / L[i * B + i];
        for (int j = 0; j < i; ++j) {
            double v = 0.0;
            for (int k = j; k < i; ++k) v -= L[i * B + k] * Li[k * B + j];
            Li[i * B + j] = v / L[i * B + i];
        }
    }
    for (int i = 0; i < B; ++i)                         // P = Sigma^-1 = Li^T Li
        for (int j = 0; j < B; ++j) {
            double v = 0.0;
            for (int k = (i > j ? i : j); k < B; ++k) v += Li[k * B + i] * Li[k * B + j];
            P[i * B + j] = v;
        }
    double quad = 0.0;
    for (int i = 0; i < B; ++i) {
        c1[i] = 0.0;
        for (int j = 0; j < B; ++j) c1[i] += P[i * B + j] * mu[j];
        quad += mu[i] * c1[i];
    }
    *c0 = -0.5 * (B * LOG2PI + logdet + quad);
    int t = 0;                                          // packed (b, b') for b <= b', in gk_sym_idx order
    for (int i = 0; i < B; ++i)
        for (int j = i; j < B; ++j) C2s[t++] = (i == j ? -0.5 : -1.0) * P[i * B + j];
    return true;
}

// host build of COt / CIt / L1 / selfc / selfa / logm into one contiguous blob
static int build_tables(gk_ctx *ctx, int K, int B, const double *log_m, const double *p, const double *mu,
                        const double *sigma, int self_mode, double beta, std::vector<double> &blob, size_t off[6]) {
    const int S = 1 + B + B * (B + 1) / 2, KS = K * S;
    off[0] = 0; off[1] = off[0] + (size_t)KS * K; off[2] = off[1] + (size_t)KS * K; off[3] = off[2] + (size_t)K * K;
    off[4] = off[3] + K; off[5] = off[4] + (size_t)S * K;
    blob.assign(off[5] + K, 0.0);
    double *COt = &blob[off[0]], *CIt = &blob[off[1]], *L1 = &blob[off[2]], *selfc = &blob[off[3]],
           *selfa = &blob[off[4]], *logm = &blob[off[5]];
    std::vector<double> cell((size_t)K * K * S, 0.0), l1m((size_t)K * K, 0.0);   // cell[(k,k')][s]
    for (int k = 0; k < K; ++k)
        for (int k2 = 0; k2 < K; ++k2) {
            double *cs = &cell[((size_t)k * K + k2) * S];
            if (p) {
                double pv = p[k * K + k2];
                if (!(pv >= 0.0 && pv <= 1.0)) return gk_fail(ctx, GK_EINVAL, "p[%d,%d]=%g outside [0,1]", k, k2, pv);
                // p == 0 / p == 1 are legal (stochastic_block_model.py:60): the reference's Bernoulli.log_probability nan_to_num's
                // x log 0 to -1.8e308 (a hard exclusion) and 0 log 0 to 0 (internals/distributions.py:59-67).  A finite -1e10 does
                // both here -- exp(-1e10) == 0 in fp64 and 0 * (-1e10) == 0 -- while the rank-one pushes of a move stay exact to 1e-6.
                double lg = pv > 0.0 ? log(pv) : GK_LOG_ZERO, l1 = pv < 1.0 ? log1p(-pv) : GK_LOG_ZERO;
                cs[0] += beta * (lg - l1);
                l1m[(size_t)k * K + k2] = beta * l1;
            }
            if (mu && B > 0) {
                double c0, c1[4], C2s[10];
                if (!gauss_coef(B, mu + ((size_t)k * K + k2) * B, sigma + ((size_t)k * K + k2) * B * B, &c0, c1, C2s))
                    return gk_fail(ctx, GK_EINVAL, "Sigma[%d,%d] is not positive definite", k, k2);
                cs[0] += beta * c0;
                for (int b = 0; b < B; ++b) cs[1 + b] += beta * c1[b];
                for (int j = 0; j < B * (B + 1) / 2; ++j) cs[1 + B + j] += beta * C2s[j];
            }
        }
    for (int k = 0; k < K; ++k)
        for (int k2 = 0; k2 < K; ++k2) {
            for (int s = 0; s < S; ++s) {
                COt[((size_t)k2 * S + s) * K + k] = cell[((size_t)k * K + k2) * S + s];
                CIt[((size_t)k2 * S + s) * K + k] = cell[((size_t)k2 * K + k) * S + s];
            }
            L1[(size_t)k * K + k2] = l1m[(size_t)k * K + k2] + l1m[(size_t)k2 * K + k];
        }
    const double mult = self_mode == GK_SELF_TWICE ? 2.0 : (self_mode == GK_SELF_ONCE ? 1.0 : 0.0);
    for (int k = 0; k < K; ++k) {
        logm[k] = log_m[k];
        if (mult == 0.0) continue;
        selfc[k] = mult * l1m[(size_t)k * K + k];
        for (int s = 0; s < S; ++s) selfa[(size_t)s * K + k] = cell[((size_t)k * K + k) * S + s];
        if (p) {   // the adjacency part of s=0 is doubled in TWICE mode; the Gaussian part is not
            const double pv = p[k * K + k];
            selfa[k] += (mult - 1.0) * beta * ((pv > 0.0 ? log(pv) : GK_LOG_ZERO) - (pv < 1.0 ? log1p(-pv) : GK_LOG_ZERO));
        }
    }
    return GK_OK;
}

template <typename T, int B>
static int launch_strip(gk_ctx *ctx, gk_net *net, const uint8_t *c8, int K, int n0, int nb_pad, int dir_first, int ndirs,
                        void **partial_out, int *nchunks_out) {
    typedef typename gk_acc<T, B>::type ACC;
    constexpr int S = gk_sdim<B>::S;
    int strips = nb_pad / GK_STRIP_COLS;
    // row chunks per strip: fill whole waves of (2 CTAs/SM) x num_sms slots -- a 1.08-wave grid costs two waves
    const int slots = 2 * ctx->num_sms, per = strips * ndirs;
    const int cmin = std::max(1, (net->nrows + GK_MAX_CHUNK_ROWS - 1) / GK_MAX_CHUNK_ROWS);
    const int cmax = std::max(cmin, std::min((net->nrows + 63) / 64, std::max(cmin + 8, slots / per)));
    int best = cmin; double best_cost = 1e30;
    for (int cnk = cmin; cnk <= cmax; ++cnk) {
        const double waves = (double)per * cnk / slots, full = ceil(waves - 1e-9);
        const double cost = full / waves * (1.0 + 0.01 * cnk);           // idle fraction, mild penalty per extra chunk
        if (cost < best_cost - 1e-12) { best_cost = cost; best = cnk; }
    }
    int chunk_rows = (net->nrows + best - 1) / best;
    chunk_rows = std::min(std::max(chunk_rows, 64), GK_MAX_CHUNK_ROWS);
    chunk_rows = (chunk_rows + 3) / 4 * 4;
    int nchunks = (net->nrows + chunk_rows - 1) / chunk_rows;
    size_t pbytes = (size_t)ndirs * nchunks * nb_pad * K * S * sizeof(ACC);
    void *partial = gk_scratch(ctx, 2, pbytes);
    if (!partial) return GK_ENOMEM;
    gk_strip_params P;
    P.Abase[0] = net->At; P.Abase[1] = net->A; P.Wbase[0] = net->Wt; P.Wbase[1] = net->W;
    P.ld = net->ld; P.row0 = net->row0; P.nrows = net->nrows; P.n0 = n0; P.nb_pad = nb_pad; P.dir_first = dir_first;
    P.c8 = c8; P.K = K; P.chunk_rows = chunk_rows; P.nchunks = nchunks; P.partial = partial;
    int U = std::max(K, GK_STRIP_WARPS);
    size_t smem = (size_t)U * S * GK_SLOT_PITCH * sizeof(ACC) + (size_t)chunk_rows * 3 + 16;
    if (smem > 160 * 1024) return gk_fail(ctx, GK_EUNSUPPORTED, "K=%d with B=%d in %s storage: the strip kernel needs %zu bytes of shared memory per CTA (limit 160 KB)", K, B, sizeof(T) == 8 ? "fp64" : "fp32", smem);
    if (gk_attr_needed(ctx, (const void *)k_strip_stats<T, B>))     // per context (= per device), not per process
        GK_CUDA(ctx, cudaFuncSetAttribute(k_strip_stats<T, B>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    double bytes = (double)net->nrows * (double)std::min(nb_pad, net->N - n0) * (1.0 + B * sizeof(T)) * ndirs;
    gk_time_begin(ctx, 0, bytes);
    k_strip_stats<T, B><<<dim3(strips, nchunks, ndirs), GK_STRIP_THREADS, smem, ctx->stream>>>(P);
    gk_time_end(ctx, 0);
    GK_CHECK_LAUNCH(ctx);
    *partial_out = partial; *nchunks_out = nchunks;
    return GK_OK;
}

template <typename T, int B>
static int strip_to_stats(gk_ctx *ctx, gk_net *net, const uint8_t *c8, int K, int n0, int nb_pad, int dir_first, int ndirs,
                          double *stats_dev) {
    typedef typename gk_acc<T, B>::type ACC;
    void *partial; int nchunks;
    int rc = launch_strip<T, B>(ctx, net, c8, K, n0, nb_pad, dir_first, ndirs, &partial, &nchunks);
    if (rc) return rc;
    size_t total = (size_t)nb_pad * ndirs * K * gk_sdim<B>::S;
    int grid = (int)std::min<size_t>((total + 255) / 256, (size_t)ctx->num_sms * 8);
    k_reduce_partials<ACC><<<grid, 256, 0, ctx->stream>>>((const ACC *)partial, stats_dev, ndirs, nchunks, nb_pad, K * gk_sdim<B>::S);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

#define GK_DISPATCH_B(net, T, CALL)                                                     \
    ((net)->B == 0 ? CALL(T, 0) : (net)->B == 1 ? CALL(T, 1) : (net)->B == 2 ? CALL(T, 2) : (net)->B == 3 ? CALL(T, 3) : CALL(T, 4))
#define GK_DISPATCH_TB(net, CALL) ((net)->dtype == GK_F32 ? GK_DISPATCH_B(net, float, CALL) : GK_DISPATCH_B(net, double, CALL))

// ---------------------------------------------------------------------------------------------
// K1
// ---------------------------------------------------------------------------------------------
extern "C" int gk_block_stats(gk_ctx *ctx, gk_net *net, const int32_t *c, int K, int64_t *n_edge, double *sum_w,
                              double *sum_ww) {
    if (!ctx || !net || !c || !n_edge) return GK_EINVAL;
    if (K < 1 || K > GK_MAX_K) return gk_fail(ctx, GK_EUNSUPPORTED, "K=%d: this build supports 1 <= K <= %d", K, GK_MAX_K);
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    const int N = net->N, B = net->B, S = 1 + B + B * (B + 1) / 2, KS = K * S;
    for (int i = 0; i < N; ++i) if (c[i] < 0 || c[i] >= K) return gk_fail(ctx, GK_EINVAL, "c[%d]=%d outside [0,%d)", i, c[i], K);
    // labels
    size_t lab_bytes = align_up((size_t)N * 4, 256) + align_up((size_t)N, 256);
    char *lab = (char *)gk_scratch(ctx, 5, lab_bytes);
    if (!lab) return GK_ENOMEM;
    int32_t *c_dev = (int32_t *)lab; uint8_t *c8 = (uint8_t *)(lab + align_up((size_t)N * 4, 256));
    GK_CUDA(ctx, cudaMemcpyAsync(c_dev, c, (size_t)N * 4, cudaMemcpyHostToDevice, ctx->stream));
    k_narrow_labels<<<(N + 255) / 256, 256, 0, ctx->stream>>>(c_dev, c8, N);
    GK_CHECK_LAUNCH(ctx);
    // cell accumulator
    const int nparts = 128;
    size_t cell_bytes = align_up((size_t)K * KS * 8, 256) + (size_t)nparts * K * KS * 8;
    char *cellbuf = (char *)gk_scratch(ctx, 7, cell_bytes);
    if (!cellbuf) return GK_ENOMEM;
    double *cell = (double *)cellbuf, *cellpart = (double *)(cellbuf + align_up((size_t)K * KS * 8, 256));
    GK_CUDA(ctx, cudaMemsetAsync(cell, 0, (size_t)K * KS * 8, ctx->stream));
    // node blocks of at most 16384 target nodes per pass (bounds the partial buffers)
    const int NBMAX = 16384;
    for (int n0 = 0; n0 < N; n0 += NBMAX) {
        int nb = std::min(NBMAX, N - n0), nb_pad = (nb + GK_STRIP_COLS - 1) / GK_STRIP_COLS * GK_STRIP_COLS;
        double *stats = (double *)gk_scratch(ctx, 3, (size_t)nb_pad * KS * 8);
        if (!stats) return GK_ENOMEM;
#define CALL_K1(T, BB) strip_to_stats<T, BB>(ctx, net, c8, K, n0, nb_pad, 1, 1, stats)
        int rc = GK_DISPATCH_TB(net, CALL_K1);
        if (rc) return rc;
        k_cell_partial<<<nparts, 128, (size_t)K * KS * 8, ctx->stream>>>(stats, c8, n0, nb, N, K, KS, cellpart);
        GK_CHECK_LAUNCH(ctx);
        k_cell_final<<<(K * KS + 127) / 128, 128, 0, ctx->stream>>>(cellpart, nparts, K, S, cell);
        GK_CHECK_LAUNCH(ctx);
    }
    std::vector<double> h((size_t)K * KS);
    GK_CUDA(ctx, cudaMemcpyAsync(h.data(), cell, (size_t)K * KS * 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int a = 0; a < K; ++a)
        for (int b = 0; b < K; ++b) {
            const double *cs = &h[((size_t)a * K + b) * S];
            n_edge[a * K + b] = (int64_t)llround(cs[0]);
            for (int i = 0; i < B; ++i) {
                if (sum_w) sum_w[((size_t)a * K + b) * B + i] = cs[1 + i];
                if (sum_ww)
                    for (int j = 0; j < B; ++j) {
                        int lo = std::min(i, j), hi = std::max(i, j);
                        sum_ww[(((size_t)a * K + b) * B + i) * B + j] = cs[gk_sym_idx(B, lo, hi)];
                    }
            }
        }
    return GK_OK;
}

// ---------------------------------------------------------------------------------------------
// K2: sweep = begin -> per node block { partial statistics ; sample } -> end
// ---------------------------------------------------------------------------------------------
__global__ void k_set_block(int32_t *c, uint8_t *c8, int *sizes, const int32_t *newc, int n0, int nb) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nb; i += gridDim.x * blockDim.x) {
        int n = n0 + i, a = c[n], b = newc[i];
        if (a != b) { c[n] = b; c8[n] = (uint8_t)b; atomicSub(&sizes[a], 1); atomicAdd(&sizes[b], 1); }
    }
}

extern "C" int gk_sbm_sweep_begin(gk_ctx *ctx, gk_net *net, const int32_t *c, int K, const double *log_m, const double *p,
                                  const double *mu, const double *sigma, int self_mode, int sampler, double beta,
                                  const double *uniforms) {
    if (!ctx || !net || !c || !log_m || !uniforms) return GK_EINVAL;
    if (K < 1 || K > GK_MAX_K) return gk_fail(ctx, GK_EUNSUPPORTED, "K=%d: this build supports 1 <= K <= %d", K, GK_MAX_K);
    if (self_mode < 0 || self_mode > 2 || sampler < 0 || sampler > 1) return gk_fail(ctx, GK_EINVAL, "bad self_mode/sampler");
    if ((mu == nullptr) != (sigma == nullptr)) return gk_fail(ctx, GK_EINVAL, "mu and sigma must be given together");
    if (mu && net->B == 0) return gk_fail(ctx, GK_EINVAL, "weight terms requested but the network has no W");
    if (self_mode != GK_SELF_SKIP && !net->have_diag) return gk_fail(ctx, GK_EINVAL, "gk_network_set_diag must be called before a sweep with self terms");
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    const int N = net->N;
    // when the weight term is off the kernels still stream W-less statistics: use B_eff = 0 tables of the net's S
    const int B = net->B, S = 1 + B + B * (B + 1) / 2;
    std::vector<int> sizes(K, 0);
    for (int i = 0; i < N; ++i) {
        if (c[i] < 0 || c[i] >= K) return gk_fail(ctx, GK_EINVAL, "c[%d]=%d outside [0,%d)", i, c[i], K);
        sizes[c[i]]++;
    }
    std::vector<double> blob; size_t off[6];
    int rc = build_tables(ctx, K, B, log_m, p, mu, sigma, self_mode, beta, blob, off);
    if (rc) return rc;
    delete ctx->sweep;
    gk_sweep *sw = ctx->sweep = new gk_sweep();
    sw->K = K; sw->S = S; sw->N = N; sw->sampler = sampler;
    sw->tables = (double *)gk_scratch(ctx, 1, blob.size() * 8);
    if (!sw->tables) return GK_ENOMEM;
    GK_CUDA(ctx, cudaMemcpyAsync(sw->tables, blob.data(), blob.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
    sw->T.COt = sw->tables + off[0]; sw->T.CIt = sw->tables + off[1]; sw->T.L1 = sw->tables + off[2];
    sw->T.selfc = sw->tables + off[3]; sw->T.selfa = sw->tables + off[4]; sw->T.logm = sw->tables + off[5];
    size_t o_c = 0, o_c8 = o_c + align_up((size_t)N * 4, 256), o_sz = o_c8 + align_up((size_t)N, 256),
           o_nc = o_sz + 256, o_u = o_nc + 256, tot = o_u + align_up((size_t)N * 8, 256);
    char *st = (char *)gk_scratch(ctx, 5, tot);
    if (!st) return GK_ENOMEM;
    sw->c = (int32_t *)(st + o_c); sw->c8 = (uint8_t *)(st + o_c8); sw->sizes = (int *)(st + o_sz);
    sw->nchanged = (unsigned long long *)(st + o_nc); sw->u = (double *)(st + o_u);
    GK_CUDA(ctx, cudaMemcpyAsync(sw->c, c, (size_t)N * 4, cudaMemcpyHostToDevice, ctx->stream));
    GK_CUDA(ctx, cudaMemcpyAsync(sw->sizes, sizes.data(), (size_t)K * 4, cudaMemcpyHostToDevice, ctx->stream));
    GK_CUDA(ctx, cudaMemcpyAsync(sw->u, uniforms, (size_t)N * 8, cudaMemcpyHostToDevice, ctx->stream));
    GK_CUDA(ctx, cudaMemsetAsync(sw->nchanged, 0, 64, ctx->stream));
    k_narrow_labels<<<(N + 255) / 256, 256, 0, ctx->stream>>>(sw->c, sw->c8, N);
    GK_CHECK_LAUNCH(ctx);
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));      // blob/sizes are host temporaries
    return GK_OK;
}

extern "C" size_t gk_sbm_partial_len(gk_ctx *ctx, int nb) {
    if (!ctx || !ctx->sweep || nb <= 0) return 0;
    size_t nb_pad = (size_t)(nb + GK_STRIP_COLS - 1) / GK_STRIP_COLS * GK_STRIP_COLS;
    return nb_pad * 2 * ctx->sweep->K * ctx->sweep->S;
}

extern "C" int gk_sbm_block_partials(gk_ctx *ctx, gk_net *net, int n0, int nb, double *stats_dev) {
    if (!ctx || !net || !ctx->sweep || !stats_dev) return GK_EINVAL;
    if (n0 < 0 || nb <= 0 || n0 + nb > net->N || n0 % GK_STRIP_COLS) return gk_fail(ctx, GK_EINVAL, "node block must start on a multiple of %d", GK_STRIP_COLS);
    gk_sweep *sw = ctx->sweep;
    int nb_pad = (nb + GK_STRIP_COLS - 1) / GK_STRIP_COLS * GK_STRIP_COLS;
#define CALL_K2(T, BB) strip_to_stats<T, BB>(ctx, net, sw->c8, sw->K, n0, nb_pad, 0, 2, stats_dev)
    return GK_DISPATCH_TB(net, CALL_K2);
}

template <typename T, int B>
static int launch_sample(gk_ctx *ctx, gk_net *net, gk_sweep *sw, int n0, int nb, const double *stats_dev) {
    {
        const int KP = gk_lp_kp(sw->K), npc = GK_LP_THREADS / KP;
        const size_t sm = gk_lp_smem(sw->K, sw->S);
        if (sm > 200 * 1024) return gk_fail(ctx, GK_EUNSUPPORTED, "K=%d with B=%d: the conditional tables need %zu bytes of shared memory (limit 200 KB)", sw->K, B, sm);
        if (sm > 48 * 1024 && gk_attr_needed(ctx, (const void *)k_lp_from_stats<B>))
            GK_CUDA(ctx, cudaFuncSetAttribute(k_lp_from_stats<B>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        k_lp_from_stats<B><<<(nb + npc - 1) / npc, GK_LP_THREADS, sm, ctx->stream>>>(stats_dev, sw->T, net->diagA, net->diagW, n0, nb,
                                                                                   sw->N, sw->K, KP, sw->lp);
        GK_CHECK_LAUNCH(ctx);
        if (ctx->fork_after_lp) {                  // the other stream's strip pass may start once the serial CTA can too
            ctx->fork_after_lp = false;
            GK_CUDA(ctx, cudaEventRecord(ctx->ev_fork, ctx->stream));
        }
    }
    gk_seq_params P;
    P.n0 = n0; P.nb = nb; P.K = sw->K; P.N = sw->N; P.ld = net->ld; P.row0 = net->row0; P.sampler = sw->sampler;
    P.c = sw->c; P.c8 = sw->c8; P.sizes = sw->sizes; P.lp = sw->lp; P.u = sw->u; P.T = sw->T;
    P.A = net->A; P.At = net->At; P.W = net->W; P.Wt = net->Wt;
    P.out_lp = sw->want_lp ? sw->out_lp : nullptr; P.nchanged = sw->nchanged;
    P.prof = ctx->sbm_prof;
    gk_time_begin(ctx, 1, 0.0);
    const size_t dyn = ((size_t)sw->K * (nb | 1) + nb) * 8 + (size_t)(nb + 15) / 16 * 16;
    // dynamic shared memory the kernel may have: 227 KB per CTA minus its static arrays (they grow with the S = 1 + B + B(B+1)/2
    // statistics per class), at most 208 KB
    static size_t dyn_max_memo[2][5] = {{0}};
    size_t &dyn_max = dyn_max_memo[sizeof(T) == 8][B];
    if (!dyn_max) {
        cudaFuncAttributes fa;
        GK_CUDA(ctx, cudaFuncGetAttributes(&fa, k_sbm_seq2<T, B>));
        dyn_max = std::min<size_t>(208 * 1024, (size_t)227 * 1024 - fa.sharedSizeBytes - 1024);
    }
    if (nb <= GK_SEQ_THREADS && dyn <= dyn_max) {          // conditionals of the block fit in shared memory
        if (gk_attr_needed(ctx, (const void *)k_sbm_seq2<T, B>))
            GK_CUDA(ctx, cudaFuncSetAttribute(k_sbm_seq2<T, B>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn_max));
        k_sbm_seq2<T, B><<<1, GK_SEQ_THREADS, dyn, ctx->stream>>>(P);
    } else {
        k_sbm_seq<T, B><<<1, GK_SEQ_THREADS, 0, ctx->stream>>>(P);
    }
    gk_time_end(ctx, 1);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

extern "C" int gk_sbm_block_sample(gk_ctx *ctx, gk_net *net, int n0, int nb, const double *stats_dev) {
    if (!ctx || !net || !ctx->sweep || !stats_dev) return GK_EINVAL;
    if (n0 < net->row0 || n0 + nb > net->row0 + net->nrows) return gk_fail(ctx, GK_EINVAL, "the sampling rank must own rows [%d,%d)", n0, n0 + nb);
    gk_sweep *sw = ctx->sweep;
    sw->lp = (double *)gk_scratch(ctx, 4, (size_t)nb * sw->K * 8);
    if (!sw->lp) return GK_ENOMEM;
#define CALL_SMP(T, BB) launch_sample<T, BB>(ctx, net, sw, n0, nb, stats_dev)
    return GK_DISPATCH_TB(net, CALL_SMP);
}

// labels of a node block: read (after the owner sampled) / write (on the other ranks)
extern "C" int gk_sbm_get_block(gk_ctx *ctx, int n0, int nb, int32_t *c_out) {
    if (!ctx || !ctx->sweep || !c_out || n0 < 0 || n0 + nb > ctx->sweep->N) return GK_EINVAL;
    GK_CUDA(ctx, cudaMemcpyAsync(c_out, ctx->sweep->c + n0, (size_t)nb * 4, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GK_OK;
}
extern "C" int gk_sbm_set_block(gk_ctx *ctx, int n0, int nb, const int32_t *c_in) {
    if (!ctx || !ctx->sweep || !c_in || n0 < 0 || n0 + nb > ctx->sweep->N) return GK_EINVAL;
    gk_sweep *sw = ctx->sweep;
    for (int i = 0; i < nb; ++i) if (c_in[i] < 0 || c_in[i] >= sw->K) return gk_fail(ctx, GK_EINVAL, "label out of range");
    int32_t *tmp = (int32_t *)gk_scratch(ctx, 6, (size_t)nb * 4);
    if (!tmp) return GK_ENOMEM;
    GK_CUDA(ctx, cudaMemcpyAsync(tmp, c_in, (size_t)nb * 4, cudaMemcpyHostToDevice, ctx->stream));
    k_set_block<<<(nb + 255) / 256, 256, 0, ctx->stream>>>(sw->c, sw->c8, sw->sizes, tmp, n0, nb);
    GK_CHECK_LAUNCH(ctx);
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GK_OK;
}

// device-side forms of get / set for the row-sharded sweep: the labels of a node block go owner -> NCCL broadcast -> the
// other ranks without touching the host (no stream synchronisation)
extern "C" int gk_sbm_export_block_dev(gk_ctx *ctx, int n0, int nb, int32_t *lab_dev) {
    if (!ctx || !ctx->sweep || !lab_dev || n0 < 0 || nb <= 0 || n0 + nb > ctx->sweep->N) return GK_EINVAL;
    GK_CUDA(ctx, cudaMemcpyAsync(lab_dev, ctx->sweep->c + n0, (size_t)nb * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    return GK_OK;
}
extern "C" int gk_sbm_import_block_dev(gk_ctx *ctx, int n0, int nb, const int32_t *lab_dev) {
    if (!ctx || !ctx->sweep || !lab_dev || n0 < 0 || nb <= 0 || n0 + nb > ctx->sweep->N) return GK_EINVAL;
    gk_sweep *sw = ctx->sweep;
    k_set_block<<<(nb + 255) / 256, 256, 0, ctx->stream>>>(sw->c, sw->c8, sw->sizes, lab_dev, n0, nb);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

extern "C" int gk_sbm_sweep_end(gk_ctx *ctx, gk_net *net, int32_t *c_out, double *out_lp, int64_t *out_nchanged) {
    if (!ctx || !ctx->sweep) return GK_EINVAL;
    gk_sweep *sw = ctx->sweep;
    unsigned long long nch = 0;
    if (ctx->sbm_prof) {                               // phase clocks of k_sbm_seq2 (shares)
        unsigned long long pc[7];
        GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        GK_CUDA(ctx, cudaMemcpy(pc, sw->nchanged, sizeof pc, cudaMemcpyDeviceToHost));
        fprintf(stderr, "[gk_sbm prof] moves %llu  draw %llu  bar1 %llu  tables+features %llu  bar2 %llu  push %llu  bar3 %llu\n",
                pc[0], pc[1], pc[2], pc[3], pc[4], pc[5], pc[6]);
    }
    if (c_out) GK_CUDA(ctx, cudaMemcpyAsync(c_out, sw->c, (size_t)sw->N * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_lp && sw->want_lp) GK_CUDA(ctx, cudaMemcpyAsync(out_lp, sw->out_lp, (size_t)sw->N * sw->K * 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaMemcpyAsync(&nch, sw->nchanged, 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (out_nchanged) *out_nchanged = (int64_t)nch;
    delete sw; ctx->sweep = nullptr;
    return GK_OK;
}

// The row-sharded sweep with the engine's own communicator (gk_comm_init): per node block, partial statistics over this
// rank's rows -> NCCL all-reduce -> the rank that owns the block's rows draws it -> NCCL broadcast of the block's labels.
// Everything is enqueued on the context stream; the host does not wait inside the sweep.
static int sbm_sweep_sharded(gk_ctx *ctx, gk_net *net, int32_t *c_inout, int K, const double *log_m, const double *p,
                             const double *mu, const double *sigma, int self_mode, int sampler, double beta,
                             const double *uniforms, int block_nodes, double *out_lp, int64_t *out_nchanged) {
    if (!ctx->comm)
        return gk_fail(ctx, GK_EINVAL, "gk_sbm_sweep on a row-sharded network needs gk_comm_init (or use the begin/partials/sample/end form)");
    if (out_lp) return gk_fail(ctx, GK_EUNSUPPORTED, "per-node log conditionals are available on single-rank networks only");
    const int world = gk_comm_world(ctx), rank = gk_comm_rank(ctx), N = net->N;
    int r0, r1;
    gk_comm_row_range(N, rank, world, &r0, &r1);
    if (r0 != net->row0 || r1 != net->row0 + net->nrows)
        return gk_fail(ctx, GK_EINVAL, "rank %d of %d must hold rows [%d, %d), the network holds [%d, %d)", rank, world, r0, r1, net->row0, net->row0 + net->nrows);
    int rc = gk_sbm_sweep_begin(ctx, net, c_inout, K, log_m, p, mu, sigma, self_mode, sampler, beta, uniforms);
    if (rc) return rc;
    gk_sweep *sw = ctx->sweep;
    int bn = block_nodes > 0 ? block_nodes : 1024;
    bn = std::max(128, bn / 128 * 128);
    double *stats = (double *)gk_scratch(ctx, 10, gk_sbm_partial_len(ctx, bn) * 8);
    int32_t *lab = (int32_t *)gk_scratch(ctx, 11, (size_t)bn * 4);
    if (!stats || !lab) return GK_ENOMEM;
    ctx->last_cells.valid = false;
    for (int rk = 0; rk < world; ++rk) {
        int s0, s1;
        gk_comm_row_range(N, rk, world, &s0, &s1);
        for (int n0 = s0; n0 < s1; n0 += bn) {
            const int nb = std::min(bn, s1 - n0);
            if ((rc = gk_sbm_block_partials(ctx, net, n0, nb, stats))) return rc;
            if ((rc = gk_comm_allreduce_f64(ctx, stats, gk_sbm_partial_len(ctx, nb)))) return rc;
            if (rank == rk) {
                if ((rc = gk_sbm_block_sample(ctx, net, n0, nb, stats))) return rc;
                GK_CUDA(ctx, cudaMemcpyAsync(lab, sw->c + n0, (size_t)nb * 4, cudaMemcpyDeviceToDevice, ctx->stream));
            }
            if ((rc = gk_comm_broadcast(ctx, lab, (size_t)nb * 4, rk))) return rc;
            if (rank != rk) {
                k_set_block<<<(nb + 255) / 256, 256, 0, ctx->stream>>>(sw->c, sw->c8, sw->sizes, lab, n0, nb);
                GK_CHECK_LAUNCH(ctx);
            }
        }
    }
    if ((rc = gk_comm_allreduce_i64(ctx, (long long *)sw->nchanged, 1))) return rc;     // every block was counted by its owner
    return gk_sbm_sweep_end(ctx, net, c_inout, nullptr, out_nchanged);
}

extern "C" int gk_sbm_sweep(gk_ctx *ctx, gk_net *net, int32_t *c_inout, int K, const double *log_m, const double *p,
                            const double *mu, const double *sigma, int self_mode, int sampler, double beta,
                            const double *uniforms, int block_nodes, double *out_lp, int64_t *out_nchanged) {
    if (!ctx || !net || !c_inout) return GK_EINVAL;
    if (K == 1) { if (out_nchanged) *out_nchanged = 0; return GK_OK; }      // C == 1: no-op (adjacency.py:657-658)
    if (net->row0 != 0 || net->nrows != net->N)
        return sbm_sweep_sharded(ctx, net, c_inout, K, log_m, p, mu, sigma, self_mode, sampler, beta, uniforms, block_nodes, out_lp, out_nchanged);
    struct SpecSelf { gk_ctx *c; SpecSelf(gk_ctx *c_) : c(c_) { c->spec.self = true; } ~SpecSelf() { c->spec.self = false; } } spec_self(ctx);
    int rc = gk_sbm_sweep_begin(ctx, net, c_inout, K, log_m, p, mu, sigma, self_mode, sampler, beta, uniforms);
    if (rc) return rc;
    gk_sweep *sw = ctx->sweep;
    const int N = net->N;
    if (out_lp) {
        sw->out_lp = (double *)gk_scratch(ctx, 6, (size_t)N * K * 8);
        if (!sw->out_lp) return GK_ENOMEM;
        sw->want_lp = true;
    }
    int bn = block_nodes > 0 ? block_nodes : 1024;
    bn = std::max(GK_STRIP_COLS, bn / GK_STRIP_COLS * GK_STRIP_COLS);
    // background strip passes left by the previous sweep (mode 3, below): matched -- or waited for and dropped -- BEFORE any
    // scratch buffer they use can be resized
    const int nblk = (N + bn - 1) / bn;
    bool spec = ctx->spec.active && ctx->spec.net == (const void *)net && ctx->spec.K == K && ctx->spec.bn == bn && ctx->spec.N == N &&
                !ctx->sbm_prof && memcmp(ctx->spec.labels.data(), c_inout, (size_t)N * 4) == 0;
    if (ctx->spec.active && !spec) gk_spec_cancel(ctx);
    ctx->spec.active = false;                     // consumed (or cancelled); re-armed at the end of this sweep
    // per-node statistics of the whole sweep stay resident ([N_pad][2][K*S] doubles): their in-edge half, folded with the labels
    // the sweep ends on, IS the K1 table of the next resample() -- no third pass over A / W (gk_sbm_sweep_stats)
    const int KS = K * sw->S;
    const size_t n_pad = (size_t)(N + bn - 1) / bn * bn;
    double *stats_all = (double *)gk_scratch(ctx, 3, n_pad * 2 * KS * 8);
    if (!stats_all) return GK_ENOMEM;
    const int CAP = GK_CELL_MOVER_CAP, nparts = 128;
    size_t o_cell = 0, o_part = align_up((size_t)K * KS * 8, 256), o_cnt = o_part + align_up((size_t)nparts * K * KS * 8, 256),
           o_list = o_cnt + 256, o_cold = o_list + align_up((size_t)CAP * 12, 256), tot = o_cold + align_up((size_t)N, 256);
    char *cb = (char *)gk_scratch(ctx, 8, tot);
    if (!cb) return GK_ENOMEM;
    double *cell = (double *)(cb + o_cell), *cellpart = (double *)(cb + o_part);
    int *mcnt = (int *)(cb + o_cnt), *mlist = (int *)(cb + o_list);
    uint8_t *c_old8 = (uint8_t *)(cb + o_cold);
    GK_CUDA(ctx, cudaMemcpyAsync(c_old8, sw->c8, (size_t)N, cudaMemcpyDeviceToDevice, ctx->stream));
    ctx->last_cells.valid = false;
    // The strip pass of block b+1 runs on a second stream beside the serial kernel of block b, on a snapshot of the labels
    // taken before block b is drawn; k_stats_fix then moves the terms of block b's movers to their new class (nothing to do
    // in the stationary regime).  The serial CTA is enqueued first and the strip grid leaves four SMs free, so the two run
    // side by side.
    // ... which pays while few nodes move (the correction costs ~2 us per mover and block): taken when the previous sweep on
    // this network moved at most 256 nodes, i.e. not during burn-in or on weakly structured data
    const bool few = ctx->last_sweep_net == (const void *)net && ctx->last_sweep_moves >= 0 && ctx->last_sweep_moves <= 256;
    // Whole-sweep speculation (mode 3): the previous sweep left the strip passes of ALL blocks of this one running on the second
    // stream, from the labels it ended on.  They are this sweep's statistics if the caller comes back with those labels.
    const bool overlap = !spec && !ctx->sbm_prof && bn <= 2048 && N > bn &&
                         (ctx->sbm_overlap_mode == 2 || (ctx->sbm_overlap_mode == 1 && few) || (ctx->sbm_overlap_mode == 3 && few));
    int *spec_cnt = nullptr, *spec_list = nullptr;
    uint8_t *spec_c = nullptr;
    if (spec || ctx->sbm_overlap_mode == 3) {
        spec_c = (uint8_t *)gk_scratch(ctx, 10, (size_t)N);
        spec_cnt = (int *)gk_scratch(ctx, 11, ((size_t)3 * N + 64) * 4);
        if (!spec_c || !spec_cnt) return GK_ENOMEM;
        spec_list = spec_cnt + 64;
    }
    if (spec) GK_CUDA(ctx, cudaMemsetAsync(spec_cnt, 0, 4, ctx->stream));
    uint8_t *c_spec = nullptr;
    if (overlap) {
        if (!ctx->aux) {
            int lo = 0, hi = 0;                    // lowest priority: when both become runnable the serial CTA is placed first
            GK_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&lo, &hi));
            GK_CUDA(ctx, cudaStreamCreateWithPriority(&ctx->aux, cudaStreamNonBlocking, lo));
            GK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
            GK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
        }
        c_spec = (uint8_t *)gk_scratch(ctx, 10, (size_t)N);
        if (!c_spec) return GK_ENOMEM;
    }
    for (int n0 = 0; n0 < N; n0 += bn) {
        int nb = std::min(bn, N - n0);
        double *stats = stats_all + (size_t)n0 * 2 * KS;
        if (spec) {
            GK_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->spec.ev[n0 / bn], 0));
            if (n0 > 0) {                          // movers of the earlier blocks of this sweep: their terms change class in this block's rows
#define CALL_FIXL(T, BB) (k_stats_fix_list<T, BB><<<(nb + 63) / 64, 128, 0, ctx->stream>>>(net->A, net->At, net->W, net->Wt, net->ld, spec_cnt, spec_list, n0, nb, K, stats), GK_OK)
                rc = GK_DISPATCH_TB(net, CALL_FIXL);
                GK_CHECK_LAUNCH(ctx);
            }
            if ((rc = gk_sbm_block_sample(ctx, net, n0, nb, stats))) return rc;
            if (n0 + bn < N) {
                k_append_movers<<<1, 1024, 0, ctx->stream>>>(spec_c, sw->c8, n0, nb, spec_cnt, spec_list);
                GK_CHECK_LAUNCH(ctx);
            }
            continue;
        }
        if (!overlap) {
            if ((rc = gk_sbm_block_partials(ctx, net, n0, nb, stats))) return rc;
            if ((rc = gk_sbm_block_sample(ctx, net, n0, nb, stats))) return rc;
            continue;
        }
        if (n0 == 0) {
            if ((rc = gk_sbm_block_partials(ctx, net, n0, nb, stats))) return rc;
        } else {                                   // statistics came from the second stream, on the labels of before block b-1
            GK_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_join, 0));
            const int p0 = n0 - bn;
#define CALL_FIX(T, BB) (k_stats_fix<T, BB><<<(nb + 63) / 64, 128, 0, ctx->stream>>>(net->A, net->At, net->W, net->Wt, net->ld, c_spec, sw->c8, p0, bn, n0, nb, K, stats), GK_OK)
            rc = GK_DISPATCH_TB(net, CALL_FIX);
            GK_CHECK_LAUNCH(ctx);
        }
        const int n1 = n0 + bn;
        if (n1 < N) {
            GK_CUDA(ctx, cudaMemcpyAsync(c_spec, sw->c8, (size_t)N, cudaMemcpyDeviceToDevice, ctx->stream));
            ctx->fork_after_lp = true;
        }
        if ((rc = gk_sbm_block_sample(ctx, net, n0, nb, stats))) return rc;          // lp + the serial CTA, enqueued first
        if (n1 < N) {
            GK_CUDA(ctx, cudaStreamWaitEvent(ctx->aux, ctx->ev_fork, 0));
            cudaStream_t keep = ctx->stream;
            ctx->stream = ctx->aux;
            const int nb1 = std::min(bn, N - n1), nb1_pad = (nb1 + GK_STRIP_COLS - 1) / GK_STRIP_COLS * GK_STRIP_COLS;
            double *stats1 = stats_all + (size_t)n1 * 2 * KS;
#define CALL_K2S(T, BB) strip_to_stats<T, BB>(ctx, net, c_spec, K, n1, nb1_pad, 0, 2, stats1)
            rc = GK_DISPATCH_TB(net, CALL_K2S);
            ctx->stream = keep;
            if (rc) return rc;
            GK_CUDA(ctx, cudaEventRecord(ctx->ev_join, ctx->aux));
        }
    }
    // K1 of the final labels: fold, then move the movers' out-edges from their old source class to the new one
    GK_CUDA(ctx, cudaMemsetAsync(cell, 0, (size_t)K * KS * 8, ctx->stream));
    GK_CUDA(ctx, cudaMemsetAsync(mcnt, 0, 4, ctx->stream));
    k_cell_partial2<<<nparts, 128, (size_t)K * KS * 8, ctx->stream>>>(stats_all, sw->c8, N, K, KS, cellpart);
    GK_CHECK_LAUNCH(ctx);
    k_cell_final<<<(K * KS + 127) / 128, 128, 0, ctx->stream>>>(cellpart, nparts, K, sw->S, cell);
    GK_CHECK_LAUNCH(ctx);
    k_find_movers<<<std::min((N + 255) / 256, ctx->num_sms * 4), 256, 0, ctx->stream>>>(c_old8, sw->c8, N, CAP, mcnt, mlist);
    GK_CHECK_LAUNCH(ctx);
#define CALL_MOV(T, BB) (k_cell_movers<T, BB><<<std::min(CAP, ctx->num_sms * 4), 64, (size_t)KS * 64 * 8, ctx->stream>>>(net->A, net->W, net->ld, sw->c8, mcnt, mlist, CAP, bn, N, K, cell), GK_OK)
#define CALL_MOVATTR(T, BB) (cudaFuncSetAttribute(k_cell_movers<T, BB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) == cudaSuccess ? GK_OK : GK_ECUDA)
    if ((size_t)KS * 64 * 8 > 200 * 1024) return gk_fail(ctx, GK_EUNSUPPORTED, "K=%d with B=%d: the mover correction needs %zu bytes of shared memory", K, net->B, (size_t)KS * 64 * 8);
    if ((size_t)KS * 64 * 8 > 48 * 1024 && (rc = GK_DISPATCH_TB(net, CALL_MOVATTR))) return gk_fail(ctx, rc, "cudaFuncSetAttribute(k_cell_movers)");
    rc = GK_DISPATCH_TB(net, CALL_MOV);
    GK_CHECK_LAUNCH(ctx);
    {
        std::vector<double> &hc = ctx->last_cells.cell;
        hc.resize((size_t)K * KS);
        int nmov = 0;
        GK_CUDA(ctx, cudaMemcpyAsync(hc.data(), cell, (size_t)K * KS * 8, cudaMemcpyDeviceToHost, ctx->stream));
        GK_CUDA(ctx, cudaMemcpyAsync(&nmov, mcnt, 4, cudaMemcpyDeviceToHost, ctx->stream));
        const int S_ = sw->S;
        int64_t moves = 0;
        const uint8_t *c8_dev = sw->c8;
        rc = gk_sbm_sweep_end(ctx, net, c_inout, out_lp, &moves);      // synchronises the stream
        if (rc) return rc;
        if (out_nchanged) *out_nchanged = moves;
        ctx->last_sweep_moves = moves; ctx->last_sweep_net = net;
        if (ctx->sbm_overlap_mode == 3 && moves <= 256 && !ctx->sbm_prof && nblk <= 64 && N > bn) {
            // strip passes of every block of the NEXT sweep, from the labels this one ended on, in the background: they overlap the
            // caller's O(K^2) host draws and the serial kernels of the next sweep; every byte of A / W is still streamed every sweep
            if (!ctx->aux) {
                int lo = 0, hi = 0;
                GK_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&lo, &hi));
                GK_CUDA(ctx, cudaStreamCreateWithPriority(&ctx->aux, cudaStreamNonBlocking, lo));
                GK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
                GK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
            }
            while ((int)ctx->spec.ev.size() < nblk) {
                cudaEvent_t e;
                GK_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                ctx->spec.ev.push_back(e);
            }
            GK_CUDA(ctx, cudaMemcpyAsync(spec_c, c8_dev, (size_t)N, cudaMemcpyDeviceToDevice, ctx->aux));
            cudaStream_t keep = ctx->stream;
            ctx->stream = ctx->aux;
            for (int n1 = 0; n1 < N && !rc; n1 += bn) {
                const int nb1 = std::min(bn, N - n1), nb1_pad = (nb1 + GK_STRIP_COLS - 1) / GK_STRIP_COLS * GK_STRIP_COLS;
                double *stats1 = stats_all + (size_t)n1 * 2 * KS;
#define CALL_K2P(T, BB) strip_to_stats<T, BB>(ctx, net, spec_c, K, n1, nb1_pad, 0, 2, stats1)
                rc = GK_DISPATCH_TB(net, CALL_K2P);
                if (!rc && cudaEventRecord(ctx->spec.ev[n1 / bn], ctx->aux) != cudaSuccess) rc = GK_ECUDA;
            }
            ctx->stream = keep;
            if (rc) { cudaStreamSynchronize(ctx->aux); return gk_fail(ctx, rc, "speculative strip passes failed to launch"); }
            ctx->spec.labels.assign(c_inout, c_inout + N);
            ctx->spec.net = net; ctx->spec.K = K; ctx->spec.bn = bn; ctx->spec.N = N; ctx->spec.nblk = nblk;
            ctx->spec.active = true;
        }
        ctx->last_cells.net = net; ctx->last_cells.K = K; ctx->last_cells.S = S_;
        ctx->last_cells.valid = nmov <= CAP;
        return GK_OK;
    }
}

// K1 statistics (off-diagonal pairs, as gk_block_stats returns them) of the labels the last unsharded gk_sbm_sweep on `net`
// ended on -- folded from the sweep's own per-node statistics, so resample() needs no separate pass over A / W.
// *valid = 0 when there is nothing to return (no sweep yet, another network or K, or the sweep moved too many nodes).
extern "C" int gk_sbm_sweep_stats(gk_ctx *ctx, gk_net *net, int K, int64_t *n_edge, double *sum_w, double *sum_ww, int *valid) {
    if (!ctx || !net || !n_edge || !valid) return GK_EINVAL;
    *valid = 0;
    if (!ctx->last_cells.valid || ctx->last_cells.net != net || ctx->last_cells.K != K) return GK_OK;
    const int B = net->B, S = ctx->last_cells.S;
    const std::vector<double> &h = ctx->last_cells.cell;
    for (int a = 0; a < K; ++a)
        for (int b = 0; b < K; ++b) {
            const double *cs = &h[((size_t)a * K + b) * S];
            n_edge[a * K + b] = (int64_t)llround(cs[0]);
            for (int i = 0; i < B; ++i) {
                if (sum_w) sum_w[((size_t)a * K + b) * B + i] = cs[1 + i];
                if (sum_ww)
                    for (int j = 0; j < B; ++j) {
                        int lo = std::min(i, j), hi = std::max(i, j);
                        sum_ww[(((size_t)a * K + b) * B + i) * B + j] = cs[gk_sym_idx(B, lo, hi)];
                    }
            }
        }
    *valid = 1;
    return GK_OK;
}
