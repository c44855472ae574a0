// gk_mf.cu -- host side of the mean-field SBM path (K10): launch plans and C-ABI entry points.
#include <math.h>
#include <string.h>
#include <algorithm>
#include "gk_mf_kernels.cuh"

static inline size_t mf_align(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct gk_mf_state {
    int K = 0, Kp = 0, KGp = 1, N = 0;
    double *m64 = nullptr; void *mT = nullptr; double *Msum = nullptr;
};

// device copies of mf_m: fp64 master [N][K], compute-type copy [N][Kp], column sums [K]   (ctx scratch slot 5)
static int mf_upload(gk_ctx *ctx, gk_net *net, const double *mf_m, int K, gk_mf_state *st) {
    const int N = net->N;
    st->K = K; st->N = N; st->Kp = (K + 3) / 4 * 4;
    int kg = st->Kp / 4; st->KGp = 1; while (st->KGp < kg) st->KGp *= 2;
    size_t o_m = 0, o_t = o_m + mf_align((size_t)N * K * 8, 256), o_s = o_t + mf_align((size_t)N * st->Kp * 8, 256),
           tot = o_s + 256;
    char *buf = (char *)gk_scratch(ctx, 5, tot);
    if (!buf) return GK_ENOMEM;
    st->m64 = (double *)(buf + o_m); st->mT = buf + o_t; st->Msum = (double *)(buf + o_s);
    std::vector<double> ms(K, 0.0);
    for (int n = 0; n < N; ++n)
        for (int k = 0; k < K; ++k) {
            double v = mf_m[(size_t)n * K + k];
            if (!(v >= 0.0 && v <= 1.0 + 1e-9)) return gk_fail(ctx, GK_EINVAL, "mf_m[%d,%d]=%g outside [0,1]", n, k, v);
            ms[k] += v;
        }
    GK_CUDA(ctx, cudaMemcpyAsync(st->m64, mf_m, (size_t)N * K * 8, cudaMemcpyHostToDevice, ctx->stream));
    GK_CUDA(ctx, cudaMemcpyAsync(st->Msum, ms.data(), (size_t)K * 8, cudaMemcpyHostToDevice, ctx->stream));
    int grid = (int)std::min<size_t>(((size_t)N * st->Kp + 255) / 256, (size_t)ctx->num_sms * 8);
    if (net->dtype == GK_F32) k_mf_pack<float><<<grid, 256, 0, ctx->stream>>>(st->m64, (float *)st->mT, N, K, st->Kp);
    else k_mf_pack<double><<<grid, 256, 0, ctx->stream>>>(st->m64, (double *)st->mT, N, K, st->Kp);
    GK_CHECK_LAUNCH(ctx);
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));          // ms is a host temporary
    return GK_OK;
}

template <typename T, int B>
static int soft_strip_to_stats(gk_ctx *ctx, gk_net *net, gk_mf_state *st, int n0, int nb_pad, int dir_first, int ndirs,
                               double *stats_dev) {
    typedef typename gk_acc<T, B>::type ACC;
    constexpr int S = gk_sdim<B>::S;
    const int K = st->K, strips = nb_pad / GK_STRIP_COLS;
    // one CTA per SM (register-heavy kernel): row chunks chosen so the grid fills whole waves
    const int slots = ctx->num_sms, per = strips * ndirs;
    const int cmin = 1, cmax = std::max(1, std::min((net->nrows + 255) / 256, std::max(8, 4 * slots / per)));
    int best = cmin; double best_cost = 1e30;
    for (int cnk = cmin; cnk <= cmax; ++cnk) {
        const double waves = (double)per * cnk / slots, full = ceil(waves - 1e-9);
        const double cost = full / waves * (1.0 + 0.01 * cnk);
        if (cost < best_cost - 1e-12) { best_cost = cost; best = cnk; }
    }
    int chunk_rows = (net->nrows + best - 1) / best;
    chunk_rows = std::max((chunk_rows + 63) / 64 * 64, 64);
    const int nchunks = (net->nrows + chunk_rows - 1) / chunk_rows;
    size_t pbytes = (size_t)ndirs * nchunks * nb_pad * K * S * sizeof(ACC);
    void *partial = gk_scratch(ctx, 2, pbytes);
    if (!partial) return GK_ENOMEM;
    gk_soft_params P;
    P.Abase[0] = net->At; P.Abase[1] = net->A; P.Wbase[0] = net->Wt; P.Wbase[1] = net->W;
    P.ld = net->ld; P.row0 = net->row0; P.nrows = net->nrows; P.n0 = n0; P.nb_pad = nb_pad; P.dir_first = dir_first;
    P.mT = st->mT; P.K = K; P.Kp = st->Kp; P.KGp = st->KGp; P.chunk_rows = chunk_rows; P.nchunks = nchunks; P.partial = partial;
    size_t smem = (size_t)st->Kp * S * GK_SLOT_PITCH * sizeof(ACC);
    if (gk_attr_needed(ctx, (const void *)k_strip_soft<T, B>)) {
        GK_CUDA(ctx, cudaFuncSetAttribute(k_strip_soft<T, B>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    }
    double bytes = (double)net->nrows * (double)std::min(nb_pad, net->N - n0) * (1.0 + B * sizeof(T)) * ndirs;
    gk_time_begin(ctx, 6, bytes);
    k_strip_soft<T, B><<<dim3(strips, nchunks, ndirs), GK_SOFT_THREADS, smem, ctx->stream>>>(P);
    gk_time_end(ctx, 6);
    GK_CHECK_LAUNCH(ctx);
    size_t total = (size_t)nb_pad * ndirs * K * S;
    int grid = (int)std::min<size_t>((total + 255) / 256, (size_t)ctx->num_sms * 8);
    k_reduce_partials<ACC><<<grid, 256, 0, ctx->stream>>>((const ACC *)partial, stats_dev, ndirs, nchunks, nb_pad, K * S);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

#define GK_MF_DISPATCH_TB(net, CALL)                                                   \
    ((net)->dtype == GK_F32                                                            \
         ? ((net)->B == 0 ? CALL(float, 0) : (net)->B == 1 ? CALL(float, 1) : CALL(float, 2))   \
         : ((net)->B == 0 ? CALL(double, 0) : (net)->B == 1 ? CALL(double, 1) : CALL(double, 2)))

// ---------------------------------------------------------------------------------------------
// soft block-pair statistics (mf_update_p, NIW natural-parameter statistics)
// ---------------------------------------------------------------------------------------------
extern "C" int gk_mf_sbm_soft_stats(gk_ctx *ctx, gk_net *net, const double *mf_m, int K, double *n_soft, double *sum_w,
                                    double *sum_ww) {
    if (!ctx || !net || !mf_m || !n_soft) return GK_EINVAL;
    if (K < 1 || K > GK_MAX_K) return gk_fail(ctx, GK_EUNSUPPORTED, "K=%d: this build supports 1 <= K <= %d", K, GK_MAX_K);
    if (net->B > 2) return gk_fail(ctx, GK_EUNSUPPORTED, "B=%d: this build supports B <= 2", net->B);
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    const int N = net->N, B = net->B, S = 1 + B + B * (B + 1) / 2, KS = K * S;
    gk_mf_state st;
    int rc = mf_upload(ctx, net, mf_m, K, &st);
    if (rc) return rc;
    const int nparts = 128;
    size_t cell_bytes = mf_align((size_t)K * KS * 8, 256) + (size_t)nparts * K * KS * 8;
    char *cellbuf = (char *)gk_scratch(ctx, 7, cell_bytes);
    if (!cellbuf) return GK_ENOMEM;
    double *cell = (double *)cellbuf, *cellpart = (double *)(cellbuf + mf_align((size_t)K * KS * 8, 256));
    GK_CUDA(ctx, cudaMemsetAsync(cell, 0, (size_t)K * KS * 8, ctx->stream));
    const int NBMAX = 8192;
    for (int n0 = 0; n0 < N; n0 += NBMAX) {
        int nb = std::min(NBMAX, N - n0), nb_pad = (nb + GK_STRIP_COLS - 1) / GK_STRIP_COLS * GK_STRIP_COLS;
        double *stats = (double *)gk_scratch(ctx, 3, (size_t)nb_pad * KS * 8);
        if (!stats) return GK_ENOMEM;
#define CALL_SOFT1(T, BB) soft_strip_to_stats<T, BB>(ctx, net, &st, n0, nb_pad, 1, 1, stats)
        rc = GK_MF_DISPATCH_TB(net, CALL_SOFT1);
        if (rc) return rc;
        k_cell_partial_soft<<<nparts, 128, (size_t)K * KS * 8, ctx->stream>>>(stats, st.m64, n0, nb, N, K, KS, cellpart);
        GK_CHECK_LAUNCH(ctx);
        k_cell_final_soft<<<(K * KS + 127) / 128, 128, 0, ctx->stream>>>(cellpart, nparts, K, S, cell);
        GK_CHECK_LAUNCH(ctx);
    }
    std::vector<double> h((size_t)K * KS);
    GK_CUDA(ctx, cudaMemcpyAsync(h.data(), cell, (size_t)K * KS * 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int a = 0; a < K; ++a)
        for (int b = 0; b < K; ++b) {
            const double *cs = &h[((size_t)a * K + b) * S];
            n_soft[a * K + b] = cs[0];
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
// sequential mean-field update of the class probabilities (mf_update_c)
// ---------------------------------------------------------------------------------------------
template <typename T, int B>
static int launch_mf_seq(gk_ctx *ctx, gk_net *net, gk_mf_state *st, const double *tables_dev, const size_t *off, int n0, int nb,
                         const double *stats, double stepsize, double *out_lp_dev) {
    constexpr int S = gk_sdim<B>::S;
    const int K = st->K, KS = K * S;
    gk_mfseq_params P;
    P.n0 = n0; P.nb = nb; P.K = K; P.Kp = st->Kp; P.N = st->N; P.ld = net->ld; P.row0 = net->row0; P.is_f32 = net->dtype == GK_F32;
    P.m64 = st->m64; P.mT = st->mT; P.Msum = st->Msum; P.stats = stats;
    P.COt = tables_dev + off[0]; P.CIt = tables_dev + off[1]; P.BOt = tables_dev + off[2]; P.BIt = tables_dev + off[3];
    P.logm = tables_dev + off[4];
    P.diagA = net->diagA; P.diagW = net->diagW; P.A = net->A; P.At = net->At; P.W = net->W; P.Wt = net->Wt;
    P.stepsize = stepsize; P.out_lp = out_lp_dev;
    size_t smem = ((size_t)GK_MF_MAX_NB * (K + 4) + 2 * (size_t)KS * K + 2 * (size_t)K * K + 2 * K + 2 * (size_t)S * K) * 8;
    if (smem > 220 * 1024) return gk_fail(ctx, GK_EUNSUPPORTED, "K=%d, B=%d needs %zu bytes of shared memory", K, B, smem);
    if (gk_attr_needed(ctx, (const void *)k_mf_seq<T, B>)) {
        GK_CUDA(ctx, cudaFuncSetAttribute(k_mf_seq<T, B>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    }
    gk_time_begin(ctx, 7, 0.0);
    k_mf_seq<T, B><<<1, GK_MFSEQ_THREADS, smem, ctx->stream>>>(P);
    gk_time_end(ctx, 7);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

extern "C" int gk_mf_sbm_update_c(gk_ctx *ctx, gk_net *net, double *mf_m_inout, int K, const double *E_log_m,
                                  const double *base_out, const double *base_in, const double *coef_out,
                                  const double *coef_in, double stepsize, int block_nodes, double *out_lp) {
    if (!ctx || !net || !mf_m_inout || !E_log_m || !base_out || !base_in || !coef_out || !coef_in) return GK_EINVAL;
    if (K < 1 || K > GK_MAX_K) return gk_fail(ctx, GK_EUNSUPPORTED, "K=%d: this build supports 1 <= K <= %d", K, GK_MAX_K);
    if (net->B > 2) return gk_fail(ctx, GK_EUNSUPPORTED, "B=%d: this build supports B <= 2", net->B);
    if (net->row0 != 0 || net->nrows != net->N) return gk_fail(ctx, GK_EINVAL, "gk_mf_sbm_update_c needs an unsharded network");
    if (!net->have_diag) return gk_fail(ctx, GK_EINVAL, "gk_network_set_diag must be called before a mean-field update");
    if (!(stepsize >= 0.0 && stepsize <= 1.0)) return gk_fail(ctx, GK_EINVAL, "stepsize=%g outside [0,1]", stepsize);
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    const int N = net->N, B = net->B, S = 1 + B + B * (B + 1) / 2, KS = K * S;
    gk_mf_state st;
    int rc = mf_upload(ctx, net, mf_m_inout, K, &st);
    if (rc) return rc;
    // tables, transposed so the candidate class k is the fastest index
    size_t off[6];
    off[0] = 0; off[1] = off[0] + (size_t)KS * K; off[2] = off[1] + (size_t)KS * K; off[3] = off[2] + (size_t)K * K;
    off[4] = off[3] + (size_t)K * K; off[5] = off[4] + K;
    std::vector<double> blob(off[5], 0.0);
    for (int k = 0; k < K; ++k) {
        for (int k2 = 0; k2 < K; ++k2) {
            for (int s = 0; s < S; ++s) {
                blob[off[0] + ((size_t)k2 * S + s) * K + k] = coef_out[((size_t)k * K + k2) * S + s];
                blob[off[1] + ((size_t)k2 * S + s) * K + k] = coef_in[((size_t)k * K + k2) * S + s];
            }
            blob[off[2] + (size_t)k2 * K + k] = base_out[(size_t)k * K + k2];
            blob[off[3] + (size_t)k2 * K + k] = base_in[(size_t)k * K + k2];
        }
        blob[off[4] + k] = E_log_m[k];
    }
    for (size_t i = 0; i < blob.size(); ++i)
        if (!std::isfinite(blob[i])) return gk_fail(ctx, GK_EINVAL, "non-finite mean-field coefficient");
    double *tables = (double *)gk_scratch(ctx, 1, blob.size() * 8);
    if (!tables) return GK_ENOMEM;
    GK_CUDA(ctx, cudaMemcpyAsync(tables, blob.data(), blob.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
    double *out_lp_dev = nullptr;
    if (out_lp) {
        out_lp_dev = (double *)gk_scratch(ctx, 6, (size_t)N * K * 8);
        if (!out_lp_dev) return GK_ENOMEM;
    }
    int bn = block_nodes > 0 ? block_nodes : GK_MF_MAX_NB;
    bn = std::min(GK_MF_MAX_NB, std::max(GK_STRIP_COLS, bn / GK_STRIP_COLS * GK_STRIP_COLS));
    double *stats = (double *)gk_scratch(ctx, 3, (size_t)bn * 2 * KS * 8);
    if (!stats) return GK_ENOMEM;
    for (int n0 = 0; n0 < N; n0 += bn) {
        int nb = std::min(bn, N - n0), nb_pad = (nb + GK_STRIP_COLS - 1) / GK_STRIP_COLS * GK_STRIP_COLS;
#define CALL_SOFT2(T, BB) soft_strip_to_stats<T, BB>(ctx, net, &st, n0, nb_pad, 0, 2, stats)
        rc = GK_MF_DISPATCH_TB(net, CALL_SOFT2);
        if (rc) return rc;
#define CALL_MFSEQ(T, BB) launch_mf_seq<T, BB>(ctx, net, &st, tables, off, n0, nb, stats, stepsize, out_lp_dev)
        rc = GK_MF_DISPATCH_TB(net, CALL_MFSEQ);
        if (rc) return rc;
    }
    GK_CUDA(ctx, cudaMemcpyAsync(mf_m_inout, st.m64, (size_t)N * K * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_lp) GK_CUDA(ctx, cudaMemcpyAsync(out_lp, out_lp_dev, (size_t)N * K * 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GK_OK;
}
