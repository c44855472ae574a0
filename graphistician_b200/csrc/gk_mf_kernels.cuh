// gk_mf_kernels.cuh -- device kernels of the mean-field SBM path (K10).
//
// Soft strip statistics: the mean-field analogue of k_strip_stats.  For a block of nodes [n0, n0+nb) and every
// class k,   R[n][k][s] = sum_{n2 != n} mf_m[n2][k] * x_s(n, n2),   x = (a, a*w_b, a*w_b*w_b'),
// summed DOWN a column strip of A^T/W^T (out-edges of n) or A/W (in-edges).  Every partner row contributes to every
// class, so a warp owns (4 classes, a residue class of the rows); lanes own 2 adjacent nodes (128-bit W loads)
// and keep 2 x 4 x S accumulators in registers.  Reference loops replaced: stochastic_block_model.py:366-389
// (mf_update_p), :340-344 (weights of the NIW updates), :400-435 (the n2 / cn2 loops of mf_update_c).
#pragma once
#include "gk_sbm_kernels.cuh"

#define GK_SOFT_THREADS 512
#define GK_SOFT_WARPS 16
#define GK_MF_MAX_NB 256

struct gk_soft_params {
    const uint8_t *Abase[2];      // [0] = A^T rows (out-edge statistics), [1] = A rows (in-edge statistics)
    const void *Wbase[2];
    int ld, row0, nrows;
    int n0, nb_pad, dir_first;
    const void *mT;               // T [N][Kp] mean-field class probabilities, zero padded to Kp columns
    int K, Kp, KGp;               // Kp = K rounded up to 4; KGp = power of two >= Kp/4 (class groups per row group)
    int chunk_rows, nchunks;
    void *partial;                // ACC [ndirs][nchunks][nb_pad][K][S]
};

template <typename T> __device__ __forceinline__ void gk_load_m4(const T *p, T (&m)[4]);
template <> __device__ __forceinline__ void gk_load_m4<float>(const float *p, float (&m)[4]) {
    float4 v = __ldg((const float4 *)p); m[0] = v.x; m[1] = v.y; m[2] = v.z; m[3] = v.w;
}
template <> __device__ __forceinline__ void gk_load_m4<double>(const double *p, double (&m)[4]) {
    double2 a = __ldg((const double2 *)p), b = __ldg((const double2 *)p + 1); m[0] = a.x; m[1] = a.y; m[2] = b.x; m[3] = b.y;
}

template <typename T, int B>
__global__ void __launch_bounds__(GK_SOFT_THREADS, 1) k_strip_soft(gk_soft_params P) {
    constexpr int S = gk_sdim<B>::S;
    typedef typename gk_acc<T, B>::type ACC;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ACC *slots = (ACC *)smem_raw;                                  // [Kp][S][GK_SLOT_PITCH]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int dir = P.dir_first + blockIdx.z, q = blockIdx.y, strip = blockIdx.x;
    const int ra = q * P.chunk_rows;
    const int nr = min(P.chunk_rows, P.nrows - ra);
    const int KGp = P.KGp, RG = GK_SOFT_WARPS / KGp, kg = warp % KGp, rg = warp / KGp;
    const bool active = kg * 4 < P.Kp;
    const int col = strip * GK_STRIP_COLS + 2 * lane, node0 = P.n0 + col;
    const uint8_t *Ab = (dir ? P.Abase[1] : P.Abase[0]) + node0;
    const T *Wb = (const T *)(dir ? P.Wbase[1] : P.Wbase[0]) + (size_t)node0 * B;
    const T *mb = (const T *)P.mT + (size_t)(P.row0 + ra) * P.Kp + kg * 4;
    const size_t ld = P.ld;

    ACC acc[2][4][S];
#pragma unroll
    for (int e = 0; e < 2; ++e)
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
#pragma unroll
            for (int s = 0; s < S; ++s) acc[e][kk][s] = ACC(0);

    if (active) {
        for (int i = rg; i < nr; i += 4 * RG) {
            int r[4]; unsigned a2[4]; T w[4][2 * B + (B == 0)]; T mv[4][4];
#pragma unroll
            for (int j = 0; j < 4; ++j) r[j] = min(i + j * RG, nr - 1);
#pragma unroll
            for (int j = 0; j < 4; ++j) a2[j] = __ldg((const unsigned short *)(Ab + (size_t)(ra + r[j]) * ld));
#pragma unroll
            for (int j = 0; j < 4; ++j) gk_load_m4<T>(mb + (size_t)r[j] * P.Kp, mv[j]);
            if constexpr (B > 0) {
#pragma unroll
                for (int j = 0; j < 4; ++j) gk_load_w2<T, B>(Wb + (size_t)(ra + r[j]) * ld * B, w[j]);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const bool valid = i + j * RG < nr;
                const int g = P.row0 + ra + r[j];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const bool edge = valid && ((a2[j] >> (8 * e)) & 0xffu) != 0u && g != node0 + e;
                    ACC x[S];
                    x[0] = edge ? ACC(1) : ACC(0);
                    if constexpr (B > 0) {
#pragma unroll
                        for (int b = 0; b < B; ++b) x[1 + b] = edge ? (ACC)w[j][e * B + b] : ACC(0);
#pragma unroll
                        for (int b = 0; b < B; ++b)
#pragma unroll
                            for (int b2 = b; b2 < B; ++b2) x[gk_sym_idx(B, b, b2)] = x[1 + b] * x[1 + b2];
                    }
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk)
#pragma unroll
                        for (int s = 0; s < S; ++s) acc[e][kk][s] += (ACC)mv[j][kk] * x[s];
                }
            }
        }
    }
    // fold the row groups in a fixed order (deterministic): group g adds into the slots after group g-1
    for (int g = 0; g < RG; ++g) {
        if (active && rg == g) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
#pragma unroll
                for (int s = 0; s < S; ++s) {
                    ACC *sl = slots + ((size_t)(kg * 4 + kk) * S + s) * GK_SLOT_PITCH + 2 * lane;
                    if (g == 0) { sl[0] = acc[0][kk][s]; sl[1] = acc[1][kk][s]; }
                    else { sl[0] += acc[0][kk][s]; sl[1] += acc[1][kk][s]; }
                }
        }
        __syncthreads();
    }
    const int K = P.K, KS = K * S;
    ACC *out = (ACC *)P.partial + (((size_t)blockIdx.z * P.nchunks + q) * P.nb_pad + (size_t)strip * GK_STRIP_COLS) * KS;
    for (int t = tid; t < GK_STRIP_COLS * KS; t += GK_SOFT_THREADS) {
        int c = t / KS, j = t - c * KS;
        out[t] = slots[(size_t)j * GK_SLOT_PITCH + c];
    }
}

// mf_m (double [N][K]) -> compute copy T [N][Kp], zero padded
template <typename T>
__global__ void k_mf_pack(const double *__restrict__ m64, T *__restrict__ mT, int N, int K, int Kp) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < (size_t)N * Kp; i += (size_t)gridDim.x * blockDim.x) {
        int k = (int)(i % Kp); size_t n = i / Kp;
        mT[i] = k < K ? (T)m64[n * K + k] : T(0);
    }
}

// soft fold of per-node in-edge statistics into cells: cellpart[cta][kd][k*S+s] = sum_n m[n][kd] * stats[n][k*S+s]
__global__ void k_cell_partial_soft(const double *__restrict__ stats, const double *__restrict__ m64, int n0, int nb, int N,
                                    int K, int KS, double *__restrict__ cellpart) {
    extern __shared__ double tbl[];                  // [K][KS]
    for (int t = threadIdx.x; t < K * KS; t += blockDim.x) tbl[t] = 0.0;
    __syncthreads();
    const int per = (nb + gridDim.x - 1) / gridDim.x;
    const int lo = blockIdx.x * per, hi = min(nb, lo + per);
    for (int j = threadIdx.x; j < KS; j += blockDim.x)
        for (int i = lo; i < hi; ++i) {
            int n = n0 + i;
            if (n >= N) continue;
            double v = stats[(size_t)i * KS + j];
            for (int kd = 0; kd < K; ++kd) tbl[kd * KS + j] += m64[(size_t)n * K + kd] * v;
        }
    __syncthreads();
    for (int t = threadIdx.x; t < K * KS; t += blockDim.x) cellpart[(size_t)blockIdx.x * K * KS + t] = tbl[t];
}

// ---------------------------------------------------------------------------------------------------------------
// Sequential mean-field update of one node block (stochastic_block_model.py:392-445).  Node n1's update reads the
// CURRENT mf_m of every other node, so the block is processed in index order by one CTA: the contributions of all N
// partners at block entry come from the soft strip statistics; after each node's update its change delta[k'] is
// pushed to the later nodes of the block (rank-1 correction per statistic) and to the running column sums Ms.
//
// Tables (double; built on the host from mf_tau / the weight models' expectations, stored TRANSPOSED so lane k reads
// consecutive addresses):   COt[(k'*S+s)*K + k] = coefficient of out-statistic s of class k' in lp[k]
//                           CIt[(k'*S+s)*K + k] = coefficient of in-statistic  s of class k' in lp[k]
//                           BOt[k'*K + k], BIt[k'*K + k] = per-partner base terms (E ln(1-p)) forward / reverse
struct gk_mfseq_params {
    int n0, nb, K, Kp, N, ld, row0, is_f32;
    double *m64; void *mT; double *Msum;       // global state (device)
    const double *stats;                       // [nb_pad][2][K*S]
    const double *COt, *CIt, *BOt, *BIt, *logm;
    const uint8_t *diagA; const double *diagW;
    const uint8_t *A, *At; const void *W, *Wt; // local rows (must contain the block's rows)
    double stepsize;
    double *out_lp;                            // optional [N][K]
};

#define GK_MFSEQ_THREADS 1024

template <typename T, int B>
__global__ void __launch_bounds__(GK_MFSEQ_THREADS, 1) k_mf_seq(gk_mfseq_params P) {
    constexpr int S = gk_sdim<B>::S;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int K = P.K, KS = K * S, LP = K + 4;
    double *lp = (double *)smem_raw;                  // [GK_MF_MAX_NB][LP]
    double *sCO = lp + (size_t)GK_MF_MAX_NB * LP;     // [KS*K]
    double *sCI = sCO + (size_t)KS * K;
    double *sBS = sCI + (size_t)KS * K;               // [K*K]  BOt + BIt
    double *sBI = sBS + K * K;                        // [K*K]
    double *Ms = sBI + K * K;                         // [K]
    double *delta = Ms + K;                           // [K]
    double *dCO = delta + K;                          // [S*K]
    double *dCI = dCO + S * K;                        // [S*K]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    for (int t = tid; t < KS * K; t += GK_MFSEQ_THREADS) { sCO[t] = P.COt[t]; sCI[t] = P.CIt[t]; }
    for (int t = tid; t < K * K; t += GK_MFSEQ_THREADS) { sBS[t] = P.BOt[t] + P.BIt[t]; sBI[t] = P.BIt[t]; }
    if (tid < K) Ms[tid] = P.Msum[tid];
    __syncthreads();
    // base conditionals from the strip statistics (partners as they were at block entry)
    for (int t = tid; t < P.nb * K; t += GK_MFSEQ_THREADS) {
        const int i = t / K, k = t - i * K;
        const double *so = P.stats + ((size_t)i * 2 + 0) * KS, *si = P.stats + ((size_t)i * 2 + 1) * KS;
        double v = P.logm[k];
        for (int j = 0; j < KS; ++j) v += so[j] * sCO[(size_t)j * K + k] + si[j] * sCI[(size_t)j * K + k];
        lp[i * LP + k] = v;
    }
    __syncthreads();

    const int i2 = tid >> 2, kq = tid & 3;            // correction target (node of the block, class residue)
    const int n2 = P.n0 + i2;
    const bool tgt_ok = i2 < P.nb && n2 < P.N;
    // warp 0: prefetched state of the node about to be updated
    double m1 = 0.0; int a11 = 0; double w11[B + (B == 0)];
    if (warp == 0 && P.n0 < P.N) {
        if (lane < K) m1 = P.m64[(size_t)P.n0 * K + lane];
        a11 = P.diagA[P.n0];
#pragma unroll
        for (int b = 0; b < B; ++b) w11[b] = P.diagW[(size_t)P.n0 * B + b];
    }
    for (int i = 0; i < P.nb; ++i) {
        const int n = P.n0 + i;
        if (n >= P.N) break;
        // phase 0: the pair (target i2, source i) -- out-edge n2 -> n in row n of A^T, in-edge n -> n2 in row n of A
        double xo[S], xi[S];
        const bool upd = tgt_ok && i2 > i;
        if (upd) {
            const size_t rowoff = (size_t)(n - P.row0) * P.ld;
            const bool eo = P.At[rowoff + n2] != 0, ei = P.A[rowoff + n2] != 0;
            xo[0] = eo ? 1.0 : 0.0; xi[0] = ei ? 1.0 : 0.0;
            if constexpr (B > 0) {
                const T *wo = (const T *)P.Wt + (rowoff + n2) * B, *wi = (const T *)P.W + (rowoff + n2) * B;
#pragma unroll
                for (int bb = 0; bb < B; ++bb) { xo[1 + bb] = eo ? (double)wo[bb] : 0.0; xi[1 + bb] = ei ? (double)wi[bb] : 0.0; }
#pragma unroll
                for (int bb = 0; bb < B; ++bb)
#pragma unroll
                    for (int b2 = bb; b2 < B; ++b2) {
                        xo[gk_sym_idx(B, bb, b2)] = xo[1 + bb] * xo[1 + b2];
                        xi[gk_sym_idx(B, bb, b2)] = xi[1 + bb] * xi[1 + b2];
                    }
            }
        }
        // phase A: warp 0 updates node n
        if (warp == 0) {
            double v = lane < K ? lp[i * LP + lane] : 0.0, v2 = 0.0;
            double xs[S]; xs[0] = 1.0;
#pragma unroll
            for (int b = 0; b < B; ++b) xs[1 + b] = w11[b];
#pragma unroll
            for (int b = 0; b < B; ++b)
#pragma unroll
                for (int b2 = b; b2 < B; ++b2) xs[gk_sym_idx(B, b, b2)] = xs[1 + b] * xs[1 + b2];
            for (int k2 = 0; k2 < K; ++k2) {
                const double mk2 = __shfl_sync(0xffffffffu, m1, k2);
                if (lane < K) {
                    v += Ms[k2] * sBS[k2 * K + lane];
                    v2 -= mk2 * sBI[k2 * K + lane];
                    if (a11) {
                        double t = 0.0;
#pragma unroll
                        for (int s = 0; s < S; ++s) t += xs[s] * sCO[(size_t)(k2 * S + s) * K + lane];
                        v2 += mk2 * t;
                    }
                }
            }
            v += v2;
            double mx = lane < K ? v : -1e300;
            for (int o = 16; o; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            double e = lane < K ? exp(v - mx) : 0.0, tot = e;
            for (int o = 16; o; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
            const double mnew = (1.0 - P.stepsize) * m1 + P.stepsize * (e / tot);
            if (lane < K) {
                const double d = mnew - m1;
                delta[lane] = d; Ms[lane] += d;
                P.m64[(size_t)n * K + lane] = mnew;
                if (P.is_f32) ((float *)P.mT)[(size_t)n * P.Kp + lane] = (float)mnew;
                else ((double *)P.mT)[(size_t)n * P.Kp + lane] = mnew;
                if (P.out_lp) P.out_lp[(size_t)n * K + lane] = v;
            }
            // prefetch the next node's state
            if (i + 1 < P.nb && n + 1 < P.N) {
                if (lane < K) m1 = P.m64[(size_t)(n + 1) * K + lane];
                a11 = P.diagA[n + 1];
#pragma unroll
                for (int b = 0; b < B; ++b) w11[b] = P.diagW[(size_t)(n + 1) * B + b];
            }
        }
        __syncthreads();
        // phase B: delta pushed through the coefficient tables
        if (tid < 2 * S * K) {
            const int which = tid / (S * K), r = tid - which * S * K, s = r / K, k = r - s * K;
            const double *tab = which ? sCI : sCO;
            double t = 0.0;
            for (int k2 = 0; k2 < K; ++k2) t += delta[k2] * tab[(size_t)(k2 * S + s) * K + k];
            (which ? dCI : dCO)[s * K + k] = t;
        }
        __syncthreads();
        // phase C: later nodes of the block see the change
        if (upd) {
            for (int k = kq; k < K; k += 4) {
                double d = 0.0;
#pragma unroll
                for (int s = 0; s < S; ++s) d += xo[s] * dCO[s * K + k] + xi[s] * dCI[s * K + k];
                lp[i2 * LP + k] += d;
            }
        }
        __syncthreads();
    }
    if (tid < K) P.Msum[tid] = Ms[tid];
}

// cell[(k_src*K + k_dst)*S + s] += sum over CTAs of cellpart[cta][k_dst][k_src*S + s]   (fixed order)
__global__ void k_cell_final_soft(const double *__restrict__ cellpart, int nparts, int K, int S, double *__restrict__ cell) {
    const int KS = K * S;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < K * KS; t += gridDim.x * blockDim.x) {
        int kdst = t / KS, j = t - kdst * KS, ksrc = j / S, s = j - ksrc * S;
        double v = 0.0;
        for (int p = 0; p < nparts; ++p) v += cellpart[(size_t)p * K * KS + t];
        cell[((size_t)ksrc * K + kdst) * S + s] += v;
    }
}
