// gk_sbm_seq.cuh -- conditionals from statistics, the exact sequential sampler, and K1 cell reduction.
#pragma once
#include "gk_sbm_kernels.cuh"

// Tables (all double, built on the host from p / mu / Sigma, see gk_sbm.cu build_tables):
//   COt[(k'*S+s)*K + k] : coefficient of out-statistic s of block k' in lp[k]     (cell (k , k'))
//   CIt[(k'*S+s)*K + k] : coefficient of in-statistic  s of block k' in lp[k]     (cell (k', k ))
//   L1 [k*K + k']       : beta*(log(1-p[k,k']) + log(1-p[k',k]))  -- multiplies the block sizes
//   selfc[k], selfa[s*K + k] : diagonal term  selfc + a_nn * sum_s x_s selfa[s]
struct gk_tables {
    const double *COt, *CIt, *L1, *selfc, *selfa, *logm;
};

// lp_base[n][k] = log m[k] + sum_{k',s} out[k'][s] COt + in[k'][s] CIt + self term   (no block-size term)
// stats: [nb_pad][2][K*S] doubles.  A CTA takes 256 / KP nodes (KP = K rounded up to 8, 16 or 32); the two coefficient
// tables and the nodes' statistics are staged in shared memory with coalesced loads, then thread (node, k) runs the sum in
// the order of the plain loop.  (The earlier one-warp-per-node form fetched every term from L2 just before its use: 37 us
// per 1024-node block for 0.4 MFLOP.)
#define GK_LP_THREADS 256
static inline int gk_lp_kp(int K) { return K > 16 ? 32 : (K > 8 ? 16 : 8); }
static inline size_t gk_lp_smem(int K, int S) { return ((size_t)2 * K * S * K + (size_t)(GK_LP_THREADS / gk_lp_kp(K)) * 2 * K * S) * 8; }

template <int B>
__global__ void __launch_bounds__(GK_LP_THREADS)
k_lp_from_stats(const double *__restrict__ stats, gk_tables Tb, const uint8_t *__restrict__ diagA,
                const double *__restrict__ diagW, int n0, int nb, int N, int K, int KP, double *__restrict__ lp) {
    constexpr int S = gk_sdim<B>::S;
    extern __shared__ double s_lp[];
    const int KS = K * S, NPC = GK_LP_THREADS / KP;
    double *s_co = s_lp, *s_ci = s_lp + (size_t)KS * K, *s_st = s_ci + (size_t)KS * K;   // [KS][K], [KS][K], [NPC][2][KS]
    const int tid = threadIdx.x, i0 = blockIdx.x * NPC;
    for (int t = tid; t < KS * K; t += GK_LP_THREADS) { s_co[t] = Tb.COt[t]; s_ci[t] = Tb.CIt[t]; }
    const int nn = min(NPC, nb - i0);
    for (int t = tid; t < nn * 2 * KS; t += GK_LP_THREADS) s_st[t] = stats[(size_t)i0 * 2 * KS + t];
    __syncthreads();
    const int il = tid / KP, k = tid % KP, i = i0 + il, n = n0 + i;
    if (il >= nn || k >= K || n >= N) return;
    const double *so = s_st + (size_t)il * 2 * KS, *si = so + KS;
    double v = Tb.logm[k] + Tb.selfc[k];
#pragma unroll 8
    for (int j = 0; j < KS; ++j) v += so[j] * s_co[j * K + k] + si[j] * s_ci[j * K + k];
    if (diagA[n]) {
        double x[S]; x[0] = 1.0;
#pragma unroll
        for (int b = 0; b < B; ++b) x[1 + b] = diagW[(size_t)n * B + b];
#pragma unroll
        for (int b = 0; b < B; ++b)
#pragma unroll
            for (int b2 = b; b2 < B; ++b2) x[gk_sym_idx(B, b, b2)] = x[1 + b] * x[1 + b2];
#pragma unroll
        for (int s = 0; s < S; ++s) v += x[s] * Tb.selfa[(size_t)s * K + k];
    }
    lp[(size_t)i * K + k] = v;
}

__device__ __forceinline__ long long gk_sbm_clock() { long long t; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory"); return t; }

struct gk_seq_params {
    int n0, nb, K, N, ld, row0, sampler;
    int32_t *c; uint8_t *c8; int *sizes;       // global state (device)
    double *lp;                                // [nb][K] working conditionals (base, then corrected in place)
    const double *u;                           // [N] injected uniforms
    gk_tables T;
    const uint8_t *A, *At; const void *W, *Wt; // local rows (must contain the block's rows)
    double *out_lp;                            // optional [N][K]
    unsigned long long *nchanged;              // [0] moves; [1..6] phase clocks when prof
    int prof;
};

#define GK_SEQ_THREADS 1024
#define GK_SEQ_WARPS 32

// Exact serial Gibbs over nodes n0..n0+nb-1, run speculatively: all pending nodes of a window draw from their
// current conditionals in parallel; everything up to and including the first node whose label changes is
// committed; that change is pushed (rank-1 correction of lp) to the later nodes of the block; repeat.
// A committed prefix saw exactly the state the serial loop would have seen, so results are identical.
template <typename T, int B>
__global__ void __launch_bounds__(GK_SEQ_THREADS, 1) k_sbm_seq(gk_seq_params P) {
    constexpr int S = gk_sdim<B>::S;
    __shared__ double T0[GK_MAX_K], dL1[GK_MAX_K], dCO[S * GK_MAX_K], dCI[S * GK_MAX_K];
    __shared__ int ssz[GK_MAX_K];
    __shared__ int s_first;
    __shared__ unsigned char s_new[4096];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, K = P.K;
    if (tid < K) ssz[tid] = P.sizes[tid];
    if (tid == 0) s_first = 0x7fffffff;
    __syncthreads();
    if (tid < K) { double t = 0.0; for (int k2 = 0; k2 < K; ++k2) t += (double)ssz[k2] * P.T.L1[tid * K + k2]; T0[tid] = t; }
    __syncthreads();
    int cursor = 0, window = 64;
    unsigned long long changed = 0;
    while (cursor < P.nb) {
        const int wend = min(P.nb, cursor + window);
        for (int i = cursor + warp; i < wend; i += GK_SEQ_WARPS) {
            const int n = P.n0 + i;
            if (n >= P.N) { if (lane == 0) s_new[i - cursor] = 255; continue; }
            const int ci = P.c[n];
            double v = lane < K ? P.lp[(size_t)i * K + lane] + T0[lane] - P.T.L1[lane * K + ci] : -1e300;
            double mx = v;
            for (int o = 16; o; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            double e = lane < K ? exp(v - mx) : 0.0;
            int newc;
            if (P.sampler == GK_SAMPLER_CUMSUM) {
                double cum = e;
                for (int o = 1; o < 32; o <<= 1) { double t = __shfl_up_sync(0xffffffffu, cum, o); if (lane >= o) cum += t; }
                double tot = __shfl_sync(0xffffffffu, cum, K - 1);
                double r = P.u[n] * tot;
                newc = __popc(__ballot_sync(0xffffffffu, lane < K && r > cum));
            } else {
                double tot = e;
                for (int o = 16; o; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
                double cum = e / tot;
                for (int o = 1; o < 32; o <<= 1) { double t = __shfl_up_sync(0xffffffffu, cum, o); if (lane >= o) cum += t; }
                newc = __popc(__ballot_sync(0xffffffffu, lane < K && cum < P.u[n]));
            }
            newc = min(newc, K - 1);
            if (P.out_lp && lane < K) P.out_lp[(size_t)n * K + lane] = v;
            if (lane == 0) { s_new[i - cursor] = (unsigned char)newc; if (newc != ci) atomicMin(&s_first, i); }
        }
        __syncthreads();
        const int first = s_first;
        __syncthreads();
        if (first >= wend) {                       // nothing in the window moved: commit it all
            cursor = wend; window = min(window * 2, 4096);
            continue;
        }
        // commit nodes cursor..first; node `first` moves a -> b
        const int ns = P.n0 + first, a = P.c[ns], b = s_new[first - cursor];
        if (tid < K) {
            dL1[tid] = P.T.L1[tid * K + b] - P.T.L1[tid * K + a];
            for (int s = 0; s < S; ++s) {
                dCO[s * K + tid] = P.T.COt[(size_t)(b * S + s) * K + tid] - P.T.COt[(size_t)(a * S + s) * K + tid];
                dCI[s * K + tid] = P.T.CIt[(size_t)(b * S + s) * K + tid] - P.T.CIt[(size_t)(a * S + s) * K + tid];
            }
        }
        __syncthreads();
        const size_t rowoff = (size_t)(ns - P.row0) * P.ld;
        const int rem = P.nb - first - 1;
        for (int t = tid; t < rem * K; t += GK_SEQ_THREADS) {
            const int i = first + 1 + t / K, k = t % K, n = P.n0 + i;
            if (n >= P.N) continue;
            // out-edge n -> ns lives in row ns of A^T; in-edge ns -> n lives in row ns of A
            double xo[S], xi[S];
            const bool eo = P.At[rowoff + n] != 0, ei = P.A[rowoff + n] != 0;
            xo[0] = eo ? 1.0 : 0.0; xi[0] = ei ? 1.0 : 0.0;
            if constexpr (B > 0) {
                const T *wo = (const T *)P.Wt + (rowoff + n) * B, *wi = (const T *)P.W + (rowoff + n) * B;
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
            double d = dL1[k];
#pragma unroll
            for (int s = 0; s < S; ++s) d += xo[s] * dCO[s * K + k] + xi[s] * dCI[s * K + k];
            P.lp[(size_t)i * K + k] += d;
        }
        if (tid == 0) {
            P.c[ns] = b; P.c8[ns] = (uint8_t)b;
            P.sizes[a] -= 1; P.sizes[b] += 1;
            s_first = 0x7fffffff;
        }
        changed++;
        cursor = first + 1; window = max(64, window / 2);
        __syncthreads();
    }
    if (tid == 0 && P.nchanged) atomicAdd(P.nchanged, changed);
}

__device__ __forceinline__ void gk_prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// Shared-memory form of k_sbm_seq for blocks whose conditionals fit on chip (nb * K * 8 B <= ~190 KB): same speculative
// windows and the same arithmetic, but lp, labels and uniforms of the block live in shared memory ([k][i], odd pitch),
// a label change is pushed by ONE pass with a thread per later node (coalesced reads of row ns of A, A^T, W, W^T), and
// the warp that finds the earliest change so far prefetches those row segments into L2 while the window finishes.
// A sweep from random labels (every node moves) costs ~1/4 of the global-memory form; a converged chain is unchanged.
template <typename T, int B>
__global__ void __launch_bounds__(GK_SEQ_THREADS, 1) k_sbm_seq2(gk_seq_params P) {
    constexpr int S = gk_sdim<B>::S;
    extern __shared__ __align__(16) double seq_dyn[];
    __shared__ __align__(16) double dL1[GK_MAX_K], dCO[S * GK_MAX_K], dCI[S * GK_MAX_K];
    __shared__ double T0[GK_MAX_K], L1s[GK_MAX_K * GK_MAX_K];
    __shared__ int s_first, ssz[GK_MAX_K];
    __shared__ unsigned char s_new[4096];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, K = P.K, nb = P.nb;
    const int pitch = nb | 1;
    const bool prof = P.prof && tid == 0;
    long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tk = 0;
    double *lp_s = seq_dyn, *u_s = lp_s + (size_t)K * pitch;
    unsigned char *c_s = (unsigned char *)(u_s + nb);
    for (int t = tid; t < nb * K; t += GK_SEQ_THREADS) lp_s[(t % K) * pitch + t / K] = P.lp[t];
    for (int i = tid; i < nb; i += GK_SEQ_THREADS) {
        const int n = P.n0 + i;
        u_s[i] = n < P.N ? P.u[n] : 0.0;
        c_s[i] = n < P.N ? P.c8[n] : 255;
    }
    for (int t = tid; t < K * K; t += GK_SEQ_THREADS) L1s[t] = P.T.L1[t];
    for (int t = tid; t < S * GK_MAX_K; t += GK_SEQ_THREADS) { dCO[t] = 0.0; dCI[t] = 0.0; }
    if (tid < GK_MAX_K) dL1[tid] = 0.0;
    if (tid == 0) s_first = 0x7fffffff;
    if (tid < K) ssz[tid] = P.sizes[tid];
    __syncthreads();
    if (tid < K) { double t = 0.0; for (int k2 = 0; k2 < K; ++k2) t += (double)ssz[k2] * L1s[tid * K + k2]; T0[tid] = t; }
    __syncthreads();
    const size_t esz = B * sizeof(T);
    const int W = K <= 16 ? 16 : 32, npw = 32 / W, sub = lane / W, l = lane % W;
    const unsigned hbits = W == 32 ? 0xffffffffu : 0xffffu, hmask = hbits << (W * sub);
    int cursor = 0, window = 64;
    unsigned long long changed = 0;
    while (cursor < nb) {
        const int wend = min(nb, cursor + window);
        if (prof) tk = gk_sbm_clock();
        // K <= 16: two nodes per warp (one per 16-lane half) -- same shuffle trees, so the same bits as one per warp
        for (int ib = cursor + npw * warp; ib < wend; ib += npw * GK_SEQ_WARPS) {
            const int i = ib + sub, n = P.n0 + i;
            const bool inw = i < wend, act = inw && n < P.N;
            const int ci = act ? c_s[i] : 0;
            double v = (act && l < K) ? lp_s[l * pitch + i] + T0[l] - L1s[l * K + ci] : -1e300;
            double mx = v;
            for (int o = W >> 1; o; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o, W));
            double e = (act && l < K) ? exp(v - mx) : 0.0;
            const double un = act ? u_s[i] : 0.0;
            unsigned bal;
            if (P.sampler == GK_SAMPLER_CUMSUM) {
                double cum = e;
                for (int o = 1; o < W; o <<= 1) { double t = __shfl_up_sync(0xffffffffu, cum, o, W); if (l >= o) cum += t; }
                double tot = __shfl_sync(0xffffffffu, cum, K - 1, W);
                double r = un * tot;
                bal = __ballot_sync(0xffffffffu, l < K && r > cum);
            } else {
                double tot = e;
                for (int o = W >> 1; o; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o, W);
                double cum = e / tot;
                for (int o = 1; o < W; o <<= 1) { double t = __shfl_up_sync(0xffffffffu, cum, o, W); if (l >= o) cum += t; }
                bal = __ballot_sync(0xffffffffu, l < K && cum < un);
            }
            const int newc = min(__popc((bal >> (W * sub)) & hbits), K - 1);
            if (act && P.out_lp && l < K) P.out_lp[(size_t)n * K + l] = v;
            if (inw && l == 0) s_new[i - cursor] = act ? (unsigned char)newc : 255;
            if (act && newc != ci) {                         // uniform over the half-warp
                int prev = 0;
                if (l == 0) prev = atomicMin(&s_first, i);
                prev = __shfl_sync(hmask, prev, 0, W);
                if (i < prev) {                              // earliest mover so far: warm L2 with its row segments
                    const size_t off = (size_t)(n - P.row0) * P.ld + P.n0;
                    for (int bb = l * 128; bb < nb; bb += W * 128) { gk_prefetch_l2(P.A + off + bb); gk_prefetch_l2(P.At + off + bb); }
                    if constexpr (B > 0)
                        for (size_t bb = (size_t)l * 128; bb < nb * esz; bb += W * 128) {
                            gk_prefetch_l2((const char *)P.W + off * esz + bb); gk_prefetch_l2((const char *)P.Wt + off * esz + bb);
                        }
                }
            }
        }
        if (prof) { long long t = gk_sbm_clock(); pc[0] += t - tk; tk = t; }
        __syncthreads();
        const int first = s_first;
        __syncthreads();
        if (prof) { long long t = gk_sbm_clock(); pc[1] += t - tk; tk = t; }
        if (first >= wend) {                       // nothing in the window moved: commit it all
            cursor = wend; window = min(window * 2, 4096);
            continue;
        }
        // commit nodes cursor..first; node `first` moves a -> b
        const int ns = P.n0 + first, a = c_s[first], b = s_new[first - cursor];
        if (tid < K) {
            dL1[tid] = L1s[tid * K + b] - L1s[tid * K + a];
            for (int s = 0; s < S; ++s) {
                dCO[s * GK_MAX_K + tid] = P.T.COt[(size_t)(b * S + s) * K + tid] - P.T.COt[(size_t)(a * S + s) * K + tid];
                dCI[s * GK_MAX_K + tid] = P.T.CIt[(size_t)(b * S + s) * K + tid] - P.T.CIt[(size_t)(a * S + s) * K + tid];
            }
        }
        const size_t rowoff = (size_t)(ns - P.row0) * P.ld;
        // features of the later nodes against ns (nb <= GK_SEQ_THREADS: one later node per thread); the loads are
        // issued before the barrier that publishes the delta tables
        const int i = first + 1 + tid, n = P.n0 + i;
        const bool act = i < nb && n < P.N;
        double xo[S], xi[S];
        if (act) {
            const bool eo = P.At[rowoff + n] != 0, ei = P.A[rowoff + n] != 0;
            xo[0] = eo ? 1.0 : 0.0; xi[0] = ei ? 1.0 : 0.0;
            if constexpr (B > 0) {
                const T *wo = (const T *)P.Wt + (rowoff + n) * B, *wi = (const T *)P.W + (rowoff + n) * B;
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
        if (prof) { long long t = gk_sbm_clock(); pc[2] += t - tk; tk = t; }
        __syncthreads();
        if (prof) { long long t = gk_sbm_clock(); pc[3] += t - tk; tk = t; }
        if (act) {
            for (int k = 0; k < K; k += 2) {
                const double2 l1 = *(const double2 *)&dL1[k];
                double d0 = l1.x, d1 = l1.y;
#pragma unroll
                for (int s = 0; s < S; ++s) {
                    const double2 co = *(const double2 *)&dCO[s * GK_MAX_K + k], cin = *(const double2 *)&dCI[s * GK_MAX_K + k];
                    d0 += xo[s] * co.x + xi[s] * cin.x;
                    d1 += xo[s] * co.y + xi[s] * cin.y;
                }
                lp_s[k * pitch + i] += d0;
                if (k + 1 < K) lp_s[(k + 1) * pitch + i] += d1;
            }
        }
        if (tid == 0) {
            P.c[ns] = b; P.c8[ns] = (uint8_t)b; c_s[first] = (unsigned char)b;
            ssz[a] -= 1; ssz[b] += 1;
            s_first = 0x7fffffff;
        }
        changed++;
        cursor = first + 1; window = max(32 * npw, window / 2);
        if (prof) { long long t = gk_sbm_clock(); pc[4] += t - tk; tk = t; }
        __syncthreads();
        if (prof) { long long t = gk_sbm_clock(); pc[5] += t - tk; tk = t; }
    }
    if (tid < K) P.sizes[tid] = ssz[tid];
    if (tid == 0 && P.nchanged) atomicAdd(P.nchanged, changed);
    if (prof) for (int q = 0; q < 6; ++q) atomicAdd(P.nchanged + 1 + q, (unsigned long long)pc[q]);
}

// K1: fold per-node in-edge statistics  stats[n1][bin = c[n2]][s]  into cells (c[n2], c[n1]).
// Each CTA owns a contiguous node range; thread j (< K*S) owns column j of a private [K][K*S] table
// (no atomics, fixed order).  cellpart: [gridDim.x][K][K*S].
__global__ void k_cell_partial(const double *__restrict__ stats, const uint8_t *__restrict__ c8, int n0, int nb, int N,
                               int K, int KS, double *__restrict__ cellpart) {
    extern __shared__ double tbl[];                  // [K][KS]
    for (int t = threadIdx.x; t < K * KS; t += blockDim.x) tbl[t] = 0.0;
    __syncthreads();
    const int per = (nb + gridDim.x - 1) / gridDim.x;
    const int lo = blockIdx.x * per, hi = min(nb, lo + per);
    for (int j = threadIdx.x; j < KS; j += blockDim.x)
        for (int i = lo; i < hi; ++i) {
            int n = n0 + i;
            if (n < N) tbl[c8[n] * KS + j] += stats[(size_t)i * KS + j];
        }
    __syncthreads();
    for (int t = threadIdx.x; t < K * KS; t += blockDim.x) cellpart[(size_t)blockIdx.x * K * KS + t] = tbl[t];
}

// cell[(k_src*K + k_dst)*S + s] += sum over CTAs  (cellpart index: [cta][k_dst][k_src*S + s])
__global__ void k_cell_final(const double *__restrict__ cellpart, int nparts, int K, int S, double *__restrict__ cell) {
    const int KS = K * S;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < K * KS; t += gridDim.x * blockDim.x) {
        int kdst = t / KS, j = t - kdst * KS, ksrc = j / S, s = j - ksrc * S;
        double v = 0.0;
        for (int p = 0; p < nparts; ++p) v += cellpart[(size_t)p * K * KS + t];
        cell[((size_t)ksrc * K + kdst) * S + s] += v;
    }
}

// The same fold over the statistics a whole SWEEP left behind (one row per node: [N_pad][2][K*S], in-edge half used), with the
// labels the sweep ended on: cell[(k_src, k_dst)] += sum_{n : c'[n] = k_dst} in_stats[n][k_src].  Exact for every pair whose source
// kept its label after the target's block was entered; k_cell_movers corrects the others.
__global__ void k_cell_partial2(const double *__restrict__ stats, const uint8_t *__restrict__ c8, int N, int K, int KS,
                                double *__restrict__ cellpart) {
    extern __shared__ double tbl[];                  // [K][KS]
    for (int t = threadIdx.x; t < K * KS; t += blockDim.x) tbl[t] = 0.0;
    __syncthreads();
    const int per = (N + gridDim.x - 1) / gridDim.x;
    const int lo = blockIdx.x * per, hi = min(N, lo + per);
    for (int j = threadIdx.x; j < KS; j += blockDim.x)
        for (int n = lo; n < hi; ++n) tbl[c8[n] * KS + j] += stats[((size_t)n * 2 + 1) * KS + j];
    __syncthreads();
    for (int t = threadIdx.x; t < K * KS; t += blockDim.x) cellpart[(size_t)blockIdx.x * K * KS + t] = tbl[t];
}

// nodes whose label changed during the sweep, in index order is not needed: (node, old, new) triples, count in cnt[0]
__global__ void k_find_movers(const uint8_t *__restrict__ c_old, const uint8_t *__restrict__ c_new, int N, int cap,
                              int *__restrict__ cnt, int *__restrict__ list) {
    for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < N; n += gridDim.x * blockDim.x)
        if (c_old[n] != c_new[n]) {
            const int i = atomicAdd(cnt, 1);
            if (i < cap) { list[3 * i] = n; list[3 * i + 1] = c_old[n]; list[3 * i + 2] = c_new[n]; }
        }
}

// Correction of the folded cell table for the sweep's movers.  Node m moved a -> b while its block [b0, b1) was drawn; every node
// n < b1 (n != m) had its in-statistics taken when m still carried a, so x(m -> n) sits in cell (a, c'[n]) and belongs in (b, c'[n]).
// One CTA per mover, 64 threads, each with a private [K][S] table in shared memory (no atomics, layout [entry][thread]); the
// tables are summed in a fixed order and applied to the global cell table with atomics (the table is tiny and movers are few:
// the host falls back to a full K1 pass when a sweep moved more than `cap` nodes).
template <typename T, int B>
__global__ void __launch_bounds__(64) k_cell_movers(const uint8_t *__restrict__ A, const void *__restrict__ Wv, int ld, const uint8_t *__restrict__ c8,
                                                    const int *__restrict__ cnt, const int *__restrict__ list, int cap, int block_nodes,
                                                    int N, int K, double *__restrict__ cell) {
    constexpr int S = gk_sdim<B>::S;
    extern __shared__ double priv[];                 // [K*S][64]
    const int nm = min(cnt[0], cap), tid = threadIdx.x, KS = K * S;
    const T *W = (const T *)Wv;
    for (int i = blockIdx.x; i < nm; i += gridDim.x) {
        const int m = list[3 * i], a = list[3 * i + 1], b = list[3 * i + 2];
        const int b1 = min(N, (m / block_nodes + 1) * block_nodes);
        for (int t = 0; t < KS; ++t) priv[t * 64 + tid] = 0.0;
        const uint8_t *Ar = A + (size_t)m * ld;
        const T *Wr = W + (size_t)m * ld * B;
        for (int n = tid; n < b1; n += 64) {
            if (n == m || !Ar[n]) continue;
            double *pt = priv + (size_t)c8[n] * S * 64 + tid;
            pt[0] += 1.0;
            if constexpr (B > 0) {
                double w[B];
#pragma unroll
                for (int q = 0; q < B; ++q) { w[q] = (double)Wr[(size_t)n * B + q]; pt[(1 + q) * 64] += w[q]; }
#pragma unroll
                for (int q = 0; q < B; ++q)
#pragma unroll
                    for (int q2 = q; q2 < B; ++q2) pt[gk_sym_idx(B, q, q2) * 64] += w[q] * w[q2];
            }
        }
        __syncthreads();
        for (int t = tid; t < KS; t += 64) {
            double v = 0.0;
            for (int j = 0; j < 64; ++j) v += priv[t * 64 + j];
            if (v != 0.0) {
                const int kd = t / S, sidx = t - kd * S;
                atomicAdd(&cell[((size_t)a * K + kd) * S + sidx], -v);
                atomicAdd(&cell[((size_t)b * K + kd) * S + sidx], v);
            }
        }
        __syncthreads();
    }
}

// Speculative strip statistics (gk_sbm_sweep overlaps the strip pass of node block b+1 with the serial kernel of block b):
// the statistics of the nodes [n0, n0+nb) were taken with the labels `c_spec` that block [p0, p0+pn) carried BEFORE it was
// drawn; every node m of that block whose label changed a -> b' moves its edge terms x(n -> m) (out half) and x(m -> n)
// (in half) from class a to class b'.  Counts stay exact integers; the fp sums change by round-off only.  One thread per
// node n (it owns its two statistics rows: no atomics); the mover list is rebuilt per CTA in shared memory, in index order.
template <typename T, int B>
__global__ void __launch_bounds__(128) k_stats_fix(const uint8_t *__restrict__ A, const uint8_t *__restrict__ At, const void *__restrict__ Wv,
                                                   const void *__restrict__ Wtv, int ld, const uint8_t *__restrict__ c_spec,
                                                   const uint8_t *__restrict__ c8, int p0, int pn, int n0, int nb, int K,
                                                   double *__restrict__ stats) {
    constexpr int S = gk_sdim<B>::S;
    __shared__ int s_cnt;
    __shared__ unsigned short s_m[2048];
    __shared__ unsigned char s_a[2048], s_b[2048];
    __shared__ int s_w[4];
    const int tid = threadIdx.x;
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    for (int base = 0; base < pn; base += 128) {                    // ordered compaction: the list is the same in every CTA
        const int i = base + tid;
        const bool mv = i < pn && c_spec[p0 + i] != c8[p0 + i];
        const unsigned bal = __ballot_sync(0xffffffffu, mv);
        if ((tid & 31) == 0) s_w[tid >> 5] = __popc(bal);
        __syncthreads();
        int off = s_cnt;
        for (int w = 0; w < (tid >> 5); ++w) off += s_w[w];
        if (mv) {
            const int pos = off + __popc(bal & ((1u << (tid & 31)) - 1));
            s_m[pos] = (unsigned short)i; s_a[pos] = c_spec[p0 + i]; s_b[pos] = c8[p0 + i];
        }
        __syncthreads();
        if (tid == 0) s_cnt += s_w[0] + s_w[1] + s_w[2] + s_w[3];
        __syncthreads();
    }
    const int nm = s_cnt;
    // thread = (node, direction): 64 nodes per CTA, each thread owns one statistics row (no atomics, fixed order)
    const int dir = tid >> 6, i = blockIdx.x * 64 + (tid & 63);
    if (nm == 0 || i >= nb) return;
    const int n = n0 + i, KS = K * S;
    const uint8_t *E = dir ? A : At;                                // dir 0: edge n -> m (out half), dir 1: edge m -> n (in half)
    const T *X = (const T *)(dir ? Wv : Wtv);
    double *row = stats + ((size_t)i * 2 + dir) * KS;
    for (int q0 = 0; q0 < nm; q0 += 8) {
        unsigned char e[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) e[t] = (q0 + t < nm) ? E[(size_t)(p0 + s_m[q0 + t]) * ld + n] : 0;
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            if (!e[t]) continue;
            const int m = p0 + s_m[q0 + t], a = s_a[q0 + t], b = s_b[q0 + t];
            double x[S]; x[0] = 1.0;
            if constexpr (B > 0) {
#pragma unroll
                for (int d = 0; d < B; ++d) x[1 + d] = (double)X[((size_t)m * ld + n) * B + d];
#pragma unroll
                for (int d = 0; d < B; ++d)
#pragma unroll
                    for (int d2 = d; d2 < B; ++d2) x[gk_sym_idx(B, d, d2)] = x[1 + d] * x[1 + d2];
            }
#pragma unroll
            for (int u = 0; u < S; ++u) { row[a * S + u] -= x[u]; row[b * S + u] += x[u]; }
        }
    }
}

// Whole-sweep speculation (gk_sbm_sweep, GK_CTX_OPT_SBM_OVERLAP = 3): the strip passes of ALL node blocks of the next sweep run
// in the background from the labels the previous sweep ended on.  k_append_movers lists, block after block, the nodes whose label
// differs from that snapshot (index order; a node moves at most once per sweep, so N entries suffice); k_stats_fix_list moves the
// terms of every listed mover in the rows of the block about to be drawn -- the list then holds exactly the movers of the earlier
// blocks of this sweep.
__global__ void __launch_bounds__(1024) k_append_movers(const uint8_t *__restrict__ c_spec, const uint8_t *__restrict__ c8, int p0, int pn,
                                                        int *__restrict__ cnt, int *__restrict__ list) {
    __shared__ int s_w[32], s_base;
    const int tid = threadIdx.x;
    if (tid == 0) s_base = *cnt;
    for (int base = 0; base < pn; base += 1024) {
        const int i = base + tid;
        const bool mv = i < pn && c_spec[p0 + i] != c8[p0 + i];
        const unsigned bal = __ballot_sync(0xffffffffu, mv);
        if ((tid & 31) == 0) s_w[tid >> 5] = __popc(bal);
        __syncthreads();
        int off = s_base;
        for (int w = 0; w < (tid >> 5); ++w) off += s_w[w];
        if (mv) {
            const int pos = off + __popc(bal & ((1u << (tid & 31)) - 1));
            list[3 * pos] = p0 + i; list[3 * pos + 1] = c_spec[p0 + i]; list[3 * pos + 2] = c8[p0 + i];
        }
        __syncthreads();
        if (tid == 0) { int t = 0; for (int w = 0; w < 32; ++w) t += s_w[w]; s_base += t; }
        __syncthreads();
    }
    if (tid == 0) *cnt = s_base;
}

template <typename T, int B>
__global__ void __launch_bounds__(128) k_stats_fix_list(const uint8_t *__restrict__ A, const uint8_t *__restrict__ At, const void *__restrict__ Wv,
                                                        const void *__restrict__ Wtv, int ld, const int *__restrict__ cnt,
                                                        const int *__restrict__ list, int n0, int nb, int K, double *__restrict__ stats) {
    constexpr int S = gk_sdim<B>::S;
    const int nm = *cnt, tid = threadIdx.x;
    const int dir = tid >> 6, i = blockIdx.x * 64 + (tid & 63);
    if (nm == 0 || i >= nb) return;
    const int n = n0 + i, KS = K * S;
    const uint8_t *E = dir ? A : At;                                // dir 0: edge n -> m (out half), dir 1: edge m -> n (in half)
    const T *X = (const T *)(dir ? Wv : Wtv);
    double *row = stats + ((size_t)i * 2 + dir) * KS;
    for (int q0 = 0; q0 < nm; q0 += 8) {
        unsigned char e[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) e[t] = (q0 + t < nm) ? E[(size_t)list[3 * (q0 + t)] * ld + n] : 0;
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            if (!e[t]) continue;
            const int m = list[3 * (q0 + t)], a = list[3 * (q0 + t) + 1], b = list[3 * (q0 + t) + 2];
            double x[S]; x[0] = 1.0;
            if constexpr (B > 0) {
#pragma unroll
                for (int d = 0; d < B; ++d) x[1 + d] = (double)X[((size_t)m * ld + n) * B + d];
#pragma unroll
                for (int d = 0; d < B; ++d)
#pragma unroll
                    for (int d2 = d; d2 < B; ++d2) x[gk_sym_idx(B, d, d2)] = x[1 + d] * x[1 + d2];
            }
#pragma unroll
            for (int u = 0; u < S; ++u) { row[a * S + u] -= x[u]; row[b * S + u] += x[u]; }
        }
    }
}

// c (int32) -> c8 and block sizes are prepared on the host; this just narrows the labels on the device
__global__ void k_narrow_labels(const int32_t *__restrict__ c, uint8_t *__restrict__ c8, int N) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) c8[i] = (uint8_t)c[i];
}
