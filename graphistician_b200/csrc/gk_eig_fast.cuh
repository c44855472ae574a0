// gk_eig_fast.cuh -- "running G" form of the sequential feature sweep (K4 + K5 + K6) for unsharded networks.
//
// eigenmodel.py:876-908 draws node n from a conditional built from  G[n] = sum_{m != n} X[n,m] Phi(f_m)  with the
// features every other node has AT THAT MOMENT (new for m < n, old for m > n).  Instead of re-contracting a slab
// of S per node block (k_eig_slab, kept for the row-sharded block form), this path keeps G for ALL nodes in HBM:
//   k_eig_gfull   once per sweep:  G[n] = sum_m X[m,n] Phi(f_m_old)          (one streaming pass over S and Asym)
//   k_eig_seq3    per node block:  the serial draws; in-block nodes j > s see node s's change immediately
//   k_eig_push    per node block:  G[n] += sum_{m in block} X[m,n] (Phi(f_m_new) - Phi(f_m_old)) for every other n
//                                   (rows of the block are contiguous in S: each CTA owns a column strip, so the
//                                    128-row contraction is CTA-local -- no atomics, no partial buffers)
//   k_eig_gstats  once per sweep:  at the end G holds the all-new contraction, from which the pair statistics of
//                                   eigenmodel.py:910-965 follow as 1/2 sum_n Phi(f_n) . G[n]
// S and Asym are each read twice per sweep (gfull + push) and nothing N x N is written.
#pragma once
#include "gk_eig_kernels.cuh"

#define GK_EF_THREADS 256
#define GK_EF_COLS 128            // columns per CTA strip (2 per thread-column)
#define GK_EF_PIECE 64            // rows of Phi staged in shared memory at a time

template <typename T> struct gk_vec2;
template <> struct gk_vec2<float> { typedef float2 type; };
template <> struct gk_vec2<double> { typedef double2 type; };

// Phi(f) = [f_a f_b (a <= b) | f_a | f_a] for one row into shared memory
template <typename ACC, int D>
__device__ __forceinline__ void gk_phi_row(const double *__restrict__ f, ACC *__restrict__ ph) {
    constexpr int Q = gk_edim<D>::Q, Pn = gk_edim<D>::P;
#pragma unroll
    for (int a = 0; a < D; ++a) {
#pragma unroll
        for (int b = a; b < D; ++b) ph[gk_qidx(D, a, b)] = (ACC)(f[a] * f[b]);
        ph[Q + a] = (ACC)f[a];
        ph[Pn + a] = (ACC)f[a];
    }
}

// One thread = (column pair cp, half of the PA output columns, row group rg of 2).  Accumulates rows
// [0, mrows) of a piece whose Phi rows sit in shared memory.  mg0 = global index of the piece's first row;
// colg = global index of the thread's first column; a pair (m, col) counts iff m > mlow[e].
template <typename T, int D, int HALF>
__device__ __forceinline__ void gk_ef_piece(const T *__restrict__ Sp, const uint8_t *__restrict__ Ap, size_t ld, int mrows,
                                            const typename gk_eacc<T>::type *__restrict__ phi, int rg, bool probit, int mg0,
                                            int colg, int mlow0, int mlow1,
                                            typename gk_eacc<T>::type (&acc)[2][gk_edim<D>::PH]) {
    typedef typename gk_eacc<T>::type ACC;
    typedef typename gk_vec2<T>::type T2;
    constexpr int Pn = gk_edim<D>::P, PA = gk_edim<D>::PA, PH = gk_edim<D>::PH;
    constexpr int p0 = HALF * PH, np = HALF == 0 ? PH : PA - PH;
    for (int m = rg; m < mrows; m += 8) {
        ACC s[4][2], a[4][2];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int mm = m + 2 * u;
            const bool ok = mm < mrows;
            const int mc = ok ? mm : m;
            const T2 sv = __ldg((const T2 *)(Sp + (size_t)mc * ld));
            const int mg = mg0 + mc;
            const bool k0 = ok && mg > mlow0, k1 = ok && mg > mlow1;
            if (probit) {
                s[u][0] = (k0 && mg != colg) ? ACC(2) : ACC(0);
                s[u][1] = (k1 && mg != colg + 1) ? ACC(2) : ACC(0);
                a[u][0] = k0 ? (ACC)sv.x : ACC(0);
                a[u][1] = k1 ? (ACC)sv.y : ACC(0);
            } else {
                const unsigned av = __ldg((const unsigned short *)(Ap + (size_t)mc * ld));
                s[u][0] = k0 ? (ACC)sv.x : ACC(0);
                s[u][1] = k1 ? (ACC)sv.y : ACC(0);
                a[u][0] = k0 ? (ACC)(av & 0xffu) : ACC(0);
                a[u][1] = k1 ? (ACC)(av >> 8) : ACC(0);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int mm = m + 2 * u;
            const ACC *ph = phi + (size_t)(mm < mrows ? mm : m) * PA + p0;
#pragma unroll
            for (int p = 0; p < np; ++p) {
                const ACC v = ph[p];
                acc[0][p] += ((p0 + p) < Pn ? s[u][0] : a[u][0]) * v;
                acc[1][p] += ((p0 + p) < Pn ? s[u][1] : a[u][1]) * v;
            }
        }
    }
}

// fold the two row groups of a CTA through shared memory; on return the rg == 0 threads hold the sums
template <typename ACC, int NP>
__device__ __forceinline__ void gk_ef_fold(ACC (&acc)[2][NP], ACC *__restrict__ scratch, int tid) {
    // scratch: [128][2*NP + 1]
    const int slot = tid & 127, rg = tid >> 7;
    __syncthreads();
    if (rg == 1) {
#pragma unroll
        for (int e = 0; e < 2; ++e)
#pragma unroll
            for (int p = 0; p < NP; ++p) scratch[(size_t)slot * (2 * NP + 1) + e * NP + p] = acc[e][p];
    }
    __syncthreads();
    if (rg == 0) {
#pragma unroll
        for (int e = 0; e < 2; ++e)
#pragma unroll
            for (int p = 0; p < NP; ++p) acc[e][p] += scratch[(size_t)slot * (2 * NP + 1) + e * NP + p];
    }
}

// ---------------------------------------------------------------------------------------------------
// k_eig_gfull: partial[chunk][n][p] = sum_{m in chunk} X[m,n] Phi(f_m)[p]   (grid: column strips x row chunks)
// ---------------------------------------------------------------------------------------------------
template <typename T, int D>
__global__ void __launch_bounds__(GK_EF_THREADS, 2) k_eig_gfull(const T *__restrict__ S, const uint8_t *__restrict__ Asym, int ld,
                                                                 int N, const double *__restrict__ F, int chunk_rows, int npad,
                                                                 typename gk_eacc<T>::type *__restrict__ partial, int probit) {
    typedef typename gk_eacc<T>::type ACC;
    constexpr int PA = gk_edim<D>::PA, PH = gk_edim<D>::PH;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ACC *phi = (ACC *)smem_raw;                                   // [GK_EF_PIECE][PA]; reused as fold scratch
    const int tid = threadIdx.x, cp = tid & 63, half = (tid >> 6) & 1, rg = tid >> 7;
    const int c0 = blockIdx.x * GK_EF_COLS, colg = c0 + 2 * cp;
    const int r0 = blockIdx.y * chunk_rows, r1 = min(N, r0 + chunk_rows);
    ACC acc[2][PH];
#pragma unroll
    for (int e = 0; e < 2; ++e)
#pragma unroll
        for (int p = 0; p < PH; ++p) acc[e][p] = ACC(0);
    for (int rp = r0; rp < r1; rp += GK_EF_PIECE) {
        const int mrows = min(GK_EF_PIECE, r1 - rp);
        __syncthreads();
        if (tid < mrows) {
            double f[D];
#pragma unroll
            for (int d = 0; d < D; ++d) f[d] = F[(size_t)(rp + tid) * D + d];
            gk_phi_row<ACC, D>(f, phi + (size_t)tid * PA);
        }
        __syncthreads();
        const T *Sp = S + (size_t)rp * ld + colg;
        const uint8_t *Ap = Asym + (size_t)rp * ld + colg;
        if (half == 0) gk_ef_piece<T, D, 0>(Sp, Ap, (size_t)ld, mrows, phi, rg, probit != 0, rp, colg, -1, -1, acc);
        else gk_ef_piece<T, D, 1>(Sp, Ap, (size_t)ld, mrows, phi, rg, probit != 0, rp, colg, -1, -1, acc);
    }
    gk_ef_fold<ACC, PH>(acc, phi, tid);
    if (rg == 0) {
        const int p0 = half * PH, np = half == 0 ? PH : PA - PH;
        ACC *out = partial + ((size_t)blockIdx.y * npad + colg) * PA + p0;
#pragma unroll
        for (int p = 0; p < PH; ++p)
            if (p < np) { out[p] = acc[0][p]; out[PA + p] = acc[1][p]; }
    }
}

// G[n][p] = sum over chunks of partial[chunk][n][p]  (fixed order)
template <typename ACC>
__global__ void k_eig_greduce(const ACC *__restrict__ partial, int nchunks, size_t per_chunk, double *__restrict__ G) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < per_chunk; i += (size_t)gridDim.x * blockDim.x) {
        double v = 0.0;
        for (int q = 0; q < nchunks; ++q) v += (double)partial[(size_t)q * per_chunk + i];
        G[i] = v;
    }
}

// ---------------------------------------------------------------------------------------------------
// k_eig_push: node block [n0, n0+nb) has just been redrawn; dphi[i][p] = Phi(new f_i) - Phi(old f_i).
// Every column n outside the block gets G[n] += sum_i X[n0+i, n] dphi[i]; columns j inside the block only take
// rows m > j (the in-block pairs with m < j were applied by k_eig_seq3 at the moment node m was drawn).
// ---------------------------------------------------------------------------------------------------
template <typename T, int D>
__global__ void __launch_bounds__(GK_EF_THREADS, 2) k_eig_push(const T *__restrict__ S, const uint8_t *__restrict__ Asym, int ld,
                                                                int N, const double *__restrict__ dphi, int n0, int nb,
                                                                double *__restrict__ G, int probit) {
    typedef typename gk_eacc<T>::type ACC;
    constexpr int PA = gk_edim<D>::PA, PH = gk_edim<D>::PH;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ACC *phi = (ACC *)smem_raw;                                   // [nb <= 128][PA]; then fold scratch
    const int tid = threadIdx.x, cp = tid & 63, half = (tid >> 6) & 1, rg = tid >> 7;
    const int c0 = blockIdx.x * GK_EF_COLS, colg = c0 + 2 * cp;
    for (int t = tid; t < nb * PA; t += GK_EF_THREADS) phi[t] = (ACC)dphi[t];
    __syncthreads();
    ACC acc[2][PH];
#pragma unroll
    for (int e = 0; e < 2; ++e)
#pragma unroll
        for (int p = 0; p < PH; ++p) acc[e][p] = ACC(0);
    const bool inb0 = colg >= n0 && colg < n0 + nb, inb1 = colg + 1 >= n0 && colg + 1 < n0 + nb;
    const int mlow0 = inb0 ? colg : -1, mlow1 = inb1 ? colg + 1 : -1;
    const T *Sp = S + (size_t)n0 * ld + colg;
    const uint8_t *Ap = Asym + (size_t)n0 * ld + colg;
    if (half == 0) gk_ef_piece<T, D, 0>(Sp, Ap, (size_t)ld, nb, phi, rg, probit != 0, n0, colg, mlow0, mlow1, acc);
    else gk_ef_piece<T, D, 1>(Sp, Ap, (size_t)ld, nb, phi, rg, probit != 0, n0, colg, mlow0, mlow1, acc);
    gk_ef_fold<ACC, PH>(acc, phi, tid);
    if (rg == 0) {
        const int p0 = half * PH, np = half == 0 ? PH : PA - PH;
        double *out = G + (size_t)colg * PA + p0;
#pragma unroll
        for (int p = 0; p < PH; ++p)
            if (p < np) { out[p] += (double)acc[0][p]; out[PA + p] += (double)acc[1][p]; }
    }
}

// ---------------------------------------------------------------------------------------------------
// k_eig_seq3: the serial draws of one node block from the running G (one CTA).
// Warp 0 = solver: conditional of node t from the completed row t, factorisation and draw, then completes row t+1
// itself (nodes t and t-1).  Warps 1..15 = updaters: while the solver works on node t they add node t-1's change
// to rows >= t+2.  One barrier per node.  On exit the block's rows of G (with every in-block pair s < j applied) and
// dphi[i] = Phi(new f_i) - Phi(old f_i) go back to global memory for k_eig_push.
// ---------------------------------------------------------------------------------------------------
struct gk_eig_seq3_params {
    int n0, nb, N;
    const void *Sblk; int ldS;           // S[n0+i][n0+j] = Sblk[i*ldS + j]
    const uint8_t *Ablk; int ldA;
    double *F; const double *z; const double *lmbda;
    double mu0, inv_sigma_F;
    double *G;                           // [npad][PA]
    double *dphi;                        // [NB][PA]
    double *glob;                        // sumF at GK_EIG_GLOB_SUMF
    double *out_prec, *out_h;
    int probit;
    long long *prof;                     // optional cycle counters (debug): [0..5] solver phases, [6] updater loop, [7] barrier wait of the solver
};

template <typename T, int D>
size_t gk_eig_seq3_smem() {
    constexpr int NB = gk_eig_nb<T>::NB, PA = gk_edim<D>::PA;
    return (size_t)NB * PA * 8 + (size_t)NB * (NB + 1) * sizeof(T) + (size_t)NB * (NB + 4) + (size_t)2 * NB * D * 8 + 64;
}

template <typename T, typename CH, int D>
__global__ void __launch_bounds__(GK_EIG_SEQ_THREADS, 1) k_eig_seq3(gk_eig_seq3_params P) {
    constexpr int NB = gk_eig_nb<T>::NB, Q = gk_edim<D>::Q, Pn = gk_edim<D>::P, PA = gk_edim<D>::PA;
    constexpr int NUPD = GK_EIG_SEQ_THREADS - 32, RUP = NUPD / PA;       // updater rows in flight
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *Gs = (double *)smem_raw;                      // [NB][PA]
    double *zb = Gs + (size_t)NB * PA;                    // [NB][D]
    double *fold = zb + (size_t)NB * D;                   // [NB][D]
    T *Sb = (T *)(fold + (size_t)NB * D);                 // [NB][NB+1]
    uint8_t *Ab = (uint8_t *)(Sb + (size_t)NB * (NB + 1));   // [NB][NB+4]
    __shared__ double Lam[gk_edim<D>::Q], hv[D], fnew[D], lam[D];
    __shared__ double dph[2][gk_edim<D>::PA];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nb = P.nb, n0 = P.n0;
    const bool probit = P.probit != 0;

    for (int t = tid; t < nb * PA; t += GK_EIG_SEQ_THREADS) Gs[t] = P.G[(size_t)n0 * PA + t];
    for (int t = tid; t < nb * D; t += GK_EIG_SEQ_THREADS) {
        zb[t] = P.z[(size_t)n0 * D + t];
        fold[t] = P.F[(size_t)n0 * D + t];
    }
    for (int t = tid; t < nb * nb; t += GK_EIG_SEQ_THREADS) {
        int i = t / nb, j = t - i * nb;
        if (j > i) {
            Sb[i * (NB + 1) + j] = ((const T *)P.Sblk)[(size_t)i * P.ldS + j];
            if (!probit) Ab[i * (NB + 4) + j] = P.Ablk[(size_t)i * P.ldA + j];
        }
    }
    auto xS = [&](int i, int j) -> double { return probit ? 2.0 : (double)Sb[i * (NB + 1) + j]; };      // i < j only
    auto xA = [&](int i, int j) -> double { return probit ? (double)Sb[i * (NB + 1) + j] : (double)Ab[i * (NB + 4) + j]; };
    if (tid < D) lam[tid] = P.lmbda[tid];
    const int p0 = lane, p1 = lane + 32;
    int a0 = 0, b0 = 0, a1 = 0, b1 = 0;
    if (p0 < PA) gk_col_ab<D>(p0, a0, b0);
    if (p1 < PA) gk_col_ab<D>(p1, a1, b1);
    const int u = tid - 32, up = u >= 0 ? u % PA : 0, ur = u >= 0 ? u / PA : RUP;    // updater: column, row lane
    double sumF = 0.0;
    if (warp == 0 && lane < D) sumF = P.glob[GK_EIG_GLOB_SUMF + lane];
    __syncthreads();

    long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, c0 = 0, c1 = 0;
    const bool prof = P.prof != nullptr;
    for (int t = 0; t <= nb; ++t) {
        if (warp == 0) {
            if (t < nb) {
                if (prof) c0 = clock64();
                const int n = n0 + t, cur = t & 1;
                const double sF = __shfl_sync(0xffffffffu, sumF, (p0 >= Q && p0 < Pn) ? a0 : ((p1 >= Q && p1 < Pn) ? a1 : 0));
                double phio0 = 0.0, phio1 = 0.0;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int p = h ? p1 : p0, pa = h ? a1 : a0, pb = h ? b1 : b0;
                    if (p < PA) {
                        const double g = Gs[(size_t)t * PA + p];
                        double po;
                        if (p < Q) {
                            const double v = (pa == pb ? P.inv_sigma_F : 0.0) + lam[pa] * lam[pb] * g;
                            Lam[p] = v;
                            po = fold[t * D + pa] * fold[t * D + pb];
                            if (P.out_prec) { P.out_prec[((size_t)n * D + pa) * D + pb] = v; P.out_prec[((size_t)n * D + pb) * D + pa] = v; }
                        } else if (p < Pn) {
                            const double gh = Gs[(size_t)t * PA + Pn + pa];
                            const double v = lam[pa] * (gh - (probit ? 0.0 : (sF - fold[t * D + pa])) - P.mu0 * g);
                            hv[pa] = v;
                            po = fold[t * D + pa];
                            if (P.out_h) P.out_h[(size_t)n * D + pa] = v;
                        } else {
                            po = fold[t * D + pa];
                        }
                        if (h) phio1 = po; else phio0 = po;
                    }
                }
                __syncwarp();
                if (prof) { c1 = clock64(); pc[0] += c1 - c0; c0 = c1; }
                if (lane == 0) gk_chol_draw<CH, D>(Lam, hv, zb + t * D, fnew);
                __syncwarp();
                if (prof) { c1 = clock64(); pc[1] += c1 - c0; c0 = c1; }
                if (lane < D) { const double f = fnew[lane]; P.F[(size_t)n * D + lane] = f; sumF += f - fold[t * D + lane]; }
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int p = h ? p1 : p0, pa = h ? a1 : a0, pb = h ? b1 : b0;
                    if (p < PA) {
                        const double d = (p < Q ? fnew[pa] * fnew[pb] : fnew[pa]) - (h ? phio1 : phio0);
                        dph[cur][p] = d;
                        P.dphi[(size_t)t * PA + p] = d;
                    }
                }
                if (prof) { c1 = clock64(); pc[2] += c1 - c0; c0 = c1; }
                // complete row t+1: node t (just drawn) and node t-1 (the updaters start at row t+2)
                if (t + 1 < nb) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int p = h ? p1 : p0;
                        if (p < PA) {
                            double x = p < Pn ? xS(t, t + 1) : xA(t, t + 1);
                            double g = x * dph[cur][p];
                            if (t > 0) {
                                x = p < Pn ? xS(t - 1, t + 1) : xA(t - 1, t + 1);
                                g += x * dph[cur ^ 1][p];
                            }
                            Gs[(size_t)(t + 1) * PA + p] += g;
                        }
                    }
                }
                if (prof) { c1 = clock64(); pc[3] += c1 - c0; c0 = c1; }
            }
        } else if (t > 0 && ur < RUP) {
            if (prof) c0 = clock64();
            const int s = t - 1;
            const double d = dph[s & 1][up];
            if (up < Pn) { for (int j = t + 2 + ur; j < nb; j += RUP) Gs[(size_t)j * PA + up] += xS(s, j) * d; }
            else { for (int j = t + 2 + ur; j < nb; j += RUP) Gs[(size_t)j * PA + up] += xA(s, j) * d; }
            if (prof) { c1 = clock64(); pc[6] += c1 - c0; }
        }
        if (prof) c0 = clock64();
        __syncthreads();
        if (prof) { c1 = clock64(); pc[7] += c1 - c0; }
    }
    if (prof && (tid == 0 || tid == 32)) for (int k = 0; k < 8; ++k) atomicAdd((unsigned long long *)P.prof + (tid ? 8 : 0) + k, (unsigned long long)pc[k]);
    if (warp == 0 && lane < D) P.glob[GK_EIG_GLOB_SUMF + lane] = sumF;
    for (int t = tid; t < nb * PA; t += GK_EIG_SEQ_THREADS) P.G[(size_t)n0 * PA + t] = Gs[t];
}

// ---------------------------------------------------------------------------------------------------
// k_eig_gstats: pair statistics of eigenmodel.py:910-965 from the all-new G (one CTA per node range, fixed order):
//   Pq[ab] = 1/2 sum_n f_na f_nb G[n][ab],  hS[a] = 1/2 sum_n f_na G[n][Q+a],
//   hK[a]  = 1/2 sum_n f_na (G[n][P+a] - (sumF[a] - f_na))        (probit: no sumF term)
// part: [gridDim.x][PA]
// ---------------------------------------------------------------------------------------------------
template <int D>
__global__ void k_eig_gstats(const double *__restrict__ G, const double *__restrict__ F, int N, const double *__restrict__ glob,
                             int probit, double *__restrict__ part) {
    constexpr int Q = gk_edim<D>::Q, Pn = gk_edim<D>::P, PA = gk_edim<D>::PA;
    __shared__ double red[8][gk_edim<D>::PA];
    const int tid = threadIdx.x, p = tid % 64, g = tid / 64;              // blockDim.x == 512: 8 node lanes x 64 columns
    const int per = (N + gridDim.x - 1) / gridDim.x, lo = blockIdx.x * per, hi = min(N, lo + per);
    double s = 0.0;
    if (p < PA) {
        int pa, pb; gk_col_ab<D>(p, pa, pb);
        const double sF = glob[GK_EIG_GLOB_SUMF + pa];
        for (int n = lo + g; n < hi; n += 8) {
            const double fa = F[(size_t)n * D + pa], gv = G[(size_t)n * PA + p];
            if (p < Q) s += fa * F[(size_t)n * D + pb] * gv;
            else if (p < Pn) s += fa * gv;
            else s += fa * (gv - (probit ? 0.0 : sF - fa));
        }
        red[g][p] = s;
    }
    __syncthreads();
    if (tid < PA) { double t = 0.0; for (int k = 0; k < 8; ++k) t += red[k][tid]; part[(size_t)blockIdx.x * PA + tid] = 0.5 * t; }
}

template <int D>
__global__ void k_eig_gstats_final(const double *__restrict__ part, int nparts, double *__restrict__ glob) {
    constexpr int Q = gk_edim<D>::Q, Pn = gk_edim<D>::P, PA = gk_edim<D>::PA;
    const int p = threadIdx.x;
    if (p >= PA) return;
    double t = 0.0;
    for (int k = 0; k < nparts; ++k) t += part[(size_t)k * PA + p];
    int pa, pb; gk_col_ab<D>(p, pa, pb);
    if (p < Q) glob[GK_EIG_GLOB_PQ + p] = t;
    else if (p < Pn) glob[GK_EIG_GLOB_HS + pa] = t;
    else glob[GK_EIG_GLOB_HK + pa] = t;
}
