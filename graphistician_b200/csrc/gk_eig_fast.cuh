// gk_eig_fast.cuh -- "running G" form of the sequential feature sweep (K4 + K5 + K6) for unsharded networks.
//
// eigenmodel.py:876-908 draws node n from a conditional built from  G[n] = sum_{m != n} X[n,m] Phi(f_m)  with the
// features every other node has AT THAT MOMENT (new for m < n, old for m > n).  Instead of re-contracting a slab
// of S per node block (k_eig_slab, kept for the row-sharded block form), this path keeps G for ALL nodes in HBM:
//   k_eig_gfull   once per sweep:  G[n] = sum_m X[m,n] Phi(f_m_old)          (one streaming pass over S and Asym)
//   k_eig_seq4    per node block:  the serial draws; in-block nodes j > s see node s's change immediately
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
template <typename T, int D, int HALF, bool SM = false>
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
            const T2 sv = SM ? *(const T2 *)(Sp + (size_t)mc * ld) : __ldg((const T2 *)(Sp + (size_t)mc * ld));
            const int mg = mg0 + mc;
            const bool k0 = ok && mg > mlow0, k1 = ok && mg > mlow1;
            if (probit) {
                s[u][0] = (k0 && mg != colg) ? ACC(2) : ACC(0);
                s[u][1] = (k1 && mg != colg + 1) ? ACC(2) : ACC(0);
                a[u][0] = k0 ? (ACC)sv.x : ACC(0);
                a[u][1] = k1 ? (ACC)sv.y : ACC(0);
            } else {
                const unsigned av = SM ? *(const unsigned short *)(Ap + (size_t)mc * ld) : __ldg((const unsigned short *)(Ap + (size_t)mc * ld));
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

// ---- asynchronous staging of a [rows x GK_EF_COLS] tile of S (and of Asym) into shared memory (cp.async, 16 bytes) ----
__device__ __forceinline__ void gk_cp16(void *smem, const void *gmem) {
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem));
}
__device__ __forceinline__ void gk_cp_commit() { asm volatile("cp.async.commit_group;"); }
__device__ __forceinline__ void gk_cp_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <typename T>
__device__ __forceinline__ void gk_ef_stage(T *Ssm, uint8_t *Asm, const T *Sg, const uint8_t *Ag, size_t ld, int rows, bool probit,
                                            int tid) {
    constexpr int VS = GK_EF_COLS * sizeof(T) / 16, VA = GK_EF_COLS / 16;      // 16-byte vectors per tile row
    for (int idx = tid; idx < rows * VS; idx += GK_EF_THREADS) {
        const int r = idx / VS, v = idx - r * VS;
        gk_cp16((char *)(Ssm + (size_t)r * GK_EF_COLS) + v * 16, (const char *)(Sg + (size_t)r * ld) + v * 16);
    }
    if (!probit)
        for (int idx = tid; idx < rows * VA; idx += GK_EF_THREADS) {
            const int r = idx / VA, v = idx - r * VA;
            gk_cp16(Asm + (size_t)r * GK_EF_COLS + v * 16, Ag + (size_t)r * ld + v * 16);
        }
}

template <typename T> struct gk_ef_piece_rows { static constexpr int R = sizeof(T) == 4 ? 64 : 32; };

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
                                                                 int N, const double *__restrict__ F, const double *__restrict__ PhiG, int chunk_rows,
                                                                 int npad, typename gk_eacc<T>::type *__restrict__ partial, int probit) {
    typedef typename gk_eacc<T>::type ACC;
    constexpr int PA = gk_edim<D>::PA, PH = gk_edim<D>::PH, PR = gk_ef_piece_rows<T>::R;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    // two stages of a [PR x 128] tile of S and of Asym (cp.async), two stages of Phi rows; the S area doubles as fold scratch
    T *Ssm = (T *)smem_raw;                                       // [2][PR][128]
    uint8_t *Asm = (uint8_t *)(Ssm + (size_t)2 * PR * GK_EF_COLS);   // [2][PR][128]
    ACC *phis = (ACC *)(Asm + (size_t)2 * PR * GK_EF_COLS);       // [2][PR][PA]
    const int tid = threadIdx.x, cp = tid & 63, half = (tid >> 6) & 1, rg = tid >> 7;
    const int c0 = blockIdx.x * GK_EF_COLS, colg = c0 + 2 * cp;
    const int r0 = blockIdx.y * chunk_rows, r1 = min(N, r0 + chunk_rows);
    ACC acc[2][PH];
#pragma unroll
    for (int e = 0; e < 2; ++e)
#pragma unroll
        for (int p = 0; p < PH; ++p) acc[e][p] = ACC(0);
    auto stage = [&](int rp, int st) {
        const int mrows = min(PR, r1 - rp);
        gk_ef_stage<T>(Ssm + (size_t)st * PR * GK_EF_COLS, Asm + (size_t)st * PR * GK_EF_COLS, S + (size_t)rp * ld + c0,
                       Asym + (size_t)rp * ld + c0, (size_t)ld, mrows, probit != 0, tid);
        gk_cp_commit();
        ACC *phi = phis + (size_t)st * PR * PA;
        if (PhiG) {                                               // mean-field form: Phi rows given (E[f f^T] | E[f] | E[f])
            for (int t = tid; t < mrows * PA; t += GK_EF_THREADS) phi[t] = (ACC)PhiG[(size_t)rp * PA + t];
        } else if (tid < mrows) {
            double f[D];
#pragma unroll
            for (int d = 0; d < D; ++d) f[d] = F[(size_t)(rp + tid) * D + d];
            gk_phi_row<ACC, D>(f, phi + (size_t)tid * PA);
        }
    };
    if (r0 < r1) stage(r0, 0);
    int k = 0;
    for (int rp = r0; rp < r1; rp += PR, ++k) {
        const int mrows = min(PR, r1 - rp), st = k & 1;
        gk_cp_wait_all();
        __syncthreads();                                          // piece k landed; everyone is done with piece k-1
        if (rp + PR < r1) stage(rp + PR, st ^ 1);
        const T *Sp = Ssm + (size_t)st * PR * GK_EF_COLS + 2 * cp;
        const uint8_t *Ap = Asm + (size_t)st * PR * GK_EF_COLS + 2 * cp;
        const ACC *phi = phis + (size_t)st * PR * PA;
        if (half == 0) gk_ef_piece<T, D, 0, true>(Sp, Ap, (size_t)GK_EF_COLS, mrows, phi, rg, probit != 0, rp, colg, -1, -1, acc);
        else gk_ef_piece<T, D, 1, true>(Sp, Ap, (size_t)GK_EF_COLS, mrows, phi, rg, probit != 0, rp, colg, -1, -1, acc);
    }
    ACC *phi = (ACC *)smem_raw;                                   // fold scratch over the (now dead) S stages
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
// rows m > j (the in-block pairs with m < j were applied by k_eig_seq4 at the moment node m was drawn).
// ---------------------------------------------------------------------------------------------------
template <typename T, int D>
__global__ void __launch_bounds__(GK_EF_THREADS, 2) k_eig_push(const T *__restrict__ S, const uint8_t *__restrict__ Asym, int ld,
                                                                int N, const typename gk_eacc<T>::type *__restrict__ dphi, int n0,
                                                                int nb, double *__restrict__ G, int probit, int skip_cb) {
    typedef typename gk_eacc<T>::type ACC;
    constexpr int PA = gk_edim<D>::PA, PH = gk_edim<D>::PH, NB = gk_eig_nb<T>::NB;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    if ((int)blockIdx.x == skip_cb) return;                      // that column strip is pushed by k_eig_push_cols
    // the whole [nb x 128] tile of S and Asym is staged at once (one exposure to HBM latency); the S area doubles as fold scratch
    T *Ssm = (T *)smem_raw;                                       // [NB][128]
    uint8_t *Asm = (uint8_t *)(Ssm + (size_t)NB * GK_EF_COLS);    // [NB][128]
    ACC *phis = (ACC *)(Asm + (size_t)NB * GK_EF_COLS);           // [NB][PA]
    const int tid = threadIdx.x, cp = tid & 63, half = (tid >> 6) & 1, rg = tid >> 7;
    const int c0 = blockIdx.x * GK_EF_COLS, colg = c0 + 2 * cp;
    gk_ef_stage<T>(Ssm, Asm, S + (size_t)n0 * ld + c0, Asym + (size_t)n0 * ld + c0, (size_t)ld, nb, probit != 0, tid);
    gk_cp_commit();
    for (int t = tid; t < nb * PA; t += GK_EF_THREADS) phis[t] = dphi[t];
    ACC acc[2][PH];
#pragma unroll
    for (int e = 0; e < 2; ++e)
#pragma unroll
        for (int p = 0; p < PH; ++p) acc[e][p] = ACC(0);
    const bool inb0 = colg >= n0 && colg < n0 + nb, inb1 = colg + 1 >= n0 && colg + 1 < n0 + nb;
    const int mlow0 = inb0 ? colg : -1, mlow1 = inb1 ? colg + 1 : -1;
    gk_cp_wait_all();
    __syncthreads();
    const T *Sp = Ssm + 2 * cp;
    const uint8_t *Ap = Asm + 2 * cp;
    if (half == 0) gk_ef_piece<T, D, 0, true>(Sp, Ap, (size_t)GK_EF_COLS, nb, phis, rg, probit != 0, n0, colg, mlow0, mlow1, acc);
    else gk_ef_piece<T, D, 1, true>(Sp, Ap, (size_t)GK_EF_COLS, nb, phis, rg, probit != 0, n0, colg, mlow0, mlow1, acc);
    ACC *phi = (ACC *)smem_raw;                                   // fold scratch over the (now dead) S tile
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
// k_eig_push_cols: the same update for a narrow column range [c0, c0+ncols) -- the columns of the NEXT node block,
// which the next k_eig_seq4 needs at once, while k_eig_push covers every other strip on a second stream, overlapped
// with that k_eig_seq4 (gk_eig.cu: sweep_F_fast).  One warp per column, lanes over the block's rows, 8 columns per
// CTA so the strip is spread over 16 SMs and finishes in a few microseconds.
// ---------------------------------------------------------------------------------------------------
#define GK_EPC_COLS 8
// recursive halving: after the steps LEN = 64, 32, 16, 8, 4 lane l holds the warp totals of entries 2l and 2l+1
template <typename ACC, int LEN>
__device__ __forceinline__ void gk_fold_step(ACC (&a)[64], int lane) {
    constexpr int H = LEN / 2, O = LEN / 4;
    const bool up = (lane & O) != 0;
#pragma unroll
    for (int p = 0; p < H; ++p) {
        const ACC send = up ? a[p] : a[p + H];
        const ACC keep = up ? a[p + H] : a[p];
        a[p] = keep + __shfl_xor_sync(0xffffffffu, send, O);
    }
}

template <typename T, int D>
__global__ void __launch_bounds__(32 * GK_EPC_COLS) k_eig_push_cols(const T *__restrict__ S, const uint8_t *__restrict__ Asym, int ld,
                                                                     int N, const typename gk_eacc<T>::type *__restrict__ dphi,
                                                                     int n0, int nb, double *__restrict__ G, int probit, int c0,
                                                                     int ncols) {
    typedef typename gk_eacc<T>::type ACC;
    constexpr int Pn = gk_edim<D>::P, PA = gk_edim<D>::PA, NB = gk_eig_nb<T>::NB, R = NB / 32;
    constexpr int VEC = 16 / sizeof(ACC);
    static_assert(PA <= 64, "gk_fold_step holds 64 entries");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // the next k_eig_seq4 may start its prologue now
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int col = c0 + blockIdx.x * GK_EPC_COLS + w;
    const bool live = w + (int)blockIdx.x * GK_EPC_COLS < ncols && col < N;
    const int mlow = (col >= n0 && col < n0 + nb) ? col : -1;
    ACC sv[R], av[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int m = lane + 32 * r, mg = n0 + m;
        const bool keep = live && m < nb && mg > mlow;
        ACC s = ACC(0), a = ACC(0);
        if (keep) {
            const ACC x = (ACC)__ldg(S + (size_t)mg * ld + col);
            if (probit) { s = mg != col ? ACC(2) : ACC(0); a = x; }
            else { s = x; a = (ACC)__ldg(Asym + (size_t)mg * ld + col); }
        }
        sv[r] = s; av[r] = a;
    }
    ACC acc[64];
#pragma unroll
    for (int p = 0; p < 64; ++p) acc[p] = ACC(0);
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const ACC *ph = dphi + (size_t)min(lane + 32 * r, nb - 1) * PA;        // rows >= nb carry zero weights
        if constexpr (PA % VEC == 0 && sizeof(ACC) == 4) {
#pragma unroll
            for (int q = 0; q < PA / 4; ++q) {
                const float4 v = __ldg((const float4 *)ph + q);
                acc[4 * q + 0] += (4 * q + 0 < Pn ? sv[r] : av[r]) * v.x;
                acc[4 * q + 1] += (4 * q + 1 < Pn ? sv[r] : av[r]) * v.y;
                acc[4 * q + 2] += (4 * q + 2 < Pn ? sv[r] : av[r]) * v.z;
                acc[4 * q + 3] += (4 * q + 3 < Pn ? sv[r] : av[r]) * v.w;
            }
        } else if constexpr (PA % VEC == 0 && sizeof(ACC) == 8) {
#pragma unroll
            for (int q = 0; q < PA / 2; ++q) {
                const double2 v = __ldg((const double2 *)ph + q);
                acc[2 * q + 0] += (2 * q + 0 < Pn ? sv[r] : av[r]) * v.x;
                acc[2 * q + 1] += (2 * q + 1 < Pn ? sv[r] : av[r]) * v.y;
            }
        } else {
#pragma unroll
            for (int p = 0; p < PA; ++p) acc[p] += (p < Pn ? sv[r] : av[r]) * __ldg(ph + p);
        }
    }
    gk_fold_step<ACC, 64>(acc, lane);
    gk_fold_step<ACC, 32>(acc, lane);
    gk_fold_step<ACC, 16>(acc, lane);
    gk_fold_step<ACC, 8>(acc, lane);
    gk_fold_step<ACC, 4>(acc, lane);
    if (live) {
        double *out = G + (size_t)col * PA;
        if (2 * lane < PA) out[2 * lane] += (double)acc[0];
        if (2 * lane + 1 < PA) out[2 * lane + 1] += (double)acc[1];
    }
}

#include "gk_eig_seq4.cuh"
#include "gk_eig_seq5.cuh"

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
