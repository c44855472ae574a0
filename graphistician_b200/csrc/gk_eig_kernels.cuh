// gk_eig_kernels.cuh -- device kernels of the eigenmodel path (K3 Polya-gamma draw, K4 slab contraction,
// K5 in-block sequential feature draw with the K6 global statistics folded in).
//
// Arithmetic being replaced (upstream eigenmodel.py):
//   :75-83   Mu = mu_0 + F diag(lmbda) F^T, diagonal forced to mu_0
//   :859-874 Omega[n,m] ~ PG(1, Mu[n,m]) for all N^2 entries
//   :876-908 for n sequentially: prec_n = I/sigma_F + sum_{m!=n} (O[n,m]+O[m,n]) (f_m f_m^T) o (lmbda lmbda^T)
//                                h_n    = sum_{m!=n} [(k[n,m]-O[n,m]mu_0) + (k[m,n]-O[m,n]mu_0)] f_m o lmbda
//   :910-932 mu_0 statistics, :934-965 lmbda statistics (pair sums over n1 < n2)
//
// Contraction form: with S = Omega + Omega^T and Asym = A + A^T (both symmetric, zero
// diagonal), Phi(f) = [f_a f_b (a<=b) | f_a],
//   G[n]  = sum_m S[n,m] Phi(f_m)      (Q + D columns)        H[n] = sum_m Asym[n,m] f_m   (D columns)
//   prec_n[a,b] = d_ab/sigma_F + lmbda_a lmbda_b G[n][ab]
//   h_n[a]      = lmbda_a ( H[n][a] - (sum_m f_m - f_n)[a] - mu_0 G[n][Q+a] )
// Because S is symmetric a row shard of S is also a column shard: every kernel below streams ROWS m of S
// and produces statistics for a block of node COLUMNS, which is what makes the path row-shardable.
#pragma once
#include "gk_internal.cuh"
#include "gk_pg.cuh"

#define GK_EIG_MAXD 8
#define GK_EIG_NBMAX 128          // nodes per sequential block (f64 storage: 64)
#define GK_EIG_SLAB_THREADS 256
#define GK_EIG_SEQ_THREADS 512
#define GK_PG_TILE 64

template <typename T> struct gk_eacc { typedef float type; };
template <> struct gk_eacc<double> { typedef double type; };

template <int D> struct gk_edim {
    static constexpr int Q = D * (D + 1) / 2;     // quadratic monomials f_a f_b, a <= b
    static constexpr int P = Q + D;               // columns multiplied by S
    static constexpr int PA = P + D;              // + columns multiplied by Asym
    static constexpr int PH = (PA + 1) / 2;       // columns per thread-half in the slab kernel
};

// packed index of the (a, b) monomial, a <= b
__host__ __device__ inline int gk_qidx(int D, int a, int b) { return a * D - a * (a - 1) / 2 + (b - a); }

// ---------------------------------------------------------------------------------------------------
// Asym = A + A^T with a zero diagonal (rows of this rank)
// ---------------------------------------------------------------------------------------------------
__global__ void k_eig_asym(const uint8_t *__restrict__ A, const uint8_t *__restrict__ At, uint8_t *__restrict__ Asym,
                           int nrows, int N, int ld, int row0) {
    size_t total = (size_t)nrows * ld;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int r = (int)(i / ld), j = (int)(i - (size_t)r * ld);
        uint8_t v = 0;
        if (j < N && j != r + row0) v = (uint8_t)((A[i] != 0) + (At[i] != 0));
        Asym[i] = v;
    }
}

// ---------------------------------------------------------------------------------------------------
// K3: S[n,m] = omega[n,m] + omega[m,n], both ~ PG(1, Mu[n,m]) (Mu is symmetric).  One CTA = one 64 x 64 tile,
// always GENERATED in its canonical orientation (smaller block index first) by per-lane generators keyed by
// (seed, sweep, canonical tile, thread), so any rank regenerates exactly the same tile; it is then written to
// the local rows directly, transposed, or both.  Tiles below the diagonal whose mirror image is local are
// skipped.  Pass 1 (uniform across the warp): logits and proposal masses into shared memory.  Pass 2: the draws,
// sorted by proposal piece so that a warp runs one code path at a time (see the comment in the kernel).
// part[tile] = sum of the tile's pair draws (each unordered pair counted by exactly one CTA over all ranks)
//              + its diagonal draws PG(1, mu_0).
// ---------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_eig_pg_tile(T *__restrict__ S, int ld, int row0, int nrows, int N, int D,
                                                     const double *__restrict__ F, const double *__restrict__ lmbda,
                                                     double mu0, uint64_t seed, uint32_t sweep,
                                                     double *__restrict__ part) {
    constexpr int TS = GK_PG_TILE, NQ = TS * TS / 256;
    __shared__ float fi[TS][GK_EIG_MAXD], fj[TS][GK_EIG_MAXD];
    __shared__ float tile[TS][TS + 1], massb[TS][TS + 1];
    extern __shared__ unsigned short pglist_dyn[];            // [8 warps][2 * NQ * 32] left-piece draw lists
    __shared__ double red[8];
    const int tid = threadIdx.x;
    const int I = row0 + blockIdx.y * TS, J = blockIdx.x * TS;          // local row block, column block
    const bool mirror_local = J >= row0 && J < row0 + nrows;
    const int pidx = blockIdx.y * gridDim.x + blockIdx.x;
    if (mirror_local && J < I) { if (tid == 0) part[pidx] = 0.0; return; }
    const bool write_mirror = mirror_local && J > I;
    const bool swapped = I > J, diag_tile = (I == J);
    const int cI = swapped ? J : I, cJ = swapped ? I : J;               // canonical tile origin
    for (int t = tid; t < TS * GK_EIG_MAXD; t += 256) {
        int r = t / GK_EIG_MAXD, d = t % GK_EIG_MAXD;
        float a = 0.f, b = 0.f;
        if (d < D) {
            if (cI + r < N) a = (float)(F[(size_t)(cI + r) * D + d] * lmbda[d]);
            if (cJ + r < N) b = (float)F[(size_t)(cJ + r) * D + d];
        }
        fi[r][d] = a; fj[r][d] = b;
    }
    __syncthreads();
    const int j = tid & (TS - 1), i0 = tid / TS;
    // pass 1: logit and proposal mass of every live entry of this lane's queue (mass < 0 marks a dead entry)
#pragma unroll 4
    for (int k = 0; k < NQ; ++k) {
        const int i = i0 + k * (256 / TS), n = cI + i, m = cJ + j;
        float z = 0.f, ms = -1.f;
        if (n < N && m < N && (!diag_tile || m >= n)) {
            z = (float)mu0;
            if (n != m) {
#pragma unroll
                for (int d = 0; d < GK_EIG_MAXD; ++d) z += fi[i][d] * fj[j][d];
            }
            ms = gk_pg_mass(0.5f * fabsf(z));
        }
        tile[i][j] = z; massb[i][j] = ms;
    }
    // pass 2: piece-sorted draws (gk_pg.cuh).  Each entry needs 2 draws (1 on the diagonal).  2a: every lane decides the
    // piece of its entries' draws and finishes the right-piece ones; left-piece draws are filed into this warp's list
    // (small-Z codes from the front, inverse-Gaussian codes from the back).  2b / 2c: the lists are drained densely, item
    // q by lane q % 32 with that lane's generator, so a tile is a pure function of (seed, sweep, canonical tile).
    // massb is reused as the accumulator of the two draws once the masses have been consumed.
    gk_xoshiro g; g.seed(seed, (uint32_t)tid | ((uint32_t)sweep << 8), (uint32_t)cI, (uint32_t)cJ);
    const int lane = tid & 31;
    const unsigned lt = (1u << lane) - 1u;
    unsigned short *lst = pglist_dyn + (tid >> 5) * (2 * NQ * 32);
    int nA = 0, nB = 0;
    for (int k = 0; k < NQ; ++k) {
        const int i = i0 + k * (256 / TS);
        const float mass = massb[i][j], Z = 0.5f * fabsf(tile[i][j]);
        const int ndraw = mass < 0.f ? 0 : ((diag_tile && i == j) ? 1 : 2);
        const bool small = Z < 1.0f / GK_PG_T;
        float accv = 0.f;
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const bool right = g.unif() < mass;             // mass < 0 for dead entries
            const bool left = q < ndraw && !right;
            if (q < ndraw && right) accv += 0.25f * gk_pg1_right(Z, g);
            const unsigned mA = __ballot_sync(0xffffffffu, left && small), mB = __ballot_sync(0xffffffffu, left && !small);
            const unsigned short code = (unsigned short)(i * TS + j);
            if (left && small) lst[nA + __popc(mA & lt)] = code;
            if (left && !small) lst[2 * NQ * 32 - 1 - nB - __popc(mB & lt)] = code;
            nA += __popc(mA); nB += __popc(mB);
        }
        massb[i][j] = accv;
    }
    __syncwarp();
    for (int q = lane; q < nA; q += 32) {
        const int code = lst[q], i = code / TS, jj = code % TS;
        atomicAdd(&massb[i][jj], 0.25f * gk_pg1_left_small(0.5f * fabsf(tile[i][jj]), g));
    }
    for (int q = lane; q < nB; q += 32) {
        const int code = lst[2 * NQ * 32 - 1 - q], i = code / TS, jj = code % TS;
        atomicAdd(&massb[i][jj], 0.25f * gk_pg1_left_ig(0.5f * fabsf(tile[i][jj]), g));
    }
    __syncwarp();
    double local = 0.0;
    for (int k = 0; k < NQ; ++k) {
        const int i = i0 + k * (256 / TS);
        const float accv = massb[i][j];                      // 0 for dead entries
        if (diag_tile && i == j) { local += (double)accv; tile[i][j] = 0.f; }
        else { tile[i][j] = accv; if (!swapped) local += (double)accv; }
    }
    __syncthreads();
    // write: own rows I.. (direct, or transposed when the canonical tile is the mirror image), mirror rows J..
    for (int i = i0; i < TS; i += 256 / TS) {
        const int n = I + i, m = J + j;
        if (n - row0 < nrows && n < N && m < ld) {
            float v;
            if (swapped) v = tile[j][i];
            else v = (diag_tile && j < i) ? tile[j][i] : tile[i][j];
            S[(size_t)(n - row0) * ld + m] = (m < N) ? (T)v : T(0);
        }
        if (write_mirror) {
            const int n2 = J + i, m2 = I + j;
            if (n2 < N && n2 - row0 < nrows && m2 < ld) S[(size_t)(n2 - row0) * ld + m2] = (m2 < N) ? (T)tile[j][i] : T(0);
        }
    }
    for (int o = 16; o; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((tid & 31) == 0) red[tid >> 5] = local;
    __syncthreads();
    if (tid == 0) { double s = 0.0; for (int w = 0; w < 8; ++w) s += red[w]; part[pidx] = s; }
}

// deterministic final sum of the per-tile partials (single CTA)
__global__ void k_sum_partials(const double *__restrict__ part, int n, double *__restrict__ out) {
    __shared__ double red[32];
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s += part[i];
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) { double t = 0.0; for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w]; *out = t; }
}

// column sums of F (N x D) -> out[D]   (single CTA, fixed order)
__global__ void k_eig_colsum(const double *__restrict__ F, int N, int D, double *__restrict__ out) {
    __shared__ double red[32][GK_EIG_MAXD];
    double s[GK_EIG_MAXD];
    for (int d = 0; d < GK_EIG_MAXD; ++d) s[d] = 0.0;
    for (int n = threadIdx.x; n < N; n += blockDim.x)
        for (int d = 0; d < D; ++d) s[d] += F[(size_t)n * D + d];
    for (int d = 0; d < D; ++d) {
        double v = s[d];
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][d] = v;
    }
    __syncthreads();
    if (threadIdx.x < D) { double t = 0.0; for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w][threadIdx.x]; out[threadIdx.x] = t; }
}

// ---------------------------------------------------------------------------------------------------
// K4 (CUDA-core form): slab contraction for the node block [n0, n0+nb).
// CTA q streams local rows [q*mc, (q+1)*mc) of S / Asym restricted to the block's nb columns (each row
// segment is contiguous: coalesced), against Phi(f_m) staged in shared memory, and adds its partial
//     acc[part][i][p] += sum_m X[m][n0+i] * Phi[m][p]       X = S for p < P, Asym for p >= P
// with part = 0 ("lower": rows m < n0, already resampled this sweep) or 1 ("upper": rows m >= n0, old values;
// includes the in-block rows, which K5 then moves over one by one).  mc divides the block stride, so a chunk
// never straddles n0.  Thread (i, half): column i of the block, half of the PA output columns in registers.
// ---------------------------------------------------------------------------------------------------
template <typename T, int D, int HALF>
__device__ __forceinline__ void gk_slab_body(const T *__restrict__ Sp, const uint8_t *__restrict__ Ap, size_t ld,
                                             int mrows, const typename gk_eacc<T>::type *__restrict__ phi,
                                             double *__restrict__ accp, bool active, bool probit, int diag_m) {
    typedef typename gk_eacc<T>::type ACC;
    constexpr int Pn = gk_edim<D>::P, PA = gk_edim<D>::PA, PH = gk_edim<D>::PH;
    constexpr int p0 = HALF * PH, np = HALF == 0 ? PH : PA - PH;
    ACC acc[np > 0 ? np : 1];
#pragma unroll
    for (int p = 0; p < np; ++p) acc[p] = ACC(0);
    for (int m = 0; m < mrows; m += 4) {
        ACC s[4], a[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const bool ok = active && (m + u < mrows);
            const ACC sv = ok ? (ACC)__ldg(Sp + (size_t)(m + u) * ld) : ACC(0);
            if (probit) {                                   // S == 2 off the diagonal; the buffer holds Z + Z^T
                s[u] = (ok && m + u != diag_m) ? ACC(2) : ACC(0);
                a[u] = sv;
            } else {
                s[u] = sv;
                a[u] = ok ? (ACC)__ldg(Ap + (size_t)(m + u) * ld) : ACC(0);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const ACC *ph = phi + (size_t)(m + u) * PA + p0;      // rows beyond mrows are zero-filled in smem
#pragma unroll
            for (int p = 0; p < np; ++p) acc[p] += ((p0 + p) < Pn ? s[u] : a[u]) * ph[p];
        }
    }
    if (active) {
#pragma unroll
        for (int p = 0; p < np; ++p) atomicAdd(accp + p0 + p, (double)acc[p]);
    }
}

template <typename T, int D>
__global__ void __launch_bounds__(GK_EIG_SLAB_THREADS) k_eig_slab(const T *__restrict__ S, const uint8_t *__restrict__ Asym,
                                                                   int ld, int row0, int nrows, const double *__restrict__ F,
                                                                   int n0, int nb, int mc, double *__restrict__ acc, int probit) {
    typedef typename gk_eacc<T>::type ACC;
    constexpr int Q = gk_edim<D>::Q, Pn = gk_edim<D>::P, PA = gk_edim<D>::PA;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ACC *phi = (ACC *)smem_raw;                                   // [mc_pad][PA]
    const int tid = threadIdx.x;
    const int r0 = blockIdx.x * mc, mrows = min(mc, nrows - r0);
    const int mc_pad = (mc + 3) / 4 * 4;
    for (int t = tid; t < mc_pad * PA; t += GK_EIG_SLAB_THREADS) phi[t] = ACC(0);
    __syncthreads();
    for (int m = tid; m < mrows; m += GK_EIG_SLAB_THREADS) {
        double f[D];
#pragma unroll
        for (int d = 0; d < D; ++d) f[d] = F[(size_t)(row0 + r0 + m) * D + d];
        ACC *ph = phi + (size_t)m * PA;
#pragma unroll
        for (int a = 0; a < D; ++a) {
#pragma unroll
            for (int b = a; b < D; ++b) ph[gk_qidx(D, a, b)] = (ACC)(f[a] * f[b]);
            ph[Q + a] = (ACC)f[a];
            ph[Pn + a] = (ACC)f[a];
        }
    }
    __syncthreads();
    const int i = tid & (GK_EIG_NBMAX - 1), half = tid / GK_EIG_NBMAX;
    const bool active = i < nb;
    const int part = (row0 + r0 >= n0) ? 1 : 0;
    const T *Sp = S + (size_t)r0 * ld + n0 + i;
    const uint8_t *Ap = Asym + (size_t)r0 * ld + n0 + i;
    double *accp = acc + ((size_t)part * GK_EIG_NBMAX + i) * PA;
    const int diag_m = n0 + i - (row0 + r0);               // local row at which this column is the diagonal
    if (half == 0) gk_slab_body<T, D, 0>(Sp, Ap, (size_t)ld, mrows, phi, accp, active, probit != 0, diag_m);
    else gk_slab_body<T, D, 1>(Sp, Ap, (size_t)ld, mrows, phi, accp, active, probit != 0, diag_m);
}

// ---------------------------------------------------------------------------------------------------
// K5: the sequential draw of the block's nodes (one CTA).  Exactly the serial Gibbs recursion of
// eigenmodel.py:884-908: node n sees the NEW features of every m < n and the OLD ones of every m > n.
//   GL[j] = contributions of already-resampled nodes (slab "lower" part + in-block nodes moved so far)
//   GU[j] = contributions of not-yet-resampled nodes (slab "upper" part, old values)
// per node i:  A) prec/h from GL+GU      B) one thread: Cholesky, f = L^-T (L^-1 h + z)
//              C) all threads: move node i from GU to GL in every later row j; threads p < PA fold the
//                 pair statistics of K6:  Pq[ab] += f_a f_b GL[i][ab],  hS[a] += f_a GL[i][Q+a],
//                 hK[a] += f_a (GL[i][P+a] - sum_{m<n} f_m[a])          (pairs (m, n), m < n, new values)
// glob layout (doubles): sumF[8] | prefixF[8] | Pq[36] | hS[8] | hK[8]
// ---------------------------------------------------------------------------------------------------
struct gk_eig_seq_params {
    int n0, nb, N;
    const void *Sblk; int ldS;           // S[n0+i][n0+j] = Sblk[i*ldS + j]
    const uint8_t *Ablk; int ldA;
    double *F; const double *z; const double *lmbda;
    double mu0, inv_sigma_F;
    double *acc;                         // [2][GK_EIG_NBMAX][PA]; consumed and zeroed
    double *glob;
    double *out_prec, *out_h;            // optional [N][D][D], [N][D]
    int probit;                          // 1: S == 2 off the diagonal, Sblk holds Z + Z^T (the "A" operand), no kappa offset
};
#define GK_EIG_GLOB_SUMF 0
#define GK_EIG_GLOB_PREF 8
#define GK_EIG_GLOB_PQ 16
#define GK_EIG_GLOB_HS 52
#define GK_EIG_GLOB_HK 60
#define GK_EIG_GLOB_LEN 68

template <typename T> struct gk_eig_nb { static constexpr int NB = sizeof(T) == 4 ? 128 : 64; };

template <typename T, int D>
size_t gk_eig_seq_smem() {
    constexpr int NB = gk_eig_nb<T>::NB, PA = gk_edim<D>::PA;
    return (size_t)2 * NB * PA * 8 + (size_t)NB * (NB + 1) * sizeof(T) + (size_t)NB * (NB + 4) + (size_t)2 * NB * D * 8 + 64;
}

// factor prec = L L^T and draw f = L^-T (L^-1 h + z); forward substitution is fused into the right-looking
// factorisation and the back substitution is column-oriented, so the dependent chain is ~3 ops per column.
template <typename CH, int D>
__device__ __forceinline__ void gk_chol_draw(const double *__restrict__ Lam, const double *__restrict__ hv,
                                             const double *__restrict__ z, double *__restrict__ fout) {
    constexpr int Q = gk_edim<D>::Q;
    CH a[Q], rinv[D], w[D];
#pragma unroll
    for (int p = 0; p < Q; ++p) a[p] = (CH)Lam[p];             // packed (r <= c); L[ii][jj] lives at (jj, ii)
#pragma unroll
    for (int d = 0; d < D; ++d) w[d] = (CH)hv[d];
#pragma unroll
    for (int jj = 0; jj < D; ++jj) {
        CH r;
        if constexpr (sizeof(CH) == 4) r = rsqrtf(a[gk_qidx(D, jj, jj)]); else r = rsqrt(a[gk_qidx(D, jj, jj)]);
        rinv[jj] = r;
        w[jj] *= r;                                            // y[jj]
#pragma unroll
        for (int ii = jj + 1; ii < D; ++ii) {
            a[gk_qidx(D, jj, ii)] *= r;                        // L[ii][jj]
            w[ii] -= a[gk_qidx(D, jj, ii)] * w[jj];
        }
#pragma unroll
        for (int kk = jj + 1; kk < D; ++kk)
#pragma unroll
            for (int ii = kk; ii < D; ++ii) a[gk_qidx(D, kk, ii)] -= a[gk_qidx(D, jj, ii)] * a[gk_qidx(D, jj, kk)];
    }
#pragma unroll
    for (int d = 0; d < D; ++d) w[d] += (CH)z[d];
#pragma unroll
    for (int jj = D - 1; jj >= 0; --jj) {
        w[jj] *= rinv[jj];
#pragma unroll
        for (int kk = 0; kk < jj; ++kk) w[kk] -= a[gk_qidx(D, kk, jj)] * w[jj];
    }
#pragma unroll
    for (int d = 0; d < D; ++d) fout[d] = (double)w[d];
}

// monomial (a, b) of packed column p (p < Q: quadratic; Q <= p < P and P <= p < PA: linear in a)
template <int D>
__device__ __forceinline__ void gk_col_ab(int p, int &pa, int &pb) {
    constexpr int Q = gk_edim<D>::Q, Pn = gk_edim<D>::P;
    if (p < Q) { int a = 0, r = p; while (r >= D - a) { r -= D - a; ++a; } pa = a; pb = a + r; }
    else if (p < Pn) pa = pb = p - Q;
    else pa = pb = p - Pn;
}

// Warp 0 = solver: conditional, factorisation and draw of node t, then completes row t+1 itself (node t and
// node t-1 contributions).  Warps 1..15 = updaters: while the solver works on node t they move node t-1 from
// "upper" to "lower" in rows >= t+2 and fold node t-1's pair statistics.  One barrier per node.
template <typename T, typename CH, int D>
__global__ void __launch_bounds__(GK_EIG_SEQ_THREADS, 1) k_eig_seq(gk_eig_seq_params P) {
    constexpr int NB = gk_eig_nb<T>::NB, Q = gk_edim<D>::Q, Pn = gk_edim<D>::P, PA = gk_edim<D>::PA;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *GL = (double *)smem_raw;                      // [NB][PA]
    double *GU = GL + (size_t)NB * PA;                    // [NB][PA]
    double *zb = GU + (size_t)NB * PA;                    // [NB][D]
    double *fold = zb + (size_t)NB * D;                   // [NB][D]
    T *Sb = (T *)(fold + (size_t)NB * D);                 // [NB][NB+1]
    uint8_t *Ab = (uint8_t *)(Sb + (size_t)NB * (NB + 1));   // [NB][NB+4]
    __shared__ double Lam[gk_edim<D>::Q], hv[D], fnew[D], lam[D];
    __shared__ double phin[2][gk_edim<D>::PA], phio[2][gk_edim<D>::PA];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nb = P.nb, n0 = P.n0;

    for (int t = tid; t < nb * PA; t += GK_EIG_SEQ_THREADS) {
        int i = t / PA, p = t - i * PA;
        GL[t] = P.acc[(size_t)i * PA + p];
        GU[t] = P.acc[((size_t)GK_EIG_NBMAX + i) * PA + p];
        P.acc[(size_t)i * PA + p] = 0.0;
        P.acc[((size_t)GK_EIG_NBMAX + i) * PA + p] = 0.0;
    }
    for (int t = tid; t < nb * D; t += GK_EIG_SEQ_THREADS) {
        zb[t] = P.z[(size_t)n0 * D + t];
        fold[t] = P.F[(size_t)n0 * D + t];
    }
    const bool probit = P.probit != 0;
    for (int t = tid; t < nb * nb; t += GK_EIG_SEQ_THREADS) {
        int i = t / nb, j = t - i * nb;
        Sb[i * (NB + 1) + j] = ((const T *)P.Sblk)[(size_t)i * P.ldS + j];
        if (!probit) Ab[i * (NB + 4) + j] = P.Ablk[(size_t)i * P.ldA + j];
    }
    // operands of the in-block moves: xS multiplies the quadratic/linear S columns, xA the "A" columns
    auto xS = [&](int i, int j) -> double { return probit ? (i != j ? 2.0 : 0.0) : (double)Sb[i * (NB + 1) + j]; };
    auto xA = [&](int i, int j) -> double { return probit ? (double)Sb[i * (NB + 1) + j] : (double)Ab[i * (NB + 4) + j]; };
    if (tid < D) lam[tid] = P.lmbda[tid];
    // solver lanes own columns p0 = lane and p1 = lane + 32; updater threads 32..32+PA-1 own statistic column tid-32
    const int p0 = lane, p1 = lane + 32;
    int a0 = 0, b0 = 0, a1 = 0, b1 = 0, sa = 0, sb = 0;
    if (p0 < PA) gk_col_ab<D>(p0, a0, b0);
    if (p1 < PA) gk_col_ab<D>(p1, a1, b1);
    const int sp = tid - 32;                               // statistic column of an updater thread
    if (sp >= 0 && sp < PA) gk_col_ab<D>(sp, sa, sb);
    double sumF = 0.0, pref = 0.0, stat = 0.0;
    if (warp == 0 && lane < D) sumF = P.glob[GK_EIG_GLOB_SUMF + lane];
    if (sp >= Pn && sp < PA) pref = P.glob[GK_EIG_GLOB_PREF + sa];
    __syncthreads();

    for (int t = 0; t <= nb; ++t) {
        if (warp == 0) {
            if (t < nb) {
                const int n = n0 + t, cur = t & 1;
                // ---- A: conditional of node n from the completed row t
                const double sF = __shfl_sync(0xffffffffu, sumF, (p0 >= Q && p0 < Pn) ? a0 : ((p1 >= Q && p1 < Pn) ? a1 : 0));
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int p = h ? p1 : p0, pa = h ? a1 : a0, pb = h ? b1 : b0;
                    if (p < PA) {
                        const double g = GL[(size_t)t * PA + p] + GU[(size_t)t * PA + p];
                        if (p < Q) {
                            const double v = (pa == pb ? P.inv_sigma_F : 0.0) + lam[pa] * lam[pb] * g;
                            Lam[p] = v;
                            phio[cur][p] = fold[t * D + pa] * fold[t * D + pb];
                            if (P.out_prec) { P.out_prec[((size_t)n * D + pa) * D + pb] = v; P.out_prec[((size_t)n * D + pb) * D + pa] = v; }
                        } else if (p < Pn) {
                            const double gh = GL[(size_t)t * PA + Pn + pa] + GU[(size_t)t * PA + Pn + pa];
                            const double v = lam[pa] * (gh - (probit ? 0.0 : (sF - fold[t * D + pa])) - P.mu0 * g);
                            hv[pa] = v;
                            phio[cur][p] = fold[t * D + pa];
                            if (P.out_h) P.out_h[(size_t)n * D + pa] = v;
                        } else {
                            phio[cur][p] = fold[t * D + pa];
                        }
                    }
                }
                __syncwarp();
                // ---- B: factor and draw (one lane: the chain is latency-bound)
                if (lane == 0) gk_chol_draw<CH, D>(Lam, hv, zb + t * D, fnew);
                __syncwarp();
                if (lane < D) { const double f = fnew[lane]; P.F[(size_t)n * D + lane] = f; sumF += f - fold[t * D + lane]; }
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int p = h ? p1 : p0, pa = h ? a1 : a0, pb = h ? b1 : b0;
                    if (p < PA) phin[cur][p] = p < Q ? fnew[pa] * fnew[pb] : fnew[pa];
                }
                __syncwarp();
                // ---- complete row t+1: node t (just drawn) and node t-1 (the updaters start at row t+2)
                if (t + 1 < nb) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int p = h ? p1 : p0;
                        if (p < PA) {
                            double x = p < Pn ? xS(t, t + 1) : xA(t, t + 1);
                            double gl = x * phin[cur][p], gu = x * phio[cur][p];
                            if (t > 0) {
                                x = p < Pn ? xS(t - 1, t + 1) : xA(t - 1, t + 1);
                                gl += x * phin[cur ^ 1][p]; gu += x * phio[cur ^ 1][p];
                            }
                            GL[(size_t)(t + 1) * PA + p] += gl;
                            GU[(size_t)(t + 1) * PA + p] -= gu;
                        }
                    }
                }
            }
        } else if (t > 0) {
            // ---- updaters: node s = t-1
            const int s = t - 1, buf = s & 1;
            if (sp < PA) {                                 // pair statistics of (m < n_s, n_s) with the new features
                const double gl = GL[(size_t)s * PA + sp];
                if (sp < Pn) stat += phin[buf][sp] * gl;
                else { stat += phin[buf][sp] * (gl - (probit ? 0.0 : pref)); pref += phin[buf][sp]; }
            }
            const int rem = (nb - t - 2) * PA;             // rows t+2 .. nb-1 (row t+1 belongs to the solver)
            for (int e = tid - 32; e < rem; e += GK_EIG_SEQ_THREADS - 32) {
                const int jr = e / PA, p = e - jr * PA, j = t + 2 + jr;
                const double x = p < Pn ? xS(s, j) : xA(s, j);
                GL[(size_t)j * PA + p] += x * phin[buf][p];
                GU[(size_t)j * PA + p] -= x * phio[buf][p];
            }
        }
        __syncthreads();
    }
    if (warp == 0 && lane < D) P.glob[GK_EIG_GLOB_SUMF + lane] = sumF;
    if (sp >= 0 && sp < Q) P.glob[GK_EIG_GLOB_PQ + sp] += stat;
    else if (sp >= Q && sp < Pn) P.glob[GK_EIG_GLOB_HS + sa] += stat;
    else if (sp >= Pn && sp < PA) { P.glob[GK_EIG_GLOB_HK + sa] += stat; P.glob[GK_EIG_GLOB_PREF + sa] = pref; }
}
