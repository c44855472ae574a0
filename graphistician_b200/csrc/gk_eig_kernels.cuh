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
// K3: S[n,m] = omega[n,m] + omega[m,n], both ~ PG(1, Mu[n,m]) (Mu is symmetric), Philox keyed by the unordered
// pair so any rank can regenerate any entry.  One CTA = one 64 x 64 tile; tiles strictly below the diagonal
// whose mirror image is also local are skipped (written by the mirror tile through shared memory).
// part[tile] = sum of the pair draws of the tile (each unordered pair once) + the diagonal draws PG(1, mu_0).
// ---------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_eig_pg_tile(T *__restrict__ S, int ld, int row0, int nrows, int N, int D,
                                                     const double *__restrict__ F, const double *__restrict__ lmbda,
                                                     double mu0, uint64_t seed, uint32_t sweep,
                                                     double *__restrict__ part) {
    constexpr int TS = GK_PG_TILE;
    __shared__ float fi[TS][GK_EIG_MAXD], fj[TS][GK_EIG_MAXD];
    __shared__ float tile[TS][TS + 1];
    __shared__ double red[8];
    const int tid = threadIdx.x;
    const int I = row0 + blockIdx.y * TS, J = blockIdx.x * TS;          // global node offsets
    const bool mirror_local = J >= row0 && J < row0 + nrows;
    const int pidx = blockIdx.y * gridDim.x + blockIdx.x;
    if (mirror_local && J < I) { if (tid == 0) part[pidx] = 0.0; return; }
    const bool write_mirror = mirror_local && J > I;
    const bool diag_tile = (I == J);
    for (int t = tid; t < TS * GK_EIG_MAXD; t += 256) {
        int r = t / GK_EIG_MAXD, d = t % GK_EIG_MAXD;
        float a = 0.f, b = 0.f;
        if (d < D) {
            if (I + r < N) a = (float)(F[(size_t)(I + r) * D + d] * lmbda[d]);
            if (J + r < N) b = (float)F[(size_t)(J + r) * D + d];
        }
        fi[r][d] = a; fj[r][d] = b;
    }
    __syncthreads();
    const int j = tid & (TS - 1);
    double local = 0.0;
    for (int i = tid / TS; i < TS; i += 256 / TS) {
        const int n = I + i, m = J + j;
        float v = 0.f;
        if (n < N && m < N && n - row0 < nrows) {
            if (n == m) {
                gk_philox g; g.init(seed, (uint32_t)n, (uint32_t)m, sweep);
                local += (double)gk_pg1_draw((float)mu0, g);
            } else if (!diag_tile || m > n) {
                float z = (float)mu0;
#pragma unroll
                for (int d = 0; d < GK_EIG_MAXD; ++d) z += fi[i][d] * fj[j][d];
                gk_philox g; g.init(seed, (uint32_t)min(n, m), (uint32_t)max(n, m), sweep);
                v = gk_pg1_draw(z, g) + gk_pg1_draw(z, g);
                if (write_mirror || diag_tile || n < m) local += (double)v;
            }
        }
        tile[i][j] = v;
    }
    __syncthreads();
    // write the tile (diagonal tile: mirror inside the tile) and, if local, its transpose
    for (int i = tid / TS; i < TS; i += 256 / TS) {
        const int n = I + i, m = J + j;
        if (n - row0 < nrows && n < N && m < ld) {
            float v = tile[i][j];
            if (diag_tile && j < i) v = tile[j][i];
            S[(size_t)(n - row0) * ld + m] = (m < N) ? (T)v : T(0);
        }
        if (write_mirror) {
            const int n2 = J + i, m2 = I + j;                           // transposed tile: row n2, column m2
            if (n2 < N && n2 - row0 < nrows && m2 < ld) S[(size_t)(n2 - row0) * ld + m2] = (m2 < N) ? (T)tile[j][i] : T(0);
        }
    }
    // tile sum (fixed order inside the CTA)
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
                                             double *__restrict__ accp, bool active) {
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
            s[u] = ok ? (ACC)__ldg(Sp + (size_t)(m + u) * ld) : ACC(0);
            a[u] = ok ? (ACC)__ldg(Ap + (size_t)(m + u) * ld) : ACC(0);
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
                                                                   int n0, int nb, int mc, double *__restrict__ acc) {
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
    if (half == 0) gk_slab_body<T, D, 0>(Sp, Ap, (size_t)ld, mrows, phi, accp, active);
    else gk_slab_body<T, D, 1>(Sp, Ap, (size_t)ld, mrows, phi, accp, active);
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

template <typename T, int D>
__global__ void __launch_bounds__(GK_EIG_SEQ_THREADS, 1) k_eig_seq(gk_eig_seq_params P) {
    constexpr int NB = gk_eig_nb<T>::NB, Q = gk_edim<D>::Q, Pn = gk_edim<D>::P, PA = gk_edim<D>::PA;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *GL = (double *)smem_raw;                      // [NB][PA]
    double *GU = GL + (size_t)NB * PA;                    // [NB][PA]
    double *zb = GU + (size_t)NB * PA;                    // [NB][D]
    double *fold = zb + (size_t)NB * D;                   // [NB][D]
    T *Sb = (T *)(fold + (size_t)NB * D);                 // [NB][NB+1]
    uint8_t *Ab = (uint8_t *)(Sb + (size_t)NB * (NB + 1));   // [NB][NB+4]
    __shared__ double Lam[gk_edim<D>::Q], hv[D], phin[gk_edim<D>::PA], phio[gk_edim<D>::PA], lam[D];
    const int tid = threadIdx.x, nb = P.nb, n0 = P.n0;

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
    for (int t = tid; t < nb * nb; t += GK_EIG_SEQ_THREADS) {
        int i = t / nb, j = t - i * nb;
        Sb[i * (NB + 1) + j] = ((const T *)P.Sblk)[(size_t)i * P.ldS + j];
        Ab[i * (NB + 4) + j] = P.Ablk[(size_t)i * P.ldA + j];
    }
    if (tid < D) lam[tid] = P.lmbda[tid];
    // per-thread role for p = tid < PA
    int pa = 0, pb = 0;                                   // monomial (a, b) for p < Q; a for the linear parts
    if (tid < Q) { int a = 0, r = tid; while (r >= D - a) { r -= D - a; ++a; } pa = a; pb = a + r; }
    else if (tid < Pn) pa = pb = tid - Q;
    else if (tid < PA) pa = pb = tid - Pn;
    double run = 0.0, stat = 0.0;                         // sumF[a] (Q <= p < P) or prefixF[a] (p >= P); statistic
    if (tid >= Q && tid < Pn) run = P.glob[GK_EIG_GLOB_SUMF + pa];
    else if (tid >= Pn && tid < PA) run = P.glob[GK_EIG_GLOB_PREF + pa];
    __syncthreads();

    for (int i = 0; i < nb; ++i) {
        const int n = n0 + i;
        // ---- A: conditional of node n
        if (tid < PA) {
            const double g = GL[(size_t)i * PA + tid] + GU[(size_t)i * PA + tid];
            if (tid < Q) {
                const double v = (pa == pb ? P.inv_sigma_F : 0.0) + lam[pa] * lam[pb] * g;
                Lam[tid] = v;
                phio[tid] = fold[i * D + pa] * fold[i * D + pb];
                if (P.out_prec) { P.out_prec[((size_t)n * D + pa) * D + pb] = v; P.out_prec[((size_t)n * D + pb) * D + pa] = v; }
            } else if (tid < Pn) {
                const double gh = GL[(size_t)i * PA + Pn + pa] + GU[(size_t)i * PA + Pn + pa];
                const double v = lam[pa] * (gh - (run - fold[i * D + pa]) - P.mu0 * g);
                hv[pa] = v;
                phio[tid] = fold[i * D + pa];
                if (P.out_h) P.out_h[(size_t)n * D + pa] = v;
            } else {
                phio[tid] = fold[i * D + pa];
            }
        }
        __syncthreads();
        // ---- B: factor and draw (single thread; the chain is latency-bound, not throughput-bound)
        if (tid == 0) {
            double a[Q], rinv[D], y[D];
#pragma unroll
            for (int p = 0; p < Q; ++p) a[p] = Lam[p];    // packed upper (r <= c) == lower by symmetry: a[(c,r)]
#pragma unroll
            for (int jj = 0; jj < D; ++jj) {
                const double r = rsqrt(a[gk_qidx(D, jj, jj)]);
                rinv[jj] = r;
#pragma unroll
                for (int ii = jj + 1; ii < D; ++ii) a[gk_qidx(D, jj, ii)] *= r;          // L[ii][jj]
#pragma unroll
                for (int kk = jj + 1; kk < D; ++kk)
#pragma unroll
                    for (int ii = kk; ii < D; ++ii) a[gk_qidx(D, kk, ii)] -= a[gk_qidx(D, jj, ii)] * a[gk_qidx(D, jj, kk)];
            }
#pragma unroll
            for (int jj = 0; jj < D; ++jj) {               // y = L^-1 h + z
                double v = hv[jj];
#pragma unroll
                for (int kk = 0; kk < jj; ++kk) v -= a[gk_qidx(D, kk, jj)] * y[kk];
                y[jj] = v * rinv[jj];
            }
#pragma unroll
            for (int jj = 0; jj < D; ++jj) y[jj] += zb[i * D + jj];
#pragma unroll
            for (int jj = D - 1; jj >= 0; --jj) {          // f = L^-T y
                double v = y[jj];
#pragma unroll
                for (int kk = jj + 1; kk < D; ++kk) v -= a[gk_qidx(D, jj, kk)] * y[kk];
                y[jj] = v * rinv[jj];
            }
#pragma unroll
            for (int aa = 0; aa < D; ++aa) {
                P.F[(size_t)n * D + aa] = y[aa];
                phin[Q + aa] = y[aa]; phin[Pn + aa] = y[aa];
#pragma unroll
                for (int bb = aa; bb < D; ++bb) phin[gk_qidx(D, aa, bb)] = y[aa] * y[bb];
            }
        }
        __syncthreads();
        // ---- C: statistics of the pairs (m < n, n), then move node i from "upper" to "lower" in later rows
        if (tid < PA) {
            const double gl = GL[(size_t)i * PA + tid];
            if (tid < Pn) stat += phin[tid] * gl;
            else stat += phin[tid] * (gl - run);
            if (tid >= Q && tid < Pn) run += phin[tid] - phio[tid];
            else if (tid >= Pn) run += phin[tid];
        }
        const int rem = (nb - i - 1) * PA;
        for (int t = tid; t < rem; t += GK_EIG_SEQ_THREADS) {
            const int jr = t / PA, p = t - jr * PA, j = i + 1 + jr;
            const double x = p < Pn ? (double)Sb[i * (NB + 1) + j] : (double)Ab[i * (NB + 4) + j];
            GL[(size_t)j * PA + p] += x * phin[p];
            GU[(size_t)j * PA + p] -= x * phio[p];
        }
        __syncthreads();
    }
    if (tid < Q) P.glob[GK_EIG_GLOB_PQ + tid] += stat;
    else if (tid < Pn) { P.glob[GK_EIG_GLOB_HS + pa] += stat; P.glob[GK_EIG_GLOB_SUMF + pa] = run; }
    else if (tid < PA) { P.glob[GK_EIG_GLOB_HK + pa] += stat; P.glob[GK_EIG_GLOB_PREF + pa] = run; }
}
