// gk_latent_kernels.cuh -- K8: fused value + gradient pass of the latent-distance adjacency model.
//
// Arithmetic being replaced (upstream adjacency.py):
//   :209-219  eta_ij = -||l_i - l_j||^2 + mu_0 + mu_self d_ij ,  P = logistic(eta)
//   :244-275  lp = sum_ij A_ij log P_ij + (1 - A_ij) log(1 - P_ij) - sum L^2 / (2 sigma) - mu_0^2/2 - mu_self^2/2
//   :370-421  its autograd gradient inside hips' HMC (11 evaluations per trajectory)
//
// Pair form.  eta and Asym = A + A^T are symmetric, so every unordered pair {i,j}, i != j, is
// visited ONCE and feeds both end points:
//     x = mu_0 - ||l_i - l_j||^2,   s = logistic(x),   w = Asym_ij - 2 s      ( = g_ij + g_ji, g = A - s )
//     G_i += w (l_i - l_j),  G_j -= w (l_i - l_j)           dlp/dl_n = -2 G_n - l_n / sigma
//     SW  += w                                               dlp/dmu_0 = SW + diagonal term - mu_0
//     SV  += Asym_ij x - 2 softplus(x)                       ll = SV + diagonal term
// The diagonal (eta_ii = mu_0 + mu_self, the same for every i) is O(1) closed form on the host side of the ABI.
//
// Tiling: the N x N pair space is cut into 128 x 128 tiles; block-row I owns the tiles (I, (I + k) mod NT) for
// k = 0 .. NT/2 (k = 0: strict upper triangle of the diagonal tile; k = NT/2 with NT even: only I < NT/2), a
// "circular half" that gives every block-row the same amount of work, so a row shard of Asym is also a balanced
// shard of the pairs.  A CTA takes work units (block-row, GK_LAT_KCH consecutive k) round-robin; a thread owns an
// 8 x 8 register block of the tile: G_i lives in registers for the whole unit, G_j is reduced through shared
// memory once per tile and leaves as 128*DIM fp64 atomics.
#pragma once
#include "gk_internal.cuh"

#define GK_LAT_TILE 128
#define GK_LAT_THREADS 256
#define GK_LAT_KCH 4
#define GK_LAT_MAXDIM 3

// accumulator layout of one pass: acc[0 .. N*DIM) = G, then the scalars
#define GK_LAT_ACC_SW 0
#define GK_LAT_ACC_SV 1
#define GK_LAT_ACC_EXTRA 4

template <typename T> struct gk_lat_math;
template <> struct gk_lat_math<float> {
    // gradient passes: MUFU ex2 + rcp (2 SFU ops per pair); value passes: full-precision expf / log1pf
    static __device__ __forceinline__ float sigmoid_from_negx(float negx) { return __fdividef(1.0f, 1.0f + __expf(negx)); }
    static __device__ __forceinline__ void sig_softplus(float x, float &s, float &sp) {
        const float e = expf(-fabsf(x)), r = 1.0f / (1.0f + e);
        s = x >= 0.0f ? r : e * r;
        sp = fmaxf(x, 0.0f) + log1pf(e);
    }
    static __device__ __forceinline__ float byte_to(uint32_t word, int c) {
        return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7650u + c)) - 8388608.0f;
    }
};
template <> struct gk_lat_math<double> {
    static __device__ __forceinline__ double sigmoid_from_negx(double negx) { return 1.0 / (1.0 + exp(negx)); }
    static __device__ __forceinline__ void sig_softplus(double x, double &s, double &sp) {
        const double e = exp(-fabs(x)), r = 1.0 / (1.0 + e);
        s = x >= 0.0 ? r : e * r;
        sp = fmax(x, 0.0) + log1p(e);
    }
    static __device__ __forceinline__ double byte_to(uint32_t word, int c) { return (double)((word >> (8 * c)) & 0xffu); }
};

struct gk_lat_pass_params {
    const uint8_t *Asym;   // nrows x ld (this rank's rows), zero diagonal
    int ld, row0, nrows, N;
    const void *Lq;        // N x DIM of the compute type: the evaluation point
    const double *mu0;     // device scalar: the evaluation point's mu_0
    double *acc;           // N*DIM + GK_LAT_ACC_EXTRA doubles, zeroed by the caller
    int NT, nkc, nunits;   // tiles per side, k-chunks per block-row, work units of this rank
};

// one 8 x 8 register block of a tile.  SPECIAL tiles (diagonal tile, ragged last block-row/column) mask per pair.
template <typename T, int DIM, bool VALUE, bool GRAD, bool SPECIAL>
__device__ __forceinline__ void gk_lat_block(const uint2 (&a)[8], const T (&li)[8][DIM], const T (&lj)[8][DIM], T mu0,
                                             T (&gi)[8][DIM], T (&gj)[8][DIM], T &tw, T &tv, int i0, int j0, int N,
                                             bool diag) {
    typedef gk_lat_math<T> M;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const T af = M::byte_to(c < 4 ? a[r].x : a[r].y, c & 3);
            T dd[DIM], d2 = T(0);
#pragma unroll
            for (int d = 0; d < DIM; ++d) { dd[d] = li[r][d] - lj[c][d]; d2 = fma(dd[d], dd[d], d2); }
            T s, v = T(0);
            if (VALUE) {
                T sp; const T x = mu0 - d2;
                M::sig_softplus(x, s, sp);
                v = fma(af, x, T(-2) * sp);
            } else {
                s = M::sigmoid_from_negx(d2 - mu0);
            }
            T w = fma(T(-2), s, af);
            if (SPECIAL) {
                const int i = i0 + r, j = j0 + c;
                const bool ok = i < N && j < N && (!diag || j > i);
                w = ok ? w : T(0); v = ok ? v : T(0);
            }
            tw += w;
            if (VALUE) tv += v;
            if (GRAD) {
#pragma unroll
                for (int d = 0; d < DIM; ++d) { gi[r][d] = fma(w, dd[d], gi[r][d]); gj[c][d] = fma(w, dd[d], gj[c][d]); }
            }
        }
    }
}

template <typename T, int DIM, bool VALUE, bool GRAD>
__global__ void __launch_bounds__(GK_LAT_THREADS, 2) k_latent_pair(gk_lat_pass_params P) {
    constexpr int TS = GK_LAT_TILE;
    __shared__ T sLj[TS * DIM];
    __shared__ T sRed[GRAD ? 8 * TS * DIM : 1];           // [warp][col * DIM + d]
    __shared__ double sScal[GK_LAT_THREADS / 32][2];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4, lane = tid & 31, warp = tid >> 5;
    const T *__restrict__ Lq = (const T *)P.Lq;
    const T mu0 = (T)__ldg(P.mu0);
    const int N = P.N, NT = P.NT;
    double dw = 0.0, dv = 0.0;

    for (int u = blockIdx.x; u < P.nunits; u += gridDim.x) {
        const int Iloc = u / P.nkc, kc = u - Iloc * P.nkc;
        const int I = P.row0 / TS + Iloc;
        int kmax = NT / 2;                                     // inclusive
        if ((NT & 1) == 0 && I >= NT / 2) kmax -= 1;           // the antipodal tile belongs to the lower block-row
        const int k0 = kc * GK_LAT_KCH, k1 = min(kmax, k0 + GK_LAT_KCH - 1);
        if (k0 > k1) continue;                                 // uniform over the CTA

        const int i0 = I * TS + ty * 8;
        T li[8][DIM], gi[8][DIM];
        const uint8_t *rowp[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const int i = i0 + r;
            const bool ok = i < P.row0 + P.nrows;
            rowp[r] = ok ? P.Asym + (size_t)(i - P.row0) * P.ld + tx * 8 : nullptr;
#pragma unroll
            for (int d = 0; d < DIM; ++d) { li[r][d] = ok ? Lq[(size_t)i * DIM + d] : T(0); gi[r][d] = T(0); }
        }
        uint2 a_cur[8], a_nxt[8];
        {
            int J = I + k0; if (J >= NT) J -= NT;
#pragma unroll
            for (int r = 0; r < 8; ++r) a_cur[r] = rowp[r] ? __ldg((const uint2 *)(rowp[r] + (size_t)J * TS)) : make_uint2(0u, 0u);
        }
        for (int k = k0; k <= k1; ++k) {
            int J = I + k; if (J >= NT) J -= NT;
            if (k < k1) {
                int Jn = I + k + 1; if (Jn >= NT) Jn -= NT;
#pragma unroll
                for (int r = 0; r < 8; ++r) a_nxt[r] = rowp[r] ? __ldg((const uint2 *)(rowp[r] + (size_t)Jn * TS)) : make_uint2(0u, 0u);
            }
            __syncthreads();                                   // previous tile's sLj / sRed readers are done
            for (int idx = tid; idx < TS * DIM; idx += GK_LAT_THREADS) {
                const size_t g = (size_t)J * TS * DIM + idx;
                sLj[idx] = g < (size_t)N * DIM ? Lq[g] : T(0);
            }
            __syncthreads();
            T lj[8][DIM], gj[8][DIM];
#pragma unroll
            for (int c = 0; c < 8; ++c)
#pragma unroll
                for (int d = 0; d < DIM; ++d) { lj[c][d] = sLj[(tx * 8 + c) * DIM + d]; gj[c][d] = T(0); }
            T tw = T(0), tv = T(0);
            const bool special = (k == 0) || ((N % TS) != 0 && (I == NT - 1 || J == NT - 1));
            if (special) gk_lat_block<T, DIM, VALUE, GRAD, true>(a_cur, li, lj, mu0, gi, gj, tw, tv, i0, J * TS + tx * 8, N, k == 0);
            else gk_lat_block<T, DIM, VALUE, GRAD, false>(a_cur, li, lj, mu0, gi, gj, tw, tv, i0, J * TS + tx * 8, N, false);
            dw += (double)tw; dv += (double)tv;
            if (GRAD) {
#pragma unroll
                for (int c = 0; c < 8; ++c)
#pragma unroll
                    for (int d = 0; d < DIM; ++d) {          // the warp's two row groups first, then one slot per warp
                        const T v = gj[c][d] + __shfl_xor_sync(0xffffffffu, gj[c][d], 16);
                        if (lane < 16) sRed[(warp * TS + tx * 8 + c) * DIM + d] = v;
                    }
                __syncthreads();
                for (int idx = tid; idx < TS * DIM; idx += GK_LAT_THREADS) {
                    T s = T(0);
#pragma unroll
                    for (int t = 0; t < 8; ++t) s += sRed[t * TS * DIM + idx];
                    const size_t g = (size_t)J * TS * DIM + idx;
                    if (g < (size_t)N * DIM) atomicAdd(&P.acc[g], -(double)s);
                }
            }
#pragma unroll
            for (int r = 0; r < 8; ++r) a_cur[r] = a_nxt[r];
        }
        if (GRAD) {
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int d = 0; d < DIM; ++d) {
                    T v = gi[r][d];
                    v += __shfl_xor_sync(0xffffffffu, v, 1); v += __shfl_xor_sync(0xffffffffu, v, 2);
                    v += __shfl_xor_sync(0xffffffffu, v, 4); v += __shfl_xor_sync(0xffffffffu, v, 8);
                    if (tx == 0 && i0 + r < N) atomicAdd(&P.acc[(size_t)(i0 + r) * DIM + d], (double)v);
                }
        }
    }
    for (int o = 16; o; o >>= 1) { dw += __shfl_xor_sync(0xffffffffu, dw, o); dv += __shfl_xor_sync(0xffffffffu, dv, o); }
    if (lane == 0) { sScal[warp][0] = dw; sScal[warp][1] = dv; }
    __syncthreads();
    if (tid == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < GK_LAT_THREADS / 32; ++w) { a += sScal[w][0]; b += sScal[w][1]; }
        atomicAdd(&P.acc[(size_t)N * DIM + GK_LAT_ACC_SW], a);
        if (VALUE) atomicAdd(&P.acc[(size_t)N * DIM + GK_LAT_ACC_SV], b);
    }
}
