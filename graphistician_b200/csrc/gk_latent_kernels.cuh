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
// shard of the pairs.  A CTA takes work units (block-row, GK_LAT_KCH consecutive k) round-robin and walks a unit in
// half tiles of 128 x 64; a thread owns an 8 x 4 register block: G_i lives in registers for the whole unit, G_j is
// reduced through (double-buffered) shared memory once per half tile and leaves as 64*DIM fp64 atomics.  fp32 mode
// works in coordinates scaled by sqrt(log2 e), so the logistic is ex2 + rcp (2 SFU ops per pair) with no multiply.
#pragma once
#include "gk_internal.cuh"

#define GK_LAT_TILE 128
#define GK_LAT_THREADS 256
#define GK_LAT_KCH 4
#define GK_LAT_HALF 64            // columns per inner step
#define GK_LAT_MAXDIM 3
#define GK_LAT_SCALE_F32 1.2011224087864498   // sqrt(log2 e): fp32 shadows of L hold L * this

// accumulator layout of one pass: acc[0 .. N*DIM) = G, then the scalars
#define GK_LAT_ACC_SW 0
#define GK_LAT_ACC_SV 1
#define GK_LAT_ACC_EXTRA 4

template <typename T> struct gk_lat_math;
template <> struct gk_lat_math<float> {
    // coordinates are pre-multiplied by sqrt(log2 e): d2s = log2(e) ||l_i - l_j||^2, mu0s = log2(e) mu_0
    static __device__ __forceinline__ float scale() { return 1.2011224087864498f; }        // sqrt(log2 e)
    static __device__ __forceinline__ float scale2() { return 1.4426950408889634f; }       // log2 e
    // gradient passes: s = 1 / (1 + 2^(d2s - mu0s)), MUFU ex2 + rcp
    static __device__ __forceinline__ float sigmoid_scaled(float t) {
        float e, r;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t));
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
        return r;
    }
    // value passes: t = -x log2(e).  e = 2^-|t| and r = 1/(1+e) on the SFU; log1p(e) by lg2 where 1 + e keeps its
    // bits and by its series where it does not (e < 2^-6), so small terms keep their RELATIVE accuracy and the
    // 1.3e8-term sum has no systematic bias.  3 SFU ops per pair.
    static __device__ __forceinline__ void sig_softplus_scaled(float t, float &x, float &s, float &sp) {
        float e, r, lg;
        const float den_arg = -fabsf(t);
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(den_arg));
        const float den = 1.0f + e;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(den));
        asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(den));
        const float series = e * fmaf(e, fmaf(e, 0.33333333f, -0.5f), 1.0f);
        const float l1p = e < 0.015625f ? series : lg * 0.6931471805599453f;
        x = t * -0.6931471805599453f;
        s = t <= 0.0f ? r : e * r;
        sp = fmaxf(x, 0.0f) + l1p;
    }
    // byte c of word -> float through the 2^23 magic number (one PRMT + one FADD, no I2F)
    static __device__ __forceinline__ float byte_to(uint32_t word, int c, uint32_t magic) {
        return __uint_as_float(__byte_perm(word, magic, 0x7650u + c)) - 8388608.0f;
    }
};
template <> struct gk_lat_math<double> {
    static __device__ __forceinline__ double scale() { return 1.0; }
    static __device__ __forceinline__ double scale2() { return 1.0; }
    static __device__ __forceinline__ double sigmoid_scaled(double t) { return 1.0 / (1.0 + exp(t)); }
    static __device__ __forceinline__ void sig_softplus_scaled(double t, double &x, double &s, double &sp) {
        x = -t;
        const double e = exp(-fabs(x)), r = 1.0 / (1.0 + e);
        s = x >= 0.0 ? r : e * r;
        sp = fmax(x, 0.0) + log1p(e);
    }
    static __device__ __forceinline__ double byte_to(uint32_t word, int c, uint32_t) { return (double)((word >> (8 * c)) & 0xffu); }
};

struct gk_lat_pass_params {
    const uint8_t *Asym;   // nrows x ld (this rank's rows), zero diagonal
    int ld, row0, nrows, N;
    const void *Lq;        // NT*128 x DIM of the compute type, zero padded, PRE-SCALED by gk_lat_math<T>::scale(): the evaluation point
    const double *mu0;     // device scalar: the evaluation point's mu_0
    double *acc;           // N*DIM + GK_LAT_ACC_EXTRA doubles, zeroed by the caller
    int NT, nkc, nunits;   // tiles per side, k-chunks per block-row, work units of this rank
    uint32_t magic;        // 0x4B000000, passed as data so that it stays in a register (PRMT selector immediate)
};

// one 8 x 4 register block.  SPECIAL blocks (diagonal tile, ragged last block-row/column) mask per pair.
//   WANT_W: the pass needs SW = sum w (mu_0 gradient);  VALUE: it needs SV;  GRAD: it needs G.
template <typename T, int DIM, bool VALUE, bool GRAD, bool WANT_W, bool SPECIAL>
__device__ __forceinline__ void gk_lat_block(const uint32_t (&a)[8], const T (&li)[8][DIM], const T (&lj)[4][DIM], T mu0s,
                                             T (&gi)[8][DIM], T (&gj)[4][DIM], T &tw, T &tv, int i0, int j0, int N,
                                             bool diag, uint32_t magic) {
    typedef gk_lat_math<T> M;
    constexpr bool AF = VALUE || GRAD || SPECIAL;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            T dd[DIM], d2 = T(0);
#pragma unroll
            for (int d = 0; d < DIM; ++d) { dd[d] = li[r][d] - lj[c][d]; d2 = fma(dd[d], dd[d], d2); }
            T s, v = T(0), w;
            if (VALUE) {
                T sp, x;
                M::sig_softplus_scaled(d2 - mu0s, x, s, sp);
                const T af = M::byte_to(a[r], c, magic);
                v = fma(af, x, T(-2) * sp);
                w = fma(T(-2), s, af);
            } else {
                s = M::sigmoid_scaled(d2 - mu0s);
                // without G only sum w is wanted: its edge part is added per word below (dp4a), not per pair
                w = AF ? fma(T(-2), s, M::byte_to(a[r], c, magic)) : T(-2) * s;
            }
            if (SPECIAL) {
                const int i = i0 + r, j = j0 + c;
                const bool ok = i < N && j < N && (!diag || j > i);
                w = ok ? w : T(0); v = ok ? v : T(0);
            }
            if (WANT_W) tw += w;
            if (VALUE) tv += v;
            if (GRAD) {
#pragma unroll
                for (int d = 0; d < DIM; ++d) { gi[r][d] = fma(w, dd[d], gi[r][d]); gj[c][d] = fma(w, dd[d], gj[c][d]); }
            }
        }
    }
    if (!AF && WANT_W) {
        int cnt = 0;
#pragma unroll
        for (int r = 0; r < 8; ++r) cnt = (int)__dp4a((unsigned)a[r], 0x01010101u, (unsigned)cnt);
        tw += (T)cnt;
    }
}

template <typename T> struct __align__(16) gk_vec16 { T v[16 / sizeof(T)]; };

template <typename T, int DIM, bool VALUE, bool GRAD>
__global__ void __launch_bounds__(GK_LAT_THREADS, VALUE ? 2 : 3) k_latent_pair(gk_lat_pass_params P) {
    typedef gk_lat_math<T> M;
    typedef gk_vec16<T> V;
    constexpr int TS = GK_LAT_TILE, HS = GK_LAT_HALF, VE = 16 / (int)sizeof(T);
    constexpr int CW = 4 * DIM;                              // values a thread owns per half tile (4 columns)
    constexpr int RW = HS * DIM;                             // values per row group of the reduction buffer
    constexpr int NBUF = sizeof(T) == 4 ? 2 : 1;             // fp64 (parity mode): one buffer, one more barrier
    constexpr bool WANT_W = VALUE || !GRAD;
    __shared__ V sRedV[GRAD ? NBUF * 16 * RW / VE : 1];      // [buffer][row group ty][col * DIM + d]
    __shared__ double sScal[GK_LAT_THREADS / 32][2];
    T *sRed = (T *)sRedV;
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4, lane = tid & 31, warp = tid >> 5;
    const T *__restrict__ Lq = (const T *)P.Lq;
    const T mu0s = (T)__ldg(P.mu0) * M::scale2();
    const T inv_scale = T(1) / M::scale();
    const int N = P.N, NT = P.NT;
    const uint32_t magic = P.magic;
    double dw = 0.0, dv = 0.0;
    int it = 0;                                              // reduced half tiles so far (buffer / reducer-warp rotation)

    for (int u = blockIdx.x; u < P.nunits; u += gridDim.x) {
        const int Iloc = u / P.nkc, kc = u - Iloc * P.nkc;
        const int I = P.row0 / TS + Iloc;
        int kmax = NT / 2;                                     // inclusive
        if ((NT & 1) == 0 && I >= NT / 2) kmax -= 1;           // the antipodal tile belongs to the lower block-row
        const int k0 = kc * GK_LAT_KCH, k1 = min(kmax, k0 + GK_LAT_KCH - 1);
        if (k0 > k1) continue;                                 // uniform over the CTA

        const int i0 = I * TS + ty * 8;
        T li[8][DIM], gi[8][DIM];
        uint32_t roff[8];                                      // byte offset of row r (clamped: ragged rows are masked)
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const int il = min(i0 + r - P.row0, P.nrows - 1);
            roff[r] = (uint32_t)il * (uint32_t)P.ld + tx * 4;  // nrows * ld < 2^32 (N <= 65536)
#pragma unroll
            for (int d = 0; d < DIM; ++d) { li[r][d] = Lq[(size_t)(i0 + r) * DIM + d]; gi[r][d] = T(0); }
        }
        const int nst = 2 * (k1 - k0 + 1);                     // half tiles of this unit
        auto colbase = [&](int st) { int J = I + k0 + (st >> 1); if (J >= NT) J -= NT; return J * TS + (st & 1) * HS; };
        uint32_t a_cur[8], a_nxt[8];
        {
            const uint8_t *base = P.Asym + colbase(0);
#pragma unroll
            for (int r = 0; r < 8; ++r) a_cur[r] = __ldg((const uint32_t *)(base + roff[r]));
        }
        for (int st = 0; st < nst; ++st) {
            const int jb = colbase(st), k = k0 + (st >> 1);
            if (st + 1 < nst) {
                const uint8_t *base = P.Asym + colbase(st + 1);
#pragma unroll
                for (int r = 0; r < 8; ++r) a_nxt[r] = __ldg((const uint32_t *)(base + roff[r]));
            }
            if (jb < N) {                                      // (a ragged last tile may have an empty second half)
                T lj[4][DIM], gj[4][DIM];
                {
                    const V *src = (const V *)(Lq + (size_t)(jb + tx * 4) * DIM);
                    T tmp[CW];
#pragma unroll
                    for (int e = 0; e < CW / VE; ++e) { V v; *(uint4 *)&v = __ldg((const uint4 *)src + e);
#pragma unroll
                        for (int q = 0; q < VE; ++q) tmp[e * VE + q] = v.v[q]; }
#pragma unroll
                    for (int c = 0; c < 4; ++c)
#pragma unroll
                        for (int d = 0; d < DIM; ++d) { lj[c][d] = tmp[c * DIM + d]; gj[c][d] = T(0); }
                }
                T tw = T(0), tv = T(0);
                const int J = jb / TS;
                const bool special = (k == 0) || ((N % TS) != 0 && (I == NT - 1 || J == NT - 1));
                if (special) gk_lat_block<T, DIM, VALUE, GRAD, WANT_W, true>(a_cur, li, lj, mu0s, gi, gj, tw, tv, i0, jb + tx * 4, N, k == 0, magic);
                else gk_lat_block<T, DIM, VALUE, GRAD, WANT_W, false>(a_cur, li, lj, mu0s, gi, gj, tw, tv, i0, jb + tx * 4, N, false, magic);
                if (WANT_W) dw += (double)tw;
                if (VALUE) dv += (double)tv;
                if (GRAD) {
                    T *red = sRed + (NBUF == 2 ? (it & 1) * (16 * RW) : 0);
                    if (NBUF == 1) __syncthreads();            // previous half tile's reducer is done
                    {
                        V *dst = (V *)(red + ty * RW + tx * CW);
#pragma unroll
                        for (int e = 0; e < CW / VE; ++e) { V v;
#pragma unroll
                            for (int q = 0; q < VE; ++q) { const int f = e * VE + q; v.v[q] = gj[f / DIM][f % DIM]; }
                            dst[e] = v; }
                    }
                    __syncthreads();                           // with two buffers: the only barrier per half tile
                    if (warp == (it & 7)) {                    // one warp (rotating) folds the 16 row groups and flushes
                        for (int g = lane; g < RW / VE; g += 32) {
                            V s = *(const V *)(red + g * VE);
#pragma unroll
                            for (int t = 1; t < 16; ++t) { const V v = *(const V *)(red + t * RW + g * VE);
#pragma unroll
                                for (int q = 0; q < VE; ++q) s.v[q] += v.v[q]; }
#pragma unroll
                            for (int q = 0; q < VE; ++q) {
                                const size_t gidx = (size_t)jb * DIM + g * VE + q;
                                if (gidx < (size_t)N * DIM) atomicAdd(&P.acc[gidx], -(double)(s.v[q] * inv_scale));
                            }
                        }
                    }
                    ++it;
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
                    if (tx == 0 && i0 + r < N) atomicAdd(&P.acc[(size_t)(i0 + r) * DIM + d], (double)(v * inv_scale));
                }
        }
    }
    for (int o = 16; o; o >>= 1) { dw += __shfl_xor_sync(0xffffffffu, dw, o); dv += __shfl_xor_sync(0xffffffffu, dv, o); }
    if (lane == 0) { sScal[warp][0] = dw; sScal[warp][1] = dv; }
    __syncthreads();
    if (tid == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < GK_LAT_THREADS / 32; ++w) { a += sScal[w][0]; b += sScal[w][1]; }
        if (WANT_W) atomicAdd(&P.acc[(size_t)N * DIM + GK_LAT_ACC_SW], a);
        if (VALUE) atomicAdd(&P.acc[(size_t)N * DIM + GK_LAT_ACC_SV], b);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Gaussian-weight sibling (upstream weights.py:613-901, LatentDistanceGaussianWeightDistribution, B = 1):
//   W_ij | A_ij = 1  ~  N(-||l_i - l_j||^2 + b, Sigma)          i != j
//   :697-731  lp(L, b) = -1/2 Lambda sum_{i != j} A_ij r_ij^2 - sum L^2 / (2 eta) - b^2 / 2 ,   r_ij = W_ij + ||l_i - l_j||^2 - b
//   :765-780  its autograd gradient inside hips' HMC (step 0.005, 10 leapfrog steps)
//   :838-873  the conjugate b and Sigma updates need  n = sum A,  sum A r,  sum A r^2  over the same pairs
// dlp/dl_i = -2 G_i - l_i / eta  with  G_i = Lambda sum_j (A_ij r_ij + A_ji r_ji)(l_i - l_j): W is not symmetric, so
// the pass streams row i of (A, W) and row i of (A^T, W^T) -- 2 x 5 B per ordered pair -- and node i's gradient is a
// row-local sum (no atomics; a row shard owns its nodes' gradients).  One warp per row, 8 rows per CTA sharing the
// column tile of L in shared memory.
// ---------------------------------------------------------------------------------------------------------------
#define GK_LATW_THREADS 256
#define GK_LATW_ROWS (GK_LATW_THREADS / 32)
#define GK_LATW_TILE 1024
#define GK_LAT_ACC_N 2

struct gk_latw_params {
    const uint8_t *A, *At;   // nrows x ld
    const void *W, *Wt;      // nrows x ld of the compute type (B = 1)
    int ld, row0, nrows, N;
    const void *Lq;          // padded N x DIM of the compute type, UNSCALED
    const double *b;         // device scalar: the evaluation point's offset b
    double lam;              // 1 / Sigma
    double *acc;             // N*DIM + GK_LAT_ACC_EXTRA doubles, zeroed by the caller
};

template <typename T> __device__ __forceinline__ void gk_load4(const T *p, T (&v)[4]);
template <> __device__ __forceinline__ void gk_load4<float>(const float *p, float (&v)[4]) {
    const float4 t = __ldcs((const float4 *)p); v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
template <> __device__ __forceinline__ void gk_load4<double>(const double *p, double (&v)[4]) {
    const double2 t = __ldcs((const double2 *)p), u = __ldcs((const double2 *)p + 1); v[0] = t.x; v[1] = t.y; v[2] = u.x; v[3] = u.y;
}

template <typename T, int DIM, bool GRAD>
__global__ void __launch_bounds__(GK_LATW_THREADS) k_latw_rows(gk_latw_params P) {
    __shared__ T sL[GK_LATW_TILE * DIM];
    __shared__ double sScal[GK_LATW_ROWS][3];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const T *__restrict__ Lq = (const T *)P.Lq;
    const T *__restrict__ W = (const T *)P.W, *__restrict__ Wt = (const T *)P.Wt;
    const int N = P.N;
    const T b = (T)__ldg(P.b);
    double s_n = 0.0, s_r = 0.0, s_rr = 0.0;
    const int ngroups = (P.nrows + GK_LATW_ROWS - 1) / GK_LATW_ROWS;
    for (int g = blockIdx.x; g < ngroups; g += gridDim.x) {
        const int il = g * GK_LATW_ROWS + warp, i = P.row0 + il;
        const bool live = il < P.nrows;
        T li[DIM], gi[DIM];
#pragma unroll
        for (int d = 0; d < DIM; ++d) { li[d] = live ? Lq[(size_t)i * DIM + d] : T(0); gi[d] = T(0); }
        T tn = T(0), tr = T(0), trr = T(0);
        const size_t rowoff = (size_t)(live ? il : 0) * P.ld;
        for (int j0 = 0; j0 < N; j0 += GK_LATW_TILE) {
            __syncthreads();
            for (int t = tid; t < GK_LATW_TILE * DIM; t += GK_LATW_THREADS) {
                const int j = j0 + t / DIM;
                sL[t] = j < N ? Lq[(size_t)j0 * DIM + t] : T(0);
            }
            __syncthreads();
            if (!live) continue;
            const int jend = min(N, j0 + GK_LATW_TILE);
            for (int j = j0 + lane * 4; j < jend; j += 128) {
                const uint32_t a_out = __ldcs((const uint32_t *)(P.A + rowoff + j));
                T w_out[4];
                gk_load4<T>(W + rowoff + j, w_out);
                uint32_t a_in = 0u;
                T w_in[4] = {T(0), T(0), T(0), T(0)};
                if (GRAD) { a_in = __ldcs((const uint32_t *)(P.At + rowoff + j)); gk_load4<T>(Wt + rowoff + j, w_in); }
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int jj = j + c;
                    const bool ok = jj < N && jj != i;
                    T dd[DIM], d2 = T(0);
#pragma unroll
                    for (int d = 0; d < DIM; ++d) { dd[d] = li[d] - sL[(jj - j0) * DIM + d]; d2 = fma(dd[d], dd[d], d2); }
                    const T x = d2 - b;
                    const T ao = ok && ((a_out >> (8 * c)) & 0xffu) ? T(1) : T(0);
                    const T ro = ao * (w_out[c] + x);
                    tn += ao; tr += ro; trr = fma(ro, ro, trr);
                    if (GRAD) {
                        const T ai = ok && ((a_in >> (8 * c)) & 0xffu) ? T(1) : T(0);
                        const T coef = fma(ai, w_in[c] + x, ro);
#pragma unroll
                        for (int d = 0; d < DIM; ++d) gi[d] = fma(coef, dd[d], gi[d]);
                    }
                }
            }
        }
        if (GRAD) {
#pragma unroll
            for (int d = 0; d < DIM; ++d) {
                double v = (double)gi[d];
                for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0 && live) P.acc[(size_t)i * DIM + d] = P.lam * v;
            }
        }
        s_n += (double)tn; s_r += (double)tr; s_rr += (double)trr;
    }
    for (int o = 16; o; o >>= 1) {
        s_n += __shfl_xor_sync(0xffffffffu, s_n, o); s_r += __shfl_xor_sync(0xffffffffu, s_r, o); s_rr += __shfl_xor_sync(0xffffffffu, s_rr, o);
    }
    if (lane == 0) { sScal[warp][0] = s_n; sScal[warp][1] = s_r; sScal[warp][2] = s_rr; }
    __syncthreads();
    if (tid == 0) {
        double a = 0.0, c = 0.0, e = 0.0;
        for (int w = 0; w < GK_LATW_ROWS; ++w) { a += sScal[w][0]; c += sScal[w][1]; e += sScal[w][2]; }
        atomicAdd(&P.acc[(size_t)N * DIM + GK_LAT_ACC_N], a);
        atomicAdd(&P.acc[(size_t)N * DIM + GK_LAT_ACC_SW], c);
        atomicAdd(&P.acc[(size_t)N * DIM + GK_LAT_ACC_SV], e);
    }
}
