// gk_sbm_kernels.cuh -- device kernels of the SBM path (K1 block-pair statistics, K2 assignment sweep).
//
// Strip formulation (DESIGN.md "SBM kernels"): for a block of nodes [n0, n0+nb) the per-node, per-block
// sufficient statistics  stats[n][k'] = sum_{n2 : c[n2]=k'} x(n, n2),  x = (a, a*w_b, a*w_b*w_b'),
// are sums DOWN a column strip of A^T/W^T (out-edges) or A/W (in-edges).  Lanes own nodes (2 adjacent
// columns each, 128-bit W loads), a warp owns one (block k', row-part) unit, so the bin is warp-uniform
// and accumulation stays in registers: no atomics, no shared-memory histogram traffic.
#pragma once
#include "gk_internal.cuh"

#define GK_STRIP_COLS 64      // nodes per CTA strip (2 per lane)
#define GK_STRIP_WARPS 16
#define GK_STRIP_THREADS (GK_STRIP_WARPS * 32)
#define GK_MAX_K 32
#define GK_MAX_CHUNK_ROWS 4096
#define GK_SLOT_PITCH 65

template <int B> struct gk_sdim { static constexpr int S = 1 + B + B * (B + 1) / 2; };

// packed index of the (b, b') second-moment statistic, b <= b'
__host__ __device__ inline int gk_sym_idx(int B, int b, int b2) { return 1 + B + b * B - b * (b - 1) / 2 + (b2 - b); }

template <typename T, int B> struct gk_acc { typedef float type; };
template <int B> struct gk_acc<double, B> { typedef double type; };

// load the 2*B weight values of two adjacent nodes (16-byte vector loads where the type allows)
template <typename T, int B>
__device__ __forceinline__ void gk_load_w2(const T *__restrict__ p, T (&w)[2 * B + (B == 0)]) {
    if constexpr (B == 0) { w[0] = T(0); }
    else if constexpr (sizeof(T) == 4 && B == 1) { float2 v = __ldg((const float2 *)p); w[0] = v.x; w[1] = v.y; }
    else if constexpr (sizeof(T) == 4 && B == 2) { float4 v = __ldg((const float4 *)p); w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w; }
    else if constexpr (sizeof(T) == 8 && B == 1) { double2 v = __ldg((const double2 *)p); w[0] = v.x; w[1] = v.y; }
    else if constexpr (sizeof(T) == 8 && B == 2) { double2 v0 = __ldg((const double2 *)p), v1 = __ldg((const double2 *)p + 1); w[0] = v0.x; w[1] = v0.y; w[2] = v1.x; w[3] = v1.y; }
    else if constexpr (sizeof(T) == 4) {          // B = 3, 4: two nodes = 2 B floats, 8-byte aligned (the node pair starts on an even column)
#pragma unroll
        for (int k = 0; k < B; ++k) { float2 v = __ldg((const float2 *)p + k); w[2 * k] = v.x; w[2 * k + 1] = v.y; }
    } else {
#pragma unroll
        for (int k = 0; k < B; ++k) { double2 v = __ldg((const double2 *)p + k); w[2 * k] = v.x; w[2 * k + 1] = v.y; }
    }
}

struct gk_strip_params {
    const uint8_t *Abase[2];      // [0] = A^T rows (out-edge statistics), [1] = A rows (in-edge statistics)
    const void *Wbase[2];
    int ld, row0, nrows;          // local rows of this rank are global nodes row0 .. row0+nrows-1
    int n0, nb_pad;               // node block start, padded node count (multiple of GK_STRIP_COLS)
    int dir_first;                // first direction handled by blockIdx.z == 0
    const uint8_t *c8;            // N block labels
    int K, chunk_rows, nchunks;
    void *partial;                // ACC [ndirs][nchunks][nb_pad][K][S]
};

template <typename T, int B>
__global__ void __launch_bounds__(GK_STRIP_THREADS, 2) k_strip_stats(gk_strip_params P) {
    constexpr int S = gk_sdim<B>::S;
    typedef typename gk_acc<T, B>::type ACC;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int K = P.K, U = K > GK_STRIP_WARPS ? K : GK_STRIP_WARPS;
    // shared-memory carve-up
    ACC *slots = (ACC *)smem_raw;                                  // [U][S][GK_SLOT_PITCH]
    uint16_t *order = (uint16_t *)(slots + (size_t)U * S * GK_SLOT_PITCH);   // [chunk_rows]
    uint8_t *sc = (uint8_t *)(order + P.chunk_rows);               // [chunk_rows]
    __shared__ int binstart[GK_MAX_K + 1];
    __shared__ int unit_bin[GK_MAX_K], unit_lo[GK_MAX_K], unit_hi[GK_MAX_K], ustart[GK_MAX_K + 1];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int dir = P.dir_first + blockIdx.z, q = blockIdx.y, strip = blockIdx.x;
    const int ra = q * P.chunk_rows;
    const int nr = min(P.chunk_rows, P.nrows - ra);                // rows in this chunk (> 0 by construction)

    // 1. labels of this chunk's rows, per-bin counts
    for (int i = tid; i < nr; i += GK_STRIP_THREADS) sc[i] = P.c8[P.row0 + ra + i];
    if (tid <= GK_MAX_K) binstart[tid] = 0;
    __syncthreads();
    // each warp counts/compacts whole bins (deterministic, ascending row order inside a bin)
    for (int k = warp; k < K; k += GK_STRIP_WARPS) {
        int cnt = 0;
        for (int g = 0; g < nr; g += 32) cnt += __popc(__ballot_sync(0xffffffffu, (g + lane < nr) && sc[g + lane] == k));
        if (lane == 0) binstart[k + 1] = cnt;
    }
    __syncthreads();
    if (tid == 0) { int a = 0; for (int k = 0; k <= K; ++k) { int t = binstart[k]; a += t; binstart[k] = a; } }
    __syncthreads();
    for (int k = warp; k < K; k += GK_STRIP_WARPS) {
        int base = binstart[k], cnt = 0;
        for (int g = 0; g < nr; g += 32) {
            bool m = (g + lane < nr) && sc[g + lane] == k;
            unsigned mask = __ballot_sync(0xffffffffu, m);
            if (m) order[base + cnt + __popc(mask & ((1u << lane) - 1u))] = (uint16_t)(g + lane);
            cnt += __popc(mask);
        }
    }
    // 2. work units: (bin, row-part).  K <= warps: split big bins over the spare warps (greedy on rows/warp)
    if (warp == 0) {
        if (K > GK_STRIP_WARPS) {
            if (lane < K) { unit_bin[lane] = lane; unit_lo[lane] = binstart[lane]; unit_hi[lane] = binstart[lane + 1]; ustart[lane] = lane; }
            if (lane == 0) ustart[K] = K;
        } else {
            int sz = lane < K ? binstart[lane + 1] - binstart[lane] : 0;
            int nk = sz > 0 ? 1 : 0;
            int used = __reduce_add_sync(0xffffffffu, nk);
            for (int it = used; it < GK_STRIP_WARPS; ++it) {
                float ratio = nk > 0 ? (float)sz / (float)nk : -1.f;
                float best = ratio; int bl = lane;
                for (int o = 16; o; o >>= 1) {
                    float ob = __shfl_xor_sync(0xffffffffu, best, o); int ol = __shfl_xor_sync(0xffffffffu, bl, o);
                    if (ob > best || (ob == best && ol < bl)) { best = ob; bl = ol; }
                }
                if (lane == bl) nk++;
            }
            int incl = nk;
            for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
            int st = incl - nk;
            if (lane < K) {
                ustart[lane] = st;
                for (int i = 0; i < nk; ++i) {
                    unit_bin[st + i] = lane;
                    unit_lo[st + i] = binstart[lane] + (int)((long long)sz * i / nk);
                    unit_hi[st + i] = binstart[lane] + (int)((long long)sz * (i + 1) / nk);
                }
            }
            if (lane == K - 1) ustart[K] = st + nk;
            int tot = __shfl_sync(0xffffffffu, incl, 31);
            if (lane >= tot && lane < GK_STRIP_WARPS) { unit_bin[lane] = 0; unit_lo[lane] = 0; unit_hi[lane] = 0; }
        }
    }
    __syncthreads();

    // 3. stream the unit's rows; lane owns nodes (2*lane, 2*lane+1) of the strip
    const int col = strip * GK_STRIP_COLS + 2 * lane;              // node offset inside the block
    const int node0 = P.n0 + col;                                  // global node id of the first column
    const uint8_t *Ab = (dir ? P.Abase[1] : P.Abase[0]) + node0;
    const T *Wb = (const T *)(dir ? P.Wbase[1] : P.Wbase[0]) + (size_t)node0 * B;
    const size_t ld = P.ld;
    for (int u = warp; u < U; u += GK_STRIP_WARPS) {
        const int lo = unit_lo[u], hi = unit_hi[u];
        ACC acc[2][S];
#pragma unroll
        for (int e = 0; e < 2; ++e)
#pragma unroll
            for (int s = 0; s < S; ++s) acc[e][s] = ACC(0);
        for (int pos = lo; pos < hi; pos += 4) {
            int r[4]; unsigned a2[4]; T w[4][2 * B + (B == 0)];
#pragma unroll
            for (int j = 0; j < 4; ++j) r[j] = ra + order[min(pos + j, hi - 1)];
#pragma unroll
            for (int j = 0; j < 4; ++j) a2[j] = __ldg((const unsigned short *)(Ab + (size_t)r[j] * ld));
            if constexpr (B > 0) {
#pragma unroll
                for (int j = 0; j < 4; ++j) gk_load_w2<T, B>(Wb + (size_t)r[j] * ld * B, w[j]);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const bool valid = pos + j < hi;
                const int g = P.row0 + r[j];                       // global id of the "other" node
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const bool edge = valid && ((a2[j] >> (8 * e)) & 0xffu) != 0u && g != node0 + e;
                    acc[e][0] += edge ? ACC(1) : ACC(0);
                    if constexpr (B > 0) {
                        ACC wv[B];
#pragma unroll
                        for (int b = 0; b < B; ++b) { wv[b] = edge ? (ACC)w[j][e * B + b] : ACC(0); acc[e][1 + b] += wv[b]; }
#pragma unroll
                        for (int b = 0; b < B; ++b)
#pragma unroll
                            for (int b2 = b; b2 < B; ++b2) acc[e][gk_sym_idx(B, b, b2)] += wv[b] * wv[b2];
                    }
                }
            }
        }
#pragma unroll
        for (int s = 0; s < S; ++s) {
            slots[((size_t)u * S + s) * GK_SLOT_PITCH + 2 * lane] = acc[0][s];
            slots[((size_t)u * S + s) * GK_SLOT_PITCH + 2 * lane + 1] = acc[1][s];
        }
    }
    __syncthreads();

    // 4. sum the parts of each bin and write this chunk's partial  [node][bin][s]
    ACC *out = (ACC *)P.partial + (((size_t)blockIdx.z * P.nchunks + q) * P.nb_pad + (size_t)strip * GK_STRIP_COLS) * K * S;
    const int KS = K * S;
    for (int t = tid; t < GK_STRIP_COLS * KS; t += GK_STRIP_THREADS) {
        int c = t / KS, j = t - c * KS, k = j / S, s = j - k * S;
        ACC v = ACC(0);
        for (int u = ustart[k]; u < ustart[k + 1]; ++u) v += slots[((size_t)u * S + s) * GK_SLOT_PITCH + c];
        out[t] = v;
    }
}

// partial (ACC) [ndirs][nchunks][nb_pad][K*S]  ->  stats (double) [nb_pad][ndirs][K*S], fixed summation order
template <typename ACC>
__global__ void k_reduce_partials(const ACC *__restrict__ partial, double *__restrict__ stats, int ndirs, int nchunks,
                                  int nb_pad, int KS) {
    size_t total = (size_t)nb_pad * ndirs * KS;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int j = (int)(i % KS); size_t t = i / KS; int d = (int)(t % ndirs); size_t node = t / ndirs;
        double v = 0.0;
        for (int q = 0; q < nchunks; ++q) v += (double)partial[(((size_t)d * nchunks + q) * nb_pad + node) * KS + j];
        stats[i] = v;
    }
}
