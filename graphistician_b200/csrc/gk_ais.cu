// gk_ais.cu -- AIS over bit-packed adjacency matrices, batched over particles.
//
// Replaces the body of the AIS driver loop of examples/demo_ais.py:10-62 for P co-resident particles:
//   StochasticBlockModel.sample_theta   internals/graph_models.py:337-371  (tempered Beta / Dirichlet posteriors: the
//                                       R x R edge counts per particle are kept current on the device)
//   naive_sample_f for n = 0..N-1       internals/graph_models.py:261-297  (beta-weighted conditional, the diagonal entry
//                                       counted in both directions, hips log_sum_exp_sample)
//   log_lkhd                            internals/graph_models.py:42-48    (from the same R x R counts)
//
// Layout: A and A^T bit-packed, one row = nw 32-bit words (N padded to 1024 bits); per particle and class a bit mask of
// the nodes carrying that label.  A node's class counts are then popc(row & mask): one row of A streamed once serves
// every co-resident particle (k_ais_rowcnt), and the serial part of the sweep is one warp per particle (k_ais_seq)
// which corrects the block-entry counts by the nodes of its own 1024-node block that have already moved.
#include <math.h>
#include <string.h>
#include "gk_internal.cuh"

#define AIS_NB 1024              // nodes per sequential block = 32 words: lane <-> word in k_ais_seq
#define AIS_MAXR 8
#define AIS_CW 512               // words of a row per column chunk of k_ais_rowcnt (16 per lane)
#define AIS_QG 32                // mask rows (particle, class) per CTA of k_ais_rowcnt: lane q owns count q
#define AIS_RW 4                 // rows per warp (2 when the grid would not fill the machine three CTAs deep)
#define AIS_WARPS 8
#define AIS_INC_MAX_MOVES 256    // a sweep in which no particle moved more nodes than this is followed by an incremental one
#define AIS_LOG0 (-1e10)         // log 0 of the conditional tables (as gk_sbm.cu: exp underflows to 0, 0 * finite = 0)

struct gk_ais {
    int N = 0, Npad = 0, nw = 0, P = 0, R = 0;
    uint32_t *Ab = nullptr, *Atb = nullptr;   // Npad x nw
    uint2 *deg = nullptr;                     // Npad: (out-degree, in-degree), diagonal included
    uint32_t *mask = nullptr;                 // P x R x nw
    uint8_t *z = nullptr;                     // P x Npad
    uint32_t *cnt = nullptr;                  // P x AIS_NB x 2 x (R-1): block-entry class counts of the current node block
    uint32_t *cnt_all = nullptr;              // P x Npad x 2 x (R-1): the same for EVERY node, kept current move by move (incremental mode; lazily allocated)
    uint32_t *mv_list = nullptr; int *mv_cnt = nullptr;   // P x AIS_NB movers (i | a << 16 | b << 24) of the block just drawn, and their number
    bool cache_valid = false; long long last_max_moves = -1; int inc_mode = 1;     // inc_mode: 0 never, 1 after a sweep with few movers, 2 always
    long long *cell = nullptr;                // P x R x R edge counts, diagonal included
    long long *sz = nullptr;                  // P x R
    long long *nchg = nullptr;                // P
    double *tab = nullptr;                    // P x (1 + R + 2 R R): beta, log pi, log B, log(1-B)
    double *u = nullptr;                      // P x Npad uniforms of the sweep: injected, or k_ais_uniforms (allocated on first use)
    cudaStream_t sx[3] = {nullptr, nullptr, nullptr}; cudaEvent_t ev_fork = nullptr, ev_join[3] = {nullptr, nullptr, nullptr};   // particle groups 2 - 4
    double *lp = nullptr;                     // P x N x R conditionals (allocated on first use)
    bool packed = false, degrees = false, labelled = false;
    uint32_t sweep = 0;
    size_t bytes = 0;
};

// ---------------------------------------------------------------------------------------------------------------
// packing
// ---------------------------------------------------------------------------------------------------------------
// rows [row0, row0+nrows) of a u8 matrix -> words of A's rows; one warp per (row, 32-column group) via ballot
__global__ void k_ais_pack_rows(const uint8_t *__restrict__ src, size_t lds, int N, int nrows, int row0, uint32_t *__restrict__ Ab, int nw) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarp = ((long long)gridDim.x * blockDim.x) >> 5;
    const int wN = (N + 31) >> 5;
    for (long long t = warp; t < (long long)nrows * wN; t += nwarp) {
        const int r = (int)(t / wN), w = (int)(t % wN);
        const int c = w * 32 + lane;
        const unsigned bit = c < N ? (src[(size_t)r * lds + c] != 0) : 0u;
        const unsigned word = __ballot_sync(0xffffffffu, bit);
        if (lane == 0) Ab[(size_t)(row0 + r) * nw + w] = word;
    }
}
// the same rows as bit COLUMNS of A^T: thread per column c, 32 source rows -> one word of At[c]
__global__ void k_ais_pack_cols(const uint8_t *__restrict__ src, size_t lds, int N, int nrows, int row0, uint32_t *__restrict__ Atb, int nw) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    const int g = blockIdx.y;                                   // group of 32 source rows
    if (c >= N) return;
    unsigned word = 0;
    for (int k = 0; k < 32; ++k) {
        const int r = g * 32 + k;
        if (r < nrows && src[(size_t)r * lds + c] != 0) word |= 1u << k;
    }
    Atb[(size_t)c * nw + (row0 >> 5) + g] = word;
}
__global__ void k_ais_degrees(const uint32_t *__restrict__ Ab, const uint32_t *__restrict__ Atb, int Npad, int nw, uint2 *__restrict__ deg) {
    const int lane = threadIdx.x & 31;
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row >= Npad) return;
    unsigned so = 0, si = 0;
    for (int w = lane; w < nw; w += 32) { so += __popc(Ab[(size_t)row * nw + w]); si += __popc(Atb[(size_t)row * nw + w]); }
    so = __reduce_add_sync(0xffffffffu, so); si = __reduce_add_sync(0xffffffffu, si);
    if (lane == 0) deg[row] = make_uint2(so, si);
}

// ---------------------------------------------------------------------------------------------------------------
// labels -> class masks and class sizes
// ---------------------------------------------------------------------------------------------------------------
__global__ void k_ais_masks(const uint8_t *__restrict__ z, int Npad, int nw, int N, int P, int R, uint32_t *__restrict__ mask) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (warp >= (long long)P * nw) return;
    const int p = (int)(warp / nw), w = (int)(warp % nw);
    const int n = w * 32 + lane;
    const int lab = n < N ? z[(size_t)p * Npad + n] : 255;
    for (int r = 0; r < R; ++r) {
        const unsigned word = __ballot_sync(0xffffffffu, lab == r);
        if (lane == 0) mask[((size_t)p * R + r) * nw + w] = word;
    }
}
__global__ void k_ais_sizes(const uint32_t *__restrict__ mask, int nw, int PR, long long *__restrict__ sz) {
    const int lane = threadIdx.x & 31;
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row >= PR) return;
    unsigned s = 0;
    for (int w = lane; w < nw; w += 32) s += __popc(mask[(size_t)row * nw + w]);
    s = __reduce_add_sync(0xffffffffu, s);
    if (lane == 0) sz[row] = s;
}

// ---------------------------------------------------------------------------------------------------------------
// K2a: class counts of a block of rows for every (particle, class) mask:  cnt[dir][i][q] = popc(row_i & mask_q)
// ---------------------------------------------------------------------------------------------------------------
// carry-save adder: h:l = a + b + c bitwise
#define AIS_CSA(h, l, a, b, c) { const uint32_t _u = (a) ^ (b); h = ((a) & (b)) | (_u & (c)); l = _u ^ (c); }

// population count of sixteen words a[k] & m[k] with 5 POPC instead of 16 (Harley-Seal: 15 carry-save adders)
__device__ __forceinline__ unsigned ais_popc16(const uint32_t *a, const uint4 m0, const uint4 m1, const uint4 m2, const uint4 m3) {
    const uint32_t x0 = a[0] & m0.x, x1 = a[1] & m0.y, x2 = a[2] & m0.z, x3 = a[3] & m0.w;
    const uint32_t x4 = a[4] & m1.x, x5 = a[5] & m1.y, x6 = a[6] & m1.z, x7 = a[7] & m1.w;
    const uint32_t x8 = a[8] & m2.x, x9 = a[9] & m2.y, x10 = a[10] & m2.z, x11 = a[11] & m2.w;
    const uint32_t x12 = a[12] & m3.x, x13 = a[13] & m3.y, x14 = a[14] & m3.z, x15 = a[15] & m3.w;
    uint32_t ones, twos, fours, eights, sixteens, t0, t1, f0, f1, e0, e1;
    AIS_CSA(t0, ones, x0, x1, x2);
    AIS_CSA(t1, ones, ones, x3, x4);
    AIS_CSA(f0, twos, t0, t1, 0u);
    AIS_CSA(t0, ones, ones, x5, x6);
    AIS_CSA(t1, ones, ones, x7, x8);
    AIS_CSA(f1, twos, twos, t0, t1);
    AIS_CSA(e0, fours, f0, f1, 0u);
    AIS_CSA(t0, ones, ones, x9, x10);
    AIS_CSA(t1, ones, ones, x11, x12);
    AIS_CSA(f0, twos, twos, t0, t1);
    AIS_CSA(t0, ones, ones, x13, x14);
    AIS_CSA(t1, ones, ones, x15, 0u);
    AIS_CSA(f1, twos, twos, t0, t1);
    AIS_CSA(e1, fours, fours, f0, f1);
    AIS_CSA(sixteens, eights, e0, e1, 0u);
    return 16u * __popc(sixteens) + 8u * __popc(eights) + 4u * __popc(fours) + 2u * __popc(twos) + __popc(ones);
}

// grid (row tiles of AIS_WARPS*AIS_RW rows, mask groups of AIS_QG, 2 directions); dynamic shared memory AIS_QG*AIS_CW words
template <int RW>
__global__ void __launch_bounds__(AIS_WARPS * 32)
k_ais_rowcnt(const uint32_t *__restrict__ Ab, const uint32_t *__restrict__ Atb, const uint32_t *__restrict__ mask, int nw, int R,
             int n0, int qbeg, int Q, uint32_t *__restrict__ cnt, size_t stride, int off) {
    extern __shared__ uint4 s_m4[];                                   // [AIS_QG][AIS_CW/4]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int q0 = qbeg + blockIdx.y * AIS_QG;
    const uint32_t *bits = blockIdx.z ? Atb : Ab;
    const int row_base = blockIdx.x * (AIS_WARPS * RW) + warp * RW;
    unsigned acc[RW];
#pragma unroll
    for (int r = 0; r < RW; ++r) acc[r] = 0;
    const int nq = min(AIS_QG, Q - q0);
    for (int c0 = 0; c0 < nw; c0 += AIS_CW) {
        __syncthreads();
        // mask rows of this group: q -> (particle q / (R-1), class q % (R-1)) -> mask row p * R + r
        for (int t = tid; t < AIS_QG * (AIS_CW / 4); t += AIS_WARPS * 32) {
            const int q = t / (AIS_CW / 4), w4 = t % (AIS_CW / 4);
            uint4 v = make_uint4(0, 0, 0, 0);
            if (q < nq && c0 + w4 * 4 < nw) {
                const int qq = q0 + q, p = qq / (R - 1), r = qq % (R - 1);
                v = *reinterpret_cast<const uint4 *>(mask + ((size_t)p * R + r) * nw + c0 + w4 * 4);
            }
            s_m4[t] = v;
        }
        __syncthreads();
#pragma unroll
        for (int rr = 0; rr < RW; ++rr) {
            const int i = row_base + rr;
            uint32_t a[16];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int w = c0 + k * 128 + lane * 4;
                uint4 v = make_uint4(0, 0, 0, 0);
                if (i < AIS_NB && w < nw) v = *reinterpret_cast<const uint4 *>(bits + (size_t)(n0 + i) * nw + w);
                a[4 * k] = v.x; a[4 * k + 1] = v.y; a[4 * k + 2] = v.z; a[4 * k + 3] = v.w;
            }
            for (int q = 0; q < nq; ++q) {
                const uint4 *m = s_m4 + q * (AIS_CW / 4) + lane;
                unsigned s = ais_popc16(a, m[0], m[32], m[64], m[96]);
                s = __reduce_add_sync(0xffffffffu, s);
                if (lane == q) acc[rr] += s;
            }
        }
    }
#pragma unroll
    for (int rr = 0; rr < RW; ++rr) {
        const int i = row_base + rr;
        if (i < AIS_NB && lane < nq) {          // cnt[particle][node][direction][class]: one contiguous record per (particle, node)
            const int qq = q0 + lane, p = qq / (R - 1), r = qq % (R - 1);
            cnt[(((size_t)p * stride + off + i) * 2 + blockIdx.z) * (R - 1) + r] = acc[rr];
        }
    }
}

// fold the out-direction counts of a block into the R x R cell table (set-up and the recompute check only)
__global__ void k_ais_fold_cells(const uint32_t *__restrict__ cnt, const uint2 *__restrict__ deg, const uint8_t *__restrict__ z,
                                 int Npad, int N, int n0, int R, long long *__restrict__ cell) {
    const int p = blockIdx.x;
    __shared__ unsigned long long s_cell[AIS_MAXR * AIS_MAXR];
    for (int t = threadIdx.x; t < R * R; t += blockDim.x) s_cell[t] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < AIS_NB; i += blockDim.x) {
        const int n = n0 + i;
        if (n >= N) break;
        const int a = z[(size_t)p * Npad + n];
        unsigned rest = deg[n].x;
        for (int r = 0; r < R - 1; ++r) {
            const unsigned v = cnt[(((size_t)p * AIS_NB + i) * 2) * (R - 1) + r];
            rest -= v;
            if (v) atomicAdd(&s_cell[a * R + r], (unsigned long long)v);
        }
        if (rest) atomicAdd(&s_cell[a * R + R - 1], (unsigned long long)rest);
    }
    __syncthreads();
    for (int t = threadIdx.x; t < R * R; t += blockDim.x) cell[(size_t)p * R * R + t] += (long long)s_cell[t];
}

// ---------------------------------------------------------------------------------------------------------------
// K2b: the serial part.  One warp per particle walks the 1024 nodes of the block in index order.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t ais_mix(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__device__ __forceinline__ double ais_uniform(uint64_t key, uint32_t n) {
    return (double)(ais_mix(key + n) >> 11) * (1.0 / 9007199254740992.0);
}

// the sweep's uniforms, one per (particle, node), when the caller injects none
__global__ void k_ais_uniforms(double *__restrict__ u, int P, int N, int Npad, uint64_t seed, uint32_t sweep) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (long long)P * Npad) return;
    const int p = (int)(t / Npad), n = (int)(t % Npad);
    const uint64_t key = ais_mix(ais_mix(seed ^ ((uint64_t)sweep << 32)) + (uint64_t)p * 0xD1B54A32D192ED03ull);
    u[t] = n < N ? ais_uniform(key, (uint32_t)n) : 0.0;
}

struct AisSeqParams {
    const uint32_t *Ab, *Atb, *cnt; const uint2 *deg;
    uint32_t *mask; uint8_t *z;
    long long *cell, *sz, *nchg;
    const double *tab, *u; double *lp;
    int N, Npad, nw, n0, p0;
    size_t cnt_stride; int cnt_off;           // cnt holds records of cnt_stride nodes per particle; this block starts at cnt_off
    uint32_t *mv_list; int *mv_cnt;           // movers of this block (incremental mode) or null
};

#define AIS_RING 64               // nodes held by the cp.async ring of the serial kernel: two windows of 32 nodes

__device__ __forceinline__ void ais_cp4(uint32_t *smem, const uint32_t *g) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(g) : "memory");
}
__device__ __forceinline__ void ais_cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N_> __device__ __forceinline__ void ais_cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N_) : "memory"); }

template <int R> static constexpr size_t ais_seq_smem() {
    return (size_t)AIS_NB * (2 * (R - 1) * 4 + 8 + 8) + (size_t)AIS_RING * 2 * 32 * 4;
}

// ln_pi_post of one node and its draw (internals/graph_models.py:261-297).  co / ci: the node's out- / in-neighbours per
// class under the current labels, the node itself included under its current label a; it is excluded here and re-added, as
// candidate k, in both directions (:275-293).  hips log_sum_exp_sample: first index whose normalised cdf reaches u -- decided
// in fp32 when u is clearly inside a cell of the cdf (error of the fp32 cdf < 1e-5 of its total, margin 1e-4), else by the
// fp64 arithmetic of the reference.
template <int R>
__device__ __forceinline__ int ais_decide(const int (&co)[R], const int (&ci)[R], const int (&sz)[R], int a, int self, double u,
                                          double beta, const double *lpi, const double *lB, const double *l1B, double (&lp)[R]) {
    double dco[R], dci[R], dsz[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int me = (r == a);
        dco[r] = (double)(co[r] - (me & self)); dci[r] = (double)(ci[r] - (me & self)); dsz[r] = (double)(sz[r] - me);
    }
    double mx = -1.7976931348623157e308;
#pragma unroll
    for (int k = 0; k < R; ++k) {
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            s0 = fma(dco[r], lB[k * R + r], s0);  s1 = fma(dsz[r] - dco[r], l1B[k * R + r], s1);
            s2 = fma(dci[r], lB[r * R + k], s2);  s3 = fma(dsz[r] - dci[r], l1B[r * R + k], s3);
        }
        const double sf = 2.0 * (self ? lB[k * R + k] : l1B[k * R + k]);
        lp[k] = lpi[k] + beta * (((s0 + s1) + (s2 + s3)) + sf);
        mx = fmax(mx, lp[k]);
    }
    int b = 0;
    float ef[R], totf = 0.f;
#pragma unroll
    for (int k = 0; k < R; ++k) { ef[k] = __expf((float)(lp[k] - mx)); totf += ef[k]; }
    const float thr = (float)u * totf, band = 1e-4f * totf;
    float cumf = 0.f; bool sure = true;
#pragma unroll
    for (int k = 0; k < R - 1; ++k) { cumf += ef[k]; b += (cumf < thr); sure = sure && (fabsf(cumf - thr) > band); }
    if (!sure) {
        double e[R], tot = 0.0, cum = 0.0;
#pragma unroll
        for (int k = 0; k < R; ++k) { e[k] = exp(lp[k] - mx); tot += e[k]; }
        b = 0;
#pragma unroll
        for (int k = 0; k < R; ++k) { cum += e[k] / tot; b += (cum < u); }
        b = min(b, R - 1);
    }
    return b;
}

// One warp per particle walks the block's 1024 nodes in index order, a window of 32 nodes at a time.  The block's small
// per-node records (entry counts, degrees, uniforms) are staged in shared memory once; the nodes' words of A / A^T (lane =
// word of the block's node bits) come through a cp.async ring of two windows.  Per window: the block-entry counts of its 32
// nodes are corrected by popc(row & current mask) - popc(row & entry mask) over the block's own words (lane = word, one
// redux per node and class; skipped while no node of the block has moved); then lane j owns node 32 w + j, all 32
// conditionals and draws are formed side by side, everything up to the first mover commits (those nodes saw exactly the
// state the serial loop would have shown them), the mover's bit goes from class a to class b in the later lanes' counts,
// and the remaining lanes draw again: (movers + 1) rounds per window, never more than the 32 of a node-by-node walk.
template <int R>
__global__ void __launch_bounds__(32) k_ais_seq(AisSeqParams P) {
    const int p = P.p0 + blockIdx.x, lane = threadIdx.x;
    constexpr int TS = 1 + R + 2 * R * R, CE = 2 * (R - 1);
    extern __shared__ __align__(16) unsigned char s_dyn[];
    uint32_t *s_ce = reinterpret_cast<uint32_t *>(s_dyn);                 // [AIS_NB][CE]
    uint2 *s_dg = reinterpret_cast<uint2 *>(s_ce + AIS_NB * CE);          // [AIS_NB]
    double *s_u = reinterpret_cast<double *>(s_dg + AIS_NB);              // [AIS_NB]
    uint32_t *s_ring = reinterpret_cast<uint32_t *>(s_u + AIS_NB);        // [AIS_RING][2][32]
    __shared__ double s_tab[TS];
    __shared__ int s_cell[R * R];                    // change of the cell table over this block
    const int nb = min(AIS_NB, P.N - P.n0), nwin = (nb + 31) >> 5;
    const uint32_t *gAo = P.Ab + (size_t)P.n0 * P.nw + (P.n0 >> 5) + lane;
    const uint32_t *gAi = P.Atb + (size_t)P.n0 * P.nw + (P.n0 >> 5) + lane;
    auto issue_window = [&](int w) {
        for (int j = 0; j < 32; ++j) {
            const int k = 32 * w + j;
            if (k < nb) {
                uint32_t *slot = s_ring + (size_t)(k & (AIS_RING - 1)) * 64;
                ais_cp4(slot + lane, gAo + (size_t)k * P.nw);
                ais_cp4(slot + 32 + lane, gAi + (size_t)k * P.nw);
            }
        }
        ais_cp_commit();
    };
    issue_window(0);
    {
        const uint4 *g = reinterpret_cast<const uint4 *>(P.cnt + ((size_t)p * P.cnt_stride + P.cnt_off) * CE);
        uint4 *d = reinterpret_cast<uint4 *>(s_ce);
        for (int t = lane; t < AIS_NB * CE / 4; t += 32) d[t] = g[t];
        const uint4 *g2 = reinterpret_cast<const uint4 *>(P.deg + P.n0);
        uint4 *d2 = reinterpret_cast<uint4 *>(s_dg);
        for (int t = lane; t < AIS_NB / 2; t += 32) d2[t] = g2[t];
        const uint4 *g3 = reinterpret_cast<const uint4 *>(P.u + (size_t)p * P.Npad + P.n0);
        uint4 *d3 = reinterpret_cast<uint4 *>(s_u);
        for (int t = lane; t < AIS_NB / 2; t += 32) d3[t] = g3[t];
    }
    for (int t = lane; t < TS; t += 32) s_tab[t] = P.tab[(size_t)p * TS + t];
    for (int t = lane; t < R * R; t += 32) s_cell[t] = 0;
    __syncwarp();
    const double beta = s_tab[0];
    const double *lpi = s_tab + 1, *lB = s_tab + 1 + R, *l1B = s_tab + 1 + R + R * R;
    uint32_t E[R], C[R];
    int sz[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        E[r] = P.mask[((size_t)p * R + r) * P.nw + (P.n0 >> 5) + lane];
        C[r] = E[r];
        sz[r] = (int)P.sz[p * R + r];
    }
    int nchg = 0;
    // one committed move a -> b of node i (uniform arguments): masks, sizes, labels, the particle's cell table
    auto commit_move = [&](int i, int a, int b, const int (&cox)[R], const int (&cix)[R], int self) {
        ++nchg;
        const uint32_t bit = (lane == (i >> 5)) ? (1u << (i & 31)) : 0u;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (r == a) { C[r] &= ~bit; sz[r] -= 1; }
            if (r == b) { C[r] |= bit; sz[r] += 1; }
        }
        if (lane == 0) {
            P.z[(size_t)p * P.Npad + P.n0 + i] = (uint8_t)b;
            if (P.mv_list) P.mv_list[(size_t)p * AIS_NB + nchg - 1] = (uint32_t)i | ((uint32_t)a << 16) | ((uint32_t)b << 24);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                s_cell[a * R + r] -= cox[r]; s_cell[b * R + r] += cox[r];          // edges n -> class r change their source class
                s_cell[r * R + a] -= cix[r]; s_cell[r * R + b] += cix[r];          // edges class r -> n change their target class
            }
            s_cell[a * R + a] -= self; s_cell[b * R + b] += self;
        }
    };
    for (int w = 0; w < nwin; ++w) {
        if (w + 1 < nwin) issue_window(w + 1); else ais_cp_commit();
        ais_cp_wait<1>();
        __syncwarp();                                 // the window path reads words other lanes copied
        const int i0 = 32 * w, cnt = min(32, nb - i0);
        {
            // lane j owns node i0 + j
            const int i = i0 + lane;
            const bool mine = lane < cnt;
            int co[R], ci[R];
#pragma unroll
            for (int r = 0; r < R - 1; ++r) { co[r] = mine ? (int)s_ce[i * CE + r] : 0; ci[r] = mine ? (int)s_ce[i * CE + R - 1 + r] : 0; }
            if (nchg) {
                for (int jj = 0; jj < cnt; ++jj) {
                    const uint32_t *slot = s_ring + (size_t)((i0 + jj) & (AIS_RING - 1)) * 64;
                    const uint32_t ao = slot[lane], ai = slot[32 + lane];
#pragma unroll
                    for (int r = 0; r < R - 1; ++r) {
                        const int d_o = __popc(ao & C[r]) - __popc(ao & E[r]) + 32;
                        const int d_i = __popc(ai & C[r]) - __popc(ai & E[r]) + 32;
                        const unsigned red = __reduce_add_sync(0xffffffffu, (unsigned)d_o | ((unsigned)d_i << 16));
                        if (lane == jj) { co[r] += (int)(red & 0xffffu) - 1024; ci[r] += (int)(red >> 16) - 1024; }
                    }
                }
            }
            {
                int so = 0, si = 0;
#pragma unroll
                for (int r = 0; r < R - 1; ++r) { so += co[r]; si += ci[r]; }
                const uint2 dg = mine ? s_dg[i] : make_uint2(0, 0);
                co[R - 1] = (int)dg.x - so;
                ci[R - 1] = (int)dg.y - si;
            }
            const uint32_t *myslot = s_ring + (size_t)(i & (AIS_RING - 1)) * 64;
            const uint32_t my_o = mine ? myslot[w] : 0u, my_i = mine ? myslot[32 + w] : 0u;   // my row / column over this window's nodes
            const int self = (my_o >> lane) & 1;
            int a = 0;
#pragma unroll
            for (int r = 1; r < R; ++r) a += r * (int)((__shfl_sync(0xffffffffu, C[r], w) >> lane) & 1u);
            const double u = mine ? s_u[i] : 0.5;
            unsigned pending = cnt == 32 ? 0xffffffffu : ((1u << cnt) - 1u);
            while (pending) {
                double lp[R];
                const int b = ais_decide<R>(co, ci, sz, a, self, u, beta, lpi, lB, l1B, lp);
                const unsigned mv = __ballot_sync(0xffffffffu, b != a) & pending;
                const int jstar = mv ? (__ffs(mv) - 1) : 32;
                const unsigned done = jstar >= 31 ? pending : (pending & ((2u << jstar) - 1u));     // committed by this round
                if (P.lp && ((done >> lane) & 1u)) {
#pragma unroll
                    for (int k = 0; k < R; ++k) P.lp[((size_t)p * P.N + P.n0 + i) * R + k] = lp[k];
                }
                pending &= ~done;
                if (!mv) break;
                const int ma = __shfl_sync(0xffffffffu, a, jstar), mb = __shfl_sync(0xffffffffu, b, jstar);
                const int mself = __shfl_sync(0xffffffffu, self, jstar);
                int cox[R], cix[R];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    cox[r] = __shfl_sync(0xffffffffu, co[r], jstar) - ((r == ma) & mself);
                    cix[r] = __shfl_sync(0xffffffffu, ci[r], jstar) - ((r == ma) & mself);
                }
                commit_move(i0 + jstar, ma, mb, cox, cix, mself);
                if (lane == jstar) a = mb;
                // the later nodes of the window: their edge to / from the mover changes class
                const int x_o = (int)((my_o >> jstar) & 1u), x_i = (int)((my_i >> jstar) & 1u);
                if (lane != jstar) {
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        if (r == ma) { co[r] -= x_o; ci[r] -= x_i; }
                        if (r == mb) { co[r] += x_o; ci[r] += x_i; }
                    }
                }
            }
        }
        __syncwarp();                                 // ring slots of this window are free for the window after next
    }
    __syncwarp();
#pragma unroll
    for (int r = 0; r < R; ++r) {
        P.mask[((size_t)p * R + r) * P.nw + (P.n0 >> 5) + lane] = C[r];
        if (lane == 0) P.sz[p * R + r] = sz[r];
    }
    for (int t = lane; t < R * R; t += 32) P.cell[(size_t)p * R * R + t] += s_cell[t];
    if (lane == 0) { P.nchg[p] += nchg; if (P.mv_cnt) P.mv_cnt[p] = nchg; }
}

// Incremental mode: the class counts of EVERY node are kept in cnt_all and moved along with the labels.  After a block has been
// drawn, each of its movers m (a -> b) changes, for every node j, the out-count of j by A[j, m] and the in-count by A[m, j]:
// one bit of row m of A^T resp. A per node.  Thread = node j (it owns its record: no atomics), the two rows are read as words
// shared by the 32 nodes of a warp.  With no mover the kernel returns at once -- and the sweep needs no counting pass at all.
template <int R>
__global__ void __launch_bounds__(256) k_ais_apply(const uint32_t *__restrict__ Ab, const uint32_t *__restrict__ Atb, int nw, int Npad, int N,
                                                   int n0, int p0, const uint32_t *__restrict__ mv_list, const int *__restrict__ mv_cnt,
                                                   uint32_t *__restrict__ cnt_all) {
    constexpr int CE = 2 * (R - 1);
    const int p = p0 + blockIdx.y, nm = mv_cnt[p];
    const int j = blockIdx.x * 256 + threadIdx.x;
    if (nm == 0 || j >= N) return;
    int d[CE];
#pragma unroll
    for (int t = 0; t < CE; ++t) d[t] = 0;
    for (int q = 0; q < nm; ++q) {
        const uint32_t e = mv_list[(size_t)p * AIS_NB + q];
        const int m = n0 + (int)(e & 0xffffu), a = (int)((e >> 16) & 0xffu), b = (int)(e >> 24);
        const int xo = (int)((Atb[(size_t)m * nw + (j >> 5)] >> (j & 31)) & 1u);      // A[j, m]
        const int xi = (int)((Ab[(size_t)m * nw + (j >> 5)] >> (j & 31)) & 1u);       // A[m, j]
#pragma unroll
        for (int r = 0; r < R - 1; ++r) {
            if (r == a) { d[r] -= xo; d[R - 1 + r] -= xi; }
            if (r == b) { d[r] += xo; d[R - 1 + r] += xi; }
        }
    }
    uint32_t *c = cnt_all + ((size_t)p * Npad + j) * CE;
#pragma unroll
    for (int t = 0; t < CE; ++t) if (d[t]) c[t] = (uint32_t)((int)c[t] + d[t]);
}

// ---------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------
static int ais_check(gk_ctx *ctx, gk_ais *a) {
    if (!ctx || !a) return GK_EINVAL;
    cudaSetDevice(ctx->device);
    return GK_OK;
}

extern "C" int gk_ais_create(gk_ctx *ctx, int N, int P, int R, gk_ais **out) {
    if (!ctx || !out || N < 1 || P < 1) return GK_EINVAL;
    *out = nullptr;
    if (R < 1 || R > AIS_MAXR) return gk_fail(ctx, GK_EUNSUPPORTED, "gk_ais_create: R = %d outside 1..%d (template instantiations)", R, AIS_MAXR);
    cudaSetDevice(ctx->device);
    gk_ais *a = new gk_ais();
    a->N = N; a->P = P; a->R = R;
    a->Npad = (N + AIS_NB - 1) / AIS_NB * AIS_NB;
    a->nw = a->Npad / 32;
    const size_t bits = ((size_t)a->Npad + 4) * a->nw * 4;       // + 4 rows: the serial kernel's read-ahead
    struct { void **p; size_t n; } allocs[] = {
        {(void **)&a->Ab, bits}, {(void **)&a->Atb, bits}, {(void **)&a->deg, ((size_t)a->Npad + 4) * 8},
        {(void **)&a->mask, (size_t)P * R * a->nw * 4}, {(void **)&a->z, (size_t)P * a->Npad + 4},
        {(void **)&a->cnt, ((size_t)P * AIS_NB + 4) * 2 * (R > 1 ? R - 1 : 1) * 4},
        {(void **)&a->u, ((size_t)P * a->Npad + 4) * 8}, {(void **)&a->cell, (size_t)P * R * R * 8},
        {(void **)&a->sz, (size_t)P * R * 8}, {(void **)&a->nchg, (size_t)P * 8},
        {(void **)&a->tab, (size_t)P * (1 + R + 2 * R * R) * 8}};
    for (auto &al : allocs) {
        if (cudaMalloc(al.p, al.n) != cudaSuccess) {
            for (auto &f : allocs) if (*f.p) cudaFree(*f.p);
            delete a;
            return gk_fail(ctx, GK_ENOMEM, "gk_ais_create: cudaMalloc(%zu) failed", al.n);
        }
        cudaMemsetAsync(*al.p, 0, al.n, ctx->stream);
        a->bytes += al.n;
    }
    bool ok = cudaEventCreateWithFlags(&a->ev_fork, cudaEventDisableTiming) == cudaSuccess;
    for (int g = 0; g < 3; ++g)
        ok = ok && cudaStreamCreateWithFlags(&a->sx[g], cudaStreamNonBlocking) == cudaSuccess &&
             cudaEventCreateWithFlags(&a->ev_join[g], cudaEventDisableTiming) == cudaSuccess;
    if (!ok)
        return gk_fail(ctx, GK_ECUDA, "gk_ais_create: stream / event creation failed");
    *out = a;
    return GK_OK;
}

extern "C" int gk_ais_destroy(gk_ctx *ctx, gk_ais *a) {
    if (ais_check(ctx, a)) return GK_EINVAL;
    cudaStreamSynchronize(ctx->stream);
    for (int g = 0; g < 3; ++g) {
        if (a->sx[g]) { cudaStreamSynchronize(a->sx[g]); cudaStreamDestroy(a->sx[g]); }
        if (a->ev_join[g]) cudaEventDestroy(a->ev_join[g]);
    }
    if (a->ev_fork) cudaEventDestroy(a->ev_fork);
    void *ptrs[] = {a->Ab, a->Atb, a->deg, a->mask, a->z, a->cnt, a->cell, a->sz, a->nchg, a->tab, a->u, a->lp, a->cnt_all, a->mv_list, a->mv_cnt};
    for (void *p : ptrs) if (p) cudaFree(p);
    delete a;
    return GK_OK;
}

extern "C" size_t gk_ais_device_bytes(gk_ais *a) { return a ? a->bytes : 0; }

extern "C" int gk_ais_set_option(gk_ctx *ctx, gk_ais *a, int option, int value) {
    if (!ctx || !a) return GK_EINVAL;
    if (option == GK_AIS_OPT_INCREMENTAL && value >= 0 && value <= 2) { a->inc_mode = value; return GK_OK; }
    return gk_fail(ctx, GK_EINVAL, "gk_ais_set_option: unknown option %d or value %d", option, value);
}

static int ais_pack(gk_ctx *ctx, gk_ais *a, const uint8_t *src, size_t lds, int row0, int nrows) {
    k_ais_pack_rows<<<ctx->num_sms * 8, 256, 0, ctx->stream>>>(src, lds, a->N, nrows, row0, a->Ab, a->nw);
    GK_CHECK_LAUNCH(ctx);
    dim3 g((a->N + 255) / 256, (nrows + 31) / 32);
    k_ais_pack_cols<<<g, 256, 0, ctx->stream>>>(src, lds, a->N, nrows, row0, a->Atb, a->nw);
    GK_CHECK_LAUNCH(ctx);
    a->packed = true; a->degrees = false; a->cache_valid = false;
    return GK_OK;
}

extern "C" int gk_ais_upload_rows(gk_ctx *ctx, gk_ais *a, const uint8_t *A_rows, int row0, int nrows, int src_on_device) {
    if (ais_check(ctx, a) || !A_rows) return GK_EINVAL;
    if (row0 < 0 || nrows < 1 || row0 + nrows > a->N || (row0 & 31))
        return gk_fail(ctx, GK_EINVAL, "gk_ais_upload_rows: rows [%d, %d) must lie in [0, N) and start on a multiple of 32", row0, row0 + nrows);
    const uint8_t *src = A_rows;
    if (!src_on_device) {
        void *tmp = gk_scratch(ctx, 0, (size_t)nrows * a->N);
        if (!tmp) return GK_ENOMEM;
        GK_CUDA(ctx, cudaMemcpyAsync(tmp, A_rows, (size_t)nrows * a->N, cudaMemcpyHostToDevice, ctx->stream));
        src = (const uint8_t *)tmp;
    }
    int rc = ais_pack(ctx, a, src, (size_t)a->N, row0, nrows);
    if (rc) return rc;
    if (!src_on_device) GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));     // the caller may reuse its buffer
    return GK_OK;
}

extern "C" int gk_ais_upload_network(gk_ctx *ctx, gk_ais *a, gk_net *net) {
    if (ais_check(ctx, a) || !net) return GK_EINVAL;
    if (net->N != a->N || net->row0 != 0 || net->nrows != net->N)
        return gk_fail(ctx, GK_EINVAL, "gk_ais_upload_network: needs an unsharded network of the same N");
    return ais_pack(ctx, a, net->A, (size_t)net->ld, 0, a->N);
}

static int ais_degrees(gk_ctx *ctx, gk_ais *a) {
    if (a->degrees) return GK_OK;
    if (!a->packed) return gk_fail(ctx, GK_EINVAL, "gk_ais: no adjacency rows uploaded yet");
    k_ais_degrees<<<(a->Npad * 32 + 255) / 256, 256, 0, ctx->stream>>>(a->Ab, a->Atb, a->Npad, a->nw, a->deg);
    GK_CHECK_LAUNCH(ctx);
    a->degrees = true;
    return GK_OK;
}

extern "C" int gk_ais_nnz(gk_ctx *ctx, gk_ais *a, int64_t *nnz) {
    if (ais_check(ctx, a) || !nnz) return GK_EINVAL;
    int rc = ais_degrees(ctx, a); if (rc) return rc;
    std::vector<uint2> d(a->Npad);
    GK_CUDA(ctx, cudaMemcpyAsync(d.data(), a->deg, (size_t)a->Npad * 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    int64_t s = 0; for (const uint2 &v : d) s += v.x;
    *nnz = s;
    return GK_OK;
}

// block-entry counts of node block n0 for the particles [p0, p1), on `st`
static int ais_rowcnt(gk_ctx *ctx, gk_ais *a, int n0, int p0, int p1, cudaStream_t st, bool to_all = false) {
    const int qbeg = p0 * (a->R - 1), Q = p1 * (a->R - 1);
    if (Q == qbeg) return GK_OK;
    const size_t sm = (size_t)AIS_QG * AIS_CW * 4;
    if (gk_attr_needed(ctx, (const void *)k_ais_rowcnt<AIS_RW>)) {
        GK_CUDA(ctx, cudaFuncSetAttribute(k_ais_rowcnt<AIS_RW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        GK_CUDA(ctx, cudaFuncSetAttribute(k_ais_rowcnt<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    }
    const int qgroups = (Q - qbeg + AIS_QG - 1) / AIS_QG;
    const bool fine = (AIS_NB / (AIS_WARPS * AIS_RW)) * qgroups * 2 < 3 * ctx->num_sms;     // three CTAs fit an SM
    dim3 g(AIS_NB / (AIS_WARPS * (fine ? 2 : AIS_RW)), qgroups, 2);
    const bool timed = st == ctx->stream;
    if (timed) gk_time_begin(ctx, 0, 2.0 * AIS_NB * a->nw * 4.0);
    uint32_t *dst = to_all ? a->cnt_all : a->cnt;
    const size_t stride = to_all ? (size_t)a->Npad : (size_t)AIS_NB;
    const int off = to_all ? n0 : 0;
    if (fine) k_ais_rowcnt<2><<<g, AIS_WARPS * 32, sm, st>>>(a->Ab, a->Atb, a->mask, a->nw, a->R, n0, qbeg, Q, dst, stride, off);
    else k_ais_rowcnt<AIS_RW><<<g, AIS_WARPS * 32, sm, st>>>(a->Ab, a->Atb, a->mask, a->nw, a->R, n0, qbeg, Q, dst, stride, off);
    if (timed) gk_time_end(ctx, 0);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

// cell table and class sizes of the current labels from scratch
static int ais_recount(gk_ctx *ctx, gk_ais *a) {
    int rc = ais_degrees(ctx, a); if (rc) return rc;
    const int PR = a->P * a->R;
    k_ais_sizes<<<(PR * 32 + 255) / 256, 256, 0, ctx->stream>>>(a->mask, a->nw, PR, a->sz);
    GK_CHECK_LAUNCH(ctx);
    GK_CUDA(ctx, cudaMemsetAsync(a->cell, 0, (size_t)a->P * a->R * a->R * 8, ctx->stream));
    for (int n0 = 0; n0 < a->N; n0 += AIS_NB) {
        if ((rc = ais_rowcnt(ctx, a, n0, 0, a->P, ctx->stream))) return rc;
        k_ais_fold_cells<<<a->P, 256, 0, ctx->stream>>>(a->cnt, a->deg, a->z, a->Npad, a->N, n0, a->R, a->cell);
        GK_CHECK_LAUNCH(ctx);
    }
    return GK_OK;
}

extern "C" int gk_ais_set_labels(gk_ctx *ctx, gk_ais *a, const int32_t *z, int broadcast) {
    if (ais_check(ctx, a) || !z) return GK_EINVAL;
    const int np = broadcast ? 1 : a->P;
    std::vector<uint8_t> h((size_t)a->P * a->Npad, 0);
    for (int p = 0; p < a->P; ++p)
        for (int n = 0; n < a->N; ++n) {
            const int v = z[(size_t)(p % np) * a->N + n];
            if (v < 0 || v >= a->R) return gk_fail(ctx, GK_EINVAL, "gk_ais_set_labels: label %d of node %d outside [0, %d)", v, n, a->R);
            h[(size_t)p * a->Npad + n] = (uint8_t)v;
        }
    GK_CUDA(ctx, cudaMemcpyAsync(a->z, h.data(), h.size(), cudaMemcpyHostToDevice, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const long long warps = (long long)a->P * a->nw;
    k_ais_masks<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, ctx->stream>>>(a->z, a->Npad, a->nw, a->N, a->P, a->R, a->mask);
    GK_CHECK_LAUNCH(ctx);
    a->labelled = true; a->cache_valid = false; a->last_max_moves = -1;
    return ais_recount(ctx, a);
}

extern "C" int gk_ais_get_labels(gk_ctx *ctx, gk_ais *a, int32_t *z) {
    if (ais_check(ctx, a) || !z) return GK_EINVAL;
    std::vector<uint8_t> h((size_t)a->P * a->Npad);
    GK_CUDA(ctx, cudaMemcpyAsync(h.data(), a->z, h.size(), cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int p = 0; p < a->P; ++p)
        for (int n = 0; n < a->N; ++n) z[(size_t)p * a->N + n] = h[(size_t)p * a->Npad + n];
    return GK_OK;
}

extern "C" int gk_ais_counts(gk_ctx *ctx, gk_ais *a, int recompute, int64_t *n_edge, int64_t *sizes) {
    if (ais_check(ctx, a)) return GK_EINVAL;
    if (!a->labelled) return gk_fail(ctx, GK_EINVAL, "gk_ais_counts: call gk_ais_set_labels first");
    if (recompute) { int rc = ais_recount(ctx, a); if (rc) return rc; }
    if (n_edge) GK_CUDA(ctx, cudaMemcpyAsync(n_edge, a->cell, (size_t)a->P * a->R * a->R * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (sizes) GK_CUDA(ctx, cudaMemcpyAsync(sizes, a->sz, (size_t)a->P * a->R * 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GK_OK;
}

template <int R> static void ais_launch_apply(gk_ais *a, int n0, int p0, int np, cudaStream_t st) {
    dim3 g((a->N + 255) / 256, np);
    k_ais_apply<R><<<g, 256, 0, st>>>(a->Ab, a->Atb, a->nw, a->Npad, a->N, n0, p0, a->mv_list, a->mv_cnt, a->cnt_all);
}

template <int R> static int ais_launch_seq(gk_ctx *ctx, int np, cudaStream_t st, const AisSeqParams &prm) {
    if (gk_attr_needed(ctx, (const void *)k_ais_seq<R>))
        GK_CUDA(ctx, cudaFuncSetAttribute(k_ais_seq<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ais_seq_smem<R>()));
    k_ais_seq<R><<<np, 32, ais_seq_smem<R>(), st>>>(prm);
    return GK_OK;
}

// One sweep.  The particles are split in up to four groups on as many streams so that one group's serial kernel (one warp per
// particle) runs beside the other group's counting pass (which fills the rest of the machine).
static int ais_sweep_device(gk_ctx *ctx, gk_ais *a, const double *tab_host, bool inject, uint64_t seed, bool want_lp) {
    const int R = a->R;
    int rc = ais_degrees(ctx, a); if (rc) return rc;
    GK_CUDA(ctx, cudaMemcpyAsync(a->tab, tab_host, (size_t)a->P * (1 + R + 2 * R * R) * 8, cudaMemcpyHostToDevice, ctx->stream));
    a->sweep++;
    if (R == 1) return GK_OK;
    if (!inject) {
        const long long tot = (long long)a->P * a->Npad;
        k_ais_uniforms<<<(unsigned)((tot + 255) / 256), 256, 0, ctx->stream>>>(a->u, a->P, a->N, a->Npad, seed, a->sweep);
        GK_CHECK_LAUNCH(ctx);
    }
    // Incremental mode (the previous sweep moved few nodes): the class counts of every node live in cnt_all and follow the
    // moves (k_ais_apply), so no counting pass runs at all; entering the mode costs one counting pass over the final masks.
    bool inc = a->inc_mode == 2;
    if (a->inc_mode == 1 && a->sweep > 1) {
        std::vector<long long> h(a->P);
        GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (a->last_max_moves == -2) {               // the previous sweep's move counts are still on the device
            GK_CUDA(ctx, cudaMemcpy(h.data(), a->nchg, (size_t)a->P * 8, cudaMemcpyDeviceToHost));
            a->last_max_moves = 0;
            for (long long v : h) a->last_max_moves = std::max(a->last_max_moves, v);
        }
        inc = a->last_max_moves >= 0 && a->last_max_moves <= AIS_INC_MAX_MOVES;
    }
    GK_CUDA(ctx, cudaMemsetAsync(a->nchg, 0, (size_t)a->P * 8, ctx->stream));
    if (inc) {
        const size_t CE = 2 * (R - 1);
        if (!a->cnt_all) {
            GK_CUDA(ctx, cudaMalloc(&a->cnt_all, ((size_t)a->P * a->Npad + 4) * CE * 4));
            GK_CUDA(ctx, cudaMalloc(&a->mv_list, (size_t)a->P * AIS_NB * 4));
            GK_CUDA(ctx, cudaMalloc(&a->mv_cnt, (size_t)a->P * 4));
            a->bytes += ((size_t)a->P * a->Npad + 4) * CE * 4 + (size_t)a->P * AIS_NB * 4 + (size_t)a->P * 4;
            a->cache_valid = false;
        }
        if (!a->cache_valid) {
            for (int n0 = 0; n0 < a->N; n0 += AIS_NB)
                if ((rc = ais_rowcnt(ctx, a, n0, 0, a->P, ctx->stream, true))) return rc;
            a->cache_valid = true;
        }
    } else {
        a->cache_valid = false;
    }
    AisSeqParams prm;
    prm.Ab = a->Ab; prm.Atb = a->Atb; prm.deg = a->deg; prm.mask = a->mask; prm.z = a->z;
    prm.cnt = inc ? a->cnt_all : a->cnt; prm.cnt_stride = inc ? (size_t)a->Npad : (size_t)AIS_NB;
    prm.mv_list = inc ? a->mv_list : nullptr; prm.mv_cnt = inc ? a->mv_cnt : nullptr;
    prm.cell = a->cell; prm.sz = a->sz; prm.nchg = a->nchg; prm.tab = a->tab; prm.u = a->u;
    prm.lp = want_lp ? a->lp : nullptr;
    prm.N = a->N; prm.Npad = a->Npad; prm.nw = a->nw;
    // particle groups on their own streams: a group's serial kernel runs beside the other groups' counting passes
    const int ngroups = ctx->timing ? 1 : (a->P >= 64 ? 4 : (a->P >= 16 ? 2 : 1));      // kernel timing brackets need one stream
    const int per = (a->P + ngroups - 1) / ngroups;
    if (ngroups > 1) {
        GK_CUDA(ctx, cudaEventRecord(a->ev_fork, ctx->stream));
        for (int g = 1; g < ngroups; ++g) GK_CUDA(ctx, cudaStreamWaitEvent(a->sx[g - 1], a->ev_fork, 0));
    }
    for (int n0 = 0; n0 < a->N; n0 += AIS_NB) {
        prm.n0 = n0; prm.cnt_off = inc ? n0 : 0;
        for (int g = 0; g < ngroups; ++g) {
            const int p0 = std::min(a->P, g * per), p1 = std::min(a->P, (g + 1) * per);
            if (p1 == p0) continue;
            cudaStream_t st = g ? a->sx[g - 1] : ctx->stream;
            if (!inc && (rc = ais_rowcnt(ctx, a, n0, p0, p1, st))) return rc;
            prm.p0 = p0;
            if (!g) gk_time_begin(ctx, 1, 0.0);
            switch (R) {
                case 2: rc = ais_launch_seq<2>(ctx, p1 - p0, st, prm); if (inc) ais_launch_apply<2>(a, n0, p0, p1 - p0, st); break;
                case 3: rc = ais_launch_seq<3>(ctx, p1 - p0, st, prm); if (inc) ais_launch_apply<3>(a, n0, p0, p1 - p0, st); break;
                case 4: rc = ais_launch_seq<4>(ctx, p1 - p0, st, prm); if (inc) ais_launch_apply<4>(a, n0, p0, p1 - p0, st); break;
                case 5: rc = ais_launch_seq<5>(ctx, p1 - p0, st, prm); if (inc) ais_launch_apply<5>(a, n0, p0, p1 - p0, st); break;
                case 6: rc = ais_launch_seq<6>(ctx, p1 - p0, st, prm); if (inc) ais_launch_apply<6>(a, n0, p0, p1 - p0, st); break;
                case 7: rc = ais_launch_seq<7>(ctx, p1 - p0, st, prm); if (inc) ais_launch_apply<7>(a, n0, p0, p1 - p0, st); break;
                default: rc = ais_launch_seq<8>(ctx, p1 - p0, st, prm); if (inc) ais_launch_apply<8>(a, n0, p0, p1 - p0, st); break;
            }
            if (!g) gk_time_end(ctx, 1);
            if (rc) return rc;
            GK_CHECK_LAUNCH(ctx);
        }
    }
    for (int g = 1; g < ngroups; ++g) {
        GK_CUDA(ctx, cudaEventRecord(a->ev_join[g - 1], a->sx[g - 1]));
        GK_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, a->ev_join[g - 1], 0));
    }
    a->last_max_moves = -2;
    return GK_OK;
}

static inline double ais_log(double x) { return x > 0.0 ? fmax(log(x), AIS_LOG0) : AIS_LOG0; }

// tab[p] = (beta, log pi[R], log B[R x R], log(1 - B)[R x R])
static void ais_fill_tab(double *tab, int R, double beta, const double *log_pi, const double *B) {
    tab[0] = beta;
    for (int r = 0; r < R; ++r) tab[1 + r] = log_pi[r];
    for (int t = 0; t < R * R; ++t) { tab[1 + R + t] = ais_log(B[t]); tab[1 + R + R * R + t] = ais_log(1.0 - B[t]); }
}

extern "C" int gk_ais_sweep(gk_ctx *ctx, gk_ais *a, const double *log_pi, const double *B, const double *beta,
                            const double *uniforms, uint64_t seed, double *out_lp, int64_t *out_nchanged) {
    if (ais_check(ctx, a) || !log_pi || !B || !beta) return GK_EINVAL;
    if (!a->labelled) return gk_fail(ctx, GK_EINVAL, "gk_ais_sweep: call gk_ais_set_labels first");
    const int R = a->R, TS = 1 + R + 2 * R * R;
    std::vector<double> tab((size_t)a->P * TS);
    for (int p = 0; p < a->P; ++p) ais_fill_tab(&tab[(size_t)p * TS], R, beta[p], log_pi + (size_t)p * R, B + (size_t)p * R * R);
    if (uniforms) {
        GK_CUDA(ctx, cudaMemcpy2DAsync(a->u, (size_t)a->Npad * 8, uniforms, (size_t)a->N * 8, (size_t)a->N * 8, a->P,
                                       cudaMemcpyHostToDevice, ctx->stream));
    }
    if (out_lp && !a->lp) { GK_CUDA(ctx, cudaMalloc(&a->lp, (size_t)a->P * a->N * R * 8)); a->bytes += (size_t)a->P * a->N * R * 8; }
    int rc = ais_sweep_device(ctx, a, tab.data(), uniforms != nullptr, seed, out_lp != nullptr);
    if (rc) return rc;
    if (out_lp) GK_CUDA(ctx, cudaMemcpyAsync(out_lp, a->lp, (size_t)a->P * a->N * R * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_nchanged) GK_CUDA(ctx, cudaMemcpyAsync(out_nchanged, a->nchg, (size_t)a->P * 8, cudaMemcpyDeviceToHost, ctx->stream));
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GK_OK;
}

// ---- host random numbers of gk_ais_run (xoshiro256**, Marsaglia-Tsang gamma) ----------------------------------------
struct AisRng {
    uint64_t s[4];
    explicit AisRng(uint64_t seed) { for (auto &v : s) { seed += 0x9E3779B97F4A7C15ull; uint64_t x = seed; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull; v = x ^ (x >> 31); } }
    static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    uint64_t next() { const uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17; s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45); return r; }
    double uniform() { return ((double)(next() >> 11) + 0.5) * (1.0 / 9007199254740992.0); }
    double normal() { const double u1 = uniform(), u2 = uniform(); return sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2); }
    double gamma(double k) {
        if (k < 1.0) return gamma(k + 1.0) * pow(uniform(), 1.0 / k);
        const double d = k - 1.0 / 3.0, c = 1.0 / sqrt(9.0 * d);
        for (;;) {
            double x, v;
            do { x = normal(); v = 1.0 + c * x; } while (v <= 0.0);
            v = v * v * v;
            const double u = uniform();
            if (u < 1.0 - 0.0331 * x * x * x * x || log(u) < 0.5 * x * x + d * (1.0 - v + log(v))) return d * v;
        }
    }
    double beta(double a, double b) { const double x = gamma(a), y = gamma(b); return x / (x + y); }
};

extern "C" int gk_ais_run(gk_ctx *ctx, gk_ais *a, int model, double b0, double b1, const double *alpha0, int n_temps,
                          int steps_per_temp, const int32_t *f0, uint64_t seed, double *log_weights) {
    if (ais_check(ctx, a) || !log_weights || n_temps < 2 || steps_per_temp < 1) return GK_EINVAL;
    if (model != GK_AIS_ER && model != GK_AIS_SBM) return gk_fail(ctx, GK_EINVAL, "gk_ais_run: model %d", model);
    const int P = a->P, R = a->R, TS = 1 + R + 2 * R * R;
    AisRng rng(seed);
    for (int p = 0; p < P; ++p) log_weights[p] = 0.0;
    const double NN = (double)a->N * (double)a->N;
    if (model == GK_AIS_ER) {
        // internals/graph_models.py:168-182: rho ~ Beta(a + beta nnz, b + beta (N^2 - nnz)); no features
        int64_t nnz = 0;
        int rc = gk_ais_nnz(ctx, a, &nnz); if (rc) return rc;
        for (int p = 0; p < P; ++p)
            for (int b = 1; b < n_temps; ++b) {
                const double beta = (double)b / (n_temps - 1), prev = (double)(b - 1) / (n_temps - 1);
                double rho = 0.5;
                for (int s = 0; s < steps_per_temp; ++s) rho = rng.beta(b1 + beta * nnz, b0 + beta * (NN - nnz));
                log_weights[p] += (beta - prev) * (nnz * ais_log(rho) + (NN - nnz) * ais_log(1.0 - rho));
            }
        return GK_OK;
    }
    if (!alpha0 || !f0) return GK_EINVAL;
    int rc = gk_ais_set_labels(ctx, a, f0, 1); if (rc) return rc;
    std::vector<int64_t> cell((size_t)P * R * R), sz((size_t)P * R);
    std::vector<double> tab((size_t)P * TS), Bm((size_t)P * R * R), lpi((size_t)P * R);
    if ((rc = gk_ais_counts(ctx, a, 0, cell.data(), sz.data()))) return rc;
    for (int b = 1; b < n_temps; ++b) {
        const double beta = (double)b / (n_temps - 1), prev = (double)(b - 1) / (n_temps - 1);
        for (int s = 0; s < steps_per_temp; ++s) {
            // sample_theta (internals/graph_models.py:337-371): pi ~ Dir(alpha0 + sizes), untempered;
            // B[r1, r2] ~ Beta(b1 + beta edges, b0 + beta non-edges), stored as float32 (:353)
            for (int p = 0; p < P; ++p) {
                double tot = 0.0;
                for (int r = 0; r < R; ++r) { lpi[p * R + r] = rng.gamma(alpha0[r] + (double)sz[p * R + r]); tot += lpi[p * R + r]; }
                for (int r = 0; r < R; ++r) lpi[p * R + r] = ais_log(lpi[p * R + r] / tot);
                for (int r1 = 0; r1 < R; ++r1)
                    for (int r2 = 0; r2 < R; ++r2) {
                        const double e = (double)cell[(p * R + r1) * R + r2], np = (double)sz[p * R + r1] * (double)sz[p * R + r2];
                        Bm[(p * R + r1) * R + r2] = (double)(float)rng.beta(b1 + beta * e, b0 + beta * (np - e));
                    }
                ais_fill_tab(&tab[(size_t)p * TS], R, beta, &lpi[p * R], &Bm[(size_t)p * R * R]);
            }
            if ((rc = ais_sweep_device(ctx, a, tab.data(), false, seed, false))) return rc;
            if ((rc = gk_ais_counts(ctx, a, 0, cell.data(), sz.data()))) return rc;
        }
        // logpr(beta_b) - logpr(beta_{b-1}) at the same (f, theta): (beta_b - beta_{b-1}) log_lkhd (demo_ais.py:49-52)
        for (int p = 0; p < P; ++p) {
            double ll = 0.0;
            for (int r1 = 0; r1 < R; ++r1)
                for (int r2 = 0; r2 < R; ++r2) {
                    const double e = (double)cell[(p * R + r1) * R + r2], np = (double)sz[p * R + r1] * (double)sz[p * R + r2];
                    const double Bv = Bm[(p * R + r1) * R + r2];
                    if (e > 0) ll += e * ais_log(Bv);
                    if (np - e > 0) ll += (np - e) * ais_log(1.0 - Bv);
                }
            log_weights[p] += (beta - prev) * ll;
        }
    }
    return GK_OK;
}
