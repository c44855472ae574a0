// gk_eig_seq5.cuh -- the serial draws of one node block from the running G (one CTA), fp32 compute path; included by
// gk_eig_fast.cuh.  Same recursion as k_eig_seq4 (eigenmodel.py:884-908 inside a block):
//   prec_t[a,b] = d_ab/sigma_F + lmbda_a lmbda_b g[ab],   h_t[a] = lmbda_a (g[P+a] - (sum_m f_m - f_t)[a] - mu_0 g[Q+a]),
//   g = G[n0+t] (all partners as they were at block entry) + sum_{s<t} X[s,t] (Phi(f_s new) - Phi(f_s old)).
//
// The sweep is bound by this kernel's per-node dependency chain, so the chain is cut to what cannot be avoided
// (profiles/micro/chol_chain.cu measures the pieces on a B200: rsqrt 17, dependent FFMA 4.5, LDS 29, SHFL 26 cycles;
// an 8x8 factor-and-draw with its operands in registers: ~420 cycles):
//   * warp 0 = solver.  It keeps the previous node's new features f in REGISTERS and adds that node's rank-one change
//     itself:  prec_t = base_t + (x lmbda o f)(lmbda o f)^T,  h_t = hbase_t + coef (lmbda o f)  -- no shared-memory round
//     trip between a draw and the next conditional.  All 32 lanes run the same arithmetic on the same (broadcast) data,
//     so the warp never diverges and its named barriers stay aligned.
//   * warps on the other three SM sub-partitions = helpers.  While the solver factors node s+1 they fold node s's change
//     into the base row of node s+2:  base_{s+2} = constants + Gb[s+2] + sum_{s' <= s} X[s',s+2] dPhi_{s'} - (the old-feature
//     half of node s+1's rank-one term).  The sum over s' <= s-1 is accumulated BEFORE f_s is published (four row classes of
//     64 column threads), so only one FMA, a 4-way reduction and the row's constants follow the publication.
//   * handshakes are named barriers: 1 = "f_s published" (solver arrives, helpers wait), 2 = "base row ready" (helpers
//     arrive, solver waits in the middle of its factorisation, then loads the row under its back substitution).
// On exit: G[n0+j] += sum_{s<j} X[s,j] dPhi_s, F <- new features, sumF, and dphi[i] = dPhi_i for k_eig_push.
#pragma once

#define GK_SEQ5_THREADS 384                       // 12 warps: warp 0 solver; warps 4, 8 (its sub-partition) stay idle
#define GK_SEQ5_HELP 288                          // warps 1-3, 5-7, 9-11: 64 finishers + 224 summers
#define GK_SEQ5_SRMAX 24                          // bound on the summers' row classes (192 threads / column quads)

__device__ int gk_seq5_debug = 0;      // timing experiments only (env GK_SEQ5_DEBUG): 1 = summers skip their sums, 2 = finishers skip the reduction
__device__ __forceinline__ void gk_bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void gk_bar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }
// %clock64 read that cannot issue before `dep` (a value loaded after a barrier) has arrived: BAR.SYNC defers its blocking to the
// first dependent memory access, so a plain clock read after bar.sync does not see the wait
__device__ __forceinline__ long long gk_clock_after(const float *smem_word) {
    float dep; long long t;
    asm volatile("ld.volatile.shared.f32 %0, [%1];" : "=f"(dep) : "r"((unsigned)__cvta_generic_to_shared(smem_word)) : "memory");
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) : "f"(dep) : "memory");
    return t;
}
// byte k of a packed word as a float without the (quarter-rate) I2F: 0x4B000000 | b is the float 2^23 + b
__device__ __forceinline__ float gk_byte_to_float(unsigned w, int k) { return __uint_as_float(0x4B000000u | ((w >> (8 * k)) & 0xffu)) - 8388608.0f; }
__device__ __forceinline__ float gk_rsq_fast(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

template <int D> struct gk_seq5_dim {
    static constexpr int Q = gk_edim<D>::Q;
    static constexpr int BROW = (Q + D + 2 + 3) / 4 * 4;          // prec base | h base | x | coef, padded to float4
};

template <int D>
size_t gk_eig_seq5_smem() {
    constexpr int NB = 128, PA = gk_edim<D>::PA;
    return (size_t)2 * NB * PA * 4 + (size_t)NB * (gk_seq5_dim<D>::BROW > PA ? gk_seq5_dim<D>::BROW : PA) * 4 + (size_t)NB * ((PA + 3) / 4 * 4) * 4 + (size_t)NB * NB * 4 + (size_t)NB * NB + (size_t)3 * NB * D * 4 + 64;
}

// PROF = true compiles the %clock64 probes in (env GK_EIG_PROF); the production instantiation carries none of them
template <int D, bool PROF>
__global__ void __launch_bounds__(GK_SEQ5_THREADS, 1) k_eig_seq5(gk_eig_seq4_params P) {
    constexpr int NB = 128, Q = gk_edim<D>::Q, Pn = gk_edim<D>::P, PA = gk_edim<D>::PA, BROW = gk_seq5_dim<D>::BROW;
    constexpr int PAp = (PA + 3) / 4 * 4, NQ = PAp / 4, SR = (GK_SEQ5_HELP - 64) / NQ < GK_SEQ5_SRMAX ? (GK_SEQ5_HELP - 64) / NQ : GK_SEQ5_SRMAX;   // dPhi row pitch, column quads, summer row classes
    static_assert(PA + D + 2 <= 64, "one finisher thread per statistic column, per h entry, and two for x and coef");
    static_assert(SR >= 1 && SR <= GK_SEQ5_SRMAX, "summer row classes");
    constexpr int CBP = BROW > PA ? BROW : PA;                 // pitch of the base-row constants
    constexpr bool HOMO = (Pn % 4 == 0) && (PA % 4 == 0);      // column quads never straddle the S / A boundary (D = 8)
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float *Gb = (float *)smem_raw;                        // [NB][PA]  G rows at block entry (A-linear column: minus sum_m f_m + f_t)
    float *Gd = Gb + (size_t)NB * PA;                     // [NB][PA]  in-block changes sum_{s<j} X[s,j] dPhi_s
    float *dph = Gd + (size_t)NB * PA;                    // [NB][PAp] dPhi of the block's nodes (rows padded to float4)
    float *Cb = dph + (size_t)NB * PAp;                   // [NB][CBP] constants of every node's base row (built at block entry)
    float *Sb = Cb + (size_t)NB * CBP;                    // [NB][NB]
    float *zc = Sb + (size_t)NB * NB;                     // [NB][D]
    float *fnS = zc + (size_t)NB * D;                     // [NB][D]   new features
    float *foldC = fnS + (size_t)NB * D;                  // [NB][D]   features at block entry
    uint8_t *Ab = (uint8_t *)(foldC + (size_t)NB * D);    // [NB][NB]
    __shared__ __align__(16) float baseS[2][BROW];
    __shared__ __align__(16) float part[2 * GK_SEQ5_SRMAX * 64];     // [2][SR][PAp] partial sums of the summers (double-buffered by phase)
    __shared__ float lamS[D];
    __shared__ double sF0[D];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nb = P.nb, n0 = P.n0;
    const bool probit = P.probit != 0;
    const float mu0 = (float)P.mu0, isf = (float)P.inv_sigma_F;
    const int dbg = gk_seq5_debug;
    const int nh1 = 64;                                       // barrier 1: the solver and the finishers only

    // ---- block entry (everything before gk_pdl_wait() is independent of the preceding kernels in the stream)
    for (int idx = tid; idx < nb * (NB / 4); idx += GK_SEQ5_THREADS) {
        const int i = idx / (NB / 4), v = idx - i * (NB / 4);
        *(int4 *)(Sb + (size_t)i * NB + v * 4) = __ldg((const int4 *)((const float *)P.Sblk + (size_t)i * P.ldS + v * 4));
    }
    if (!probit)
        for (int idx = tid; idx < nb * (NB / 16); idx += GK_SEQ5_THREADS) {
            const int i = idx / (NB / 16), v = idx - i * (NB / 16);
            *(int4 *)(Ab + (size_t)i * NB + v * 16) = __ldg((const int4 *)(P.Ablk + (size_t)i * P.ldA + v * 16));
        }
    for (int t = tid; t < nb * D; t += GK_SEQ5_THREADS) {
        zc[t] = (float)P.z[(size_t)n0 * D + t];
        foldC[t] = (float)P.F[(size_t)n0 * D + t];
    }
    if (tid < D) lamS[tid] = (float)P.lmbda[tid];
    gk_pdl_wait();                                        // G and the running column sums come from the previous kernels
    if (tid < D) sF0[tid] = P.glob[GK_EIG_GLOB_SUMF + tid];
    __syncthreads();
    for (int t = tid; t < nb * PA; t += GK_SEQ5_THREADS) {
        const int i = t / PA, p = t - i * PA;
        double g = P.G[(size_t)n0 * PA + t];
        if (p >= Pn && !probit) g -= sF0[p - Pn] - P.F[(size_t)(n0 + i) * D + (p - Pn)];
        Gb[t] = (float)g; Gd[t] = 0.f;
        if (p == 0) { for (int e = PA; e < PAp; ++e) dph[i * PAp + e] = 0.f; }
    }
    __syncthreads();
    // constants of every node's base row: Cb[j] = [ k0 + k1 (Gb[j] - x_j Phi(old f_{j-1})) | lmbda (Gb_A - mu_0 Gb_S) - coef_j lmbda f_old_{j-1} | x_j | coef_j ]
    // with x_j = X_S[j-1, j], coef_j = X_A[j-1, j] - mu_0 x_j - 1 (zero for j = 0): everything in base_j that does not depend on the draws
    for (int t = tid; t < nb * BROW; t += GK_SEQ5_THREADS) {
        const int j = t / BROW, p = t - j * BROW;
        const float xs = j > 0 ? (probit ? 2.f : Sb[(j - 1) * NB + j]) : 0.f;
        const float xa = j > 0 ? (probit ? Sb[(j - 1) * NB + j] : (float)Ab[(j - 1) * NB + j]) : 0.f;
        const float coef = j > 0 ? xa - mu0 * xs - (probit ? 0.f : 1.f) : 0.f;
        float v = 0.f;
        if (p < Q) {
            int a, b; gk_col_ab<D>(p, a, b);
            const float old = j > 0 ? xs * foldC[(j - 1) * D + a] * foldC[(j - 1) * D + b] : 0.f;
            v = (a == b ? isf : 0.f) + lamS[a] * lamS[b] * (Gb[j * PA + p] - old);
        } else if (p < Q + D) {
            const int a = p - Q;
            const float fo = j > 0 ? foldC[(j - 1) * D + a] : 0.f;
            v = lamS[a] * (Gb[j * PA + Pn + a] - mu0 * Gb[j * PA + Q + a]) - coef * lamS[a] * fo;
        } else if (p == Q + D) v = xs;
        else if (p == Q + D + 1) v = coef;
        Cb[j * CBP + p] = v;
    }
    for (int t = tid; t < 2 * GK_SEQ5_SRMAX * 64; t += GK_SEQ5_THREADS) part[t] = 0.f;
    __syncthreads();
    if (tid < BROW) { baseS[0][tid] = Cb[tid]; if (nb > 1) baseS[1][tid] = Cb[CBP + tid]; }   // base rows of nodes 0 and 1 carry no in-block sums
    __syncthreads();

    if (warp == 0) {
        // ================= solver: every lane runs the same arithmetic =================
        float lam[D], f[D], br[BROW];
#pragma unroll
        for (int d = 0; d < D; ++d) { lam[d] = lamS[d]; f[d] = 0.f; }
#pragma unroll
        for (int q = 0; q < BROW / 4; ++q) { const float4 v = ((const float4 *)baseS[0])[q]; br[4 * q] = v.x; br[4 * q + 1] = v.y; br[4 * q + 2] = v.z; br[4 * q + 3] = v.w; }
        const bool sprof = PROF && P.prof != nullptr;
        long long pcw = 0, ck = 0, pa_ = 0, pb_ = 0, pf_ = 0, pr_ = 0, c0 = 0;
        const long long l0 = sprof ? gk_clock() : 0;
        for (int t = 0; t < nb; ++t) {
            const int n = n0 + t;
            if (sprof) c0 = gk_clock();
            float zz[D];
#pragma unroll
            for (int d = 0; d < D; ++d) zz[d] = zc[t * D + d];
            // conditional: base row + the previous node's new-feature half
            float a[Q], w[D], uu[D], vv[D];
            const float xs = br[Q + D], coef = br[Q + D + 1];
#pragma unroll
            for (int d = 0; d < D; ++d) { uu[d] = lam[d] * f[d]; vv[d] = xs * uu[d]; }
#pragma unroll
            for (int aa = 0; aa < D; ++aa) {
#pragma unroll
                for (int bb = aa; bb < D; ++bb) a[gk_qidx(D, aa, bb)] = br[gk_qidx(D, aa, bb)] + vv[aa] * uu[bb];
                w[aa] = br[Q + aa] + coef * uu[aa];
            }
            if (P.out_prec && lane == 0) {
#pragma unroll
                for (int aa = 0; aa < D; ++aa)
#pragma unroll
                    for (int bb = aa; bb < D; ++bb) {
                        P.out_prec[((size_t)n * D + aa) * D + bb] = (double)a[gk_qidx(D, aa, bb)];
                        P.out_prec[((size_t)n * D + bb) * D + aa] = (double)a[gk_qidx(D, aa, bb)];
                    }
            }
            if (P.out_h && lane == 0) {
#pragma unroll
                for (int aa = 0; aa < D; ++aa) P.out_h[(size_t)n * D + aa] = (double)w[aa];
            }
            // prec = L L^T (right-looking, forward substitution fused in)
            float rinv[D];
#pragma unroll
            for (int jj = 0; jj < D; ++jj) {
                const float r = gk_rsq_fast(a[gk_qidx(D, jj, jj)]);
                rinv[jj] = r;
                w[jj] *= r;
#pragma unroll
                for (int ii = jj + 1; ii < D; ++ii) {
                    a[gk_qidx(D, jj, ii)] *= r;
                    w[ii] -= a[gk_qidx(D, jj, ii)] * w[jj];
                }
#pragma unroll
                for (int kk = jj + 1; kk < D; ++kk)
#pragma unroll
                    for (int ii = kk; ii < D; ++ii) a[gk_qidx(D, kk, ii)] -= a[gk_qidx(D, jj, ii)] * a[gk_qidx(D, jj, kk)];
            }
            // next node's base row: ready by now (the helpers had the whole factorisation); its loads fly under the back substitution
            __syncwarp();
            if (sprof) { ck = gk_clock(); pa_ += ck - c0; }
            if (t >= 1) gk_bar_sync(2, 96);
            if (sprof) { c0 = gk_clock_after(baseS[0]); pcw += c0 - ck; }
            // (br is dead once a[] and w[] are built: the next row lands in the same registers; volatile loads stay here,
            //  right after the barrier, so that their latency is covered by the back substitution)
            if (t + 1 < nb) {
                const unsigned baddr = (unsigned)__cvta_generic_to_shared(baseS[(t + 1) & 1]);
#pragma unroll
                for (int q = 0; q < BROW / 4; ++q)
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(br[4 * q]), "=f"(br[4 * q + 1]), "=f"(br[4 * q + 2]), "=f"(br[4 * q + 3]) : "r"(baddr + 16 * q) : "memory");
            }
#pragma unroll
            for (int d = 0; d < D; ++d) w[d] += zz[d];
#pragma unroll
            for (int jj = D - 1; jj >= 0; --jj) {
                w[jj] *= rinv[jj];
#pragma unroll
                for (int kk = 0; kk < jj; ++kk) w[kk] -= a[gk_qidx(D, kk, jj)] * w[jj];
            }
#pragma unroll
            for (int d = 0; d < D; ++d) f[d] = w[d];
            if (lane == 0) {
#pragma unroll
                for (int d = 0; d < D; ++d) fnS[t * D + d] = f[d];
            }
            if (sprof) { ck = gk_clock(); pb_ += ck - c0; }
            // (bar.arrive / bar.sync order the prior shared-memory accesses of the participants: no fence needed)
            if (sprof) { c0 = gk_clock(); pf_ += c0 - ck; }
            __syncwarp();
            gk_bar_arrive(1, 32 + nh1);                    // f_t published
            if (sprof) { pr_ += gk_clock() - c0; }
        }
        if (sprof && lane == 0) { atomicAdd((unsigned long long *)P.prof + 0, (unsigned long long)(gk_clock() - l0)); atomicAdd((unsigned long long *)P.prof + 1, (unsigned long long)pcw);
            atomicAdd((unsigned long long *)P.prof + 2, (unsigned long long)pa_); atomicAdd((unsigned long long *)P.prof + 3, (unsigned long long)pb_);
            atomicAdd((unsigned long long *)P.prof + 4, (unsigned long long)pf_); atomicAdd((unsigned long long *)P.prof + 5, (unsigned long long)pr_); }
    } else if ((warp & 3) != 0) {
        // ================= helpers =================
        const int u = (warp - 1 - (warp >> 2)) * 32 + lane;       // 0..287
        if (u < 64) {
            // ---- finishers: thread p < PA owns statistic column p (dPhi, the in-block sums; for p < Q also prec entry p of the
            //      base row); thread PA + a owns h entry a of the base row (it needs columns Q+a and P+a and the running sum of
            //      the block's feature changes, all of which it can form itself: no exchange between finishers); two threads
            //      copy x and coef.  After f_s is published a finisher does: two loads of f, its partial sums, a few FMAs,
            //      one store -- then hands barrier 2 to the solver.
            const int p = u;
            const bool col = p < PA, isq = p < Q, ish = p >= PA && p < PA + D, isx = p >= PA + D && p < PA + D + 2;
            int pa = 0, pb = 0;
            if (col) gk_col_ab<D>(p, pa, pb);
            if (ish) pa = pb = p - PA;
            const int c1 = ish ? Q + pa : (col ? p : 0), c2 = Pn + pa;        // columns of the partial sums this thread reduces
            const int brow = ish ? Q + pa : (isx ? Q + D + (p - PA - D) : p); // entry of the base row it writes
            // X[s', j] of column c1 (and of c2 for the h threads) as wS * S + wA * A + w1: per-thread constants, no divergence
            const bool c1S = c1 < Pn;
            const float wS = (c1S && !probit) || (!c1S && probit) ? 1.f : 0.f, wA = (!c1S && !probit) ? 1.f : 0.f, w1 = (c1S && probit) ? 2.f : 0.f;
            const float vS = probit ? 1.f : 0.f, vA = probit ? 0.f : 1.f;    // column c2 is an A column
            float k1 = 0.f, lamh = 0.f;
            if (isq) { k1 = lamS[pa] * lamS[pb]; }
            if (ish) lamh = lamS[pa];
            float run = 0.f;
            long long pc0 = 0, pc1 = 0, ck = 0, pcA = 0, pcB = 0, pcC = 0, c1k = 0;
            const bool prof = PROF && P.prof != nullptr && u == 0;
            for (int s = 0; s < nb; ++s) {
                const int j = s + 2;
                const bool haveJ = j < nb;
                // ---- everything that does not depend on f_s is fetched before the wait
                const int jj = haveJ ? j : s;                       // (a valid address when there is no node s + 2)
                const float sx = Sb[s * NB + jj], ax = (float)Ab[s * NB + jj];
                const float xj1 = fmaf(wS, sx, fmaf(wA, ax, w1)), xj2 = fmaf(vS, sx, vA * ax);
                const int s1 = s + 1 < nb ? s + 1 : s;
                const float x1 = fmaf(wS, Sb[s * NB + s1], fmaf(wA, (float)Ab[s * NB + s1], w1));
                const float foa = foldC[s * D + pa], fob = foldC[s * D + pb];
                const float phs = isq ? foa * fob : foa;
                const float cpre = Cb[jj * CBP + brow];
                if (isx && haveJ) baseS[j & 1][brow] = cpre;       // (the solver finished with this buffer two phases ago)
                if (prof) ck = gk_clock();
                __syncwarp();
                gk_bar_sync(1, 32 + nh1);                // f_s published
                if (prof) { const long long c2k = gk_clock_after(fnS); pc0 += c2k - ck; ck = c2k; }
                const float fa = fnS[s * D + pa], fb = fnS[s * D + pb];
                float t1 = 0.f, t2 = 0.f;
                if (s >= 1 && !(dbg & 4)) gk_bar_sync(4, GK_SEQ5_HELP);      // the summers' sums of phase s - 1 are in shared memory
                if (haveJ && !(dbg & 2)) {
                    const float *pp = part + (size_t)(s & 1) * SR * PAp;
                    float u0 = 0.f, u1 = 0.f, w0 = 0.f, w1s = 0.f;
#pragma unroll
                    for (int q = 0; q + 1 < SR; q += 2) { u0 += pp[q * PAp + c1]; u1 += pp[(q + 1) * PAp + c1]; }
                    if (SR & 1) u0 += pp[(SR - 1) * PAp + c1];
                    if (ish) {
#pragma unroll
                        for (int q = 0; q + 1 < SR; q += 2) { w0 += pp[q * PAp + c2]; w1s += pp[(q + 1) * PAp + c2]; }
                        if (SR & 1) w0 += pp[(SR - 1) * PAp + c2];
                    }
                    t1 = u0 + u1; t2 = w0 + w1s;
                }
                const float d = isq ? fa * fb - phs : fa - phs;
                if (col) dph[s * PAp + p] = d;
                if (ish) run += d;
                t1 = fmaf(xj1, d, t1); t2 = fmaf(xj2, d, t2);
                if (haveJ) {
                    if (isq) baseS[j & 1][brow] = fmaf(k1, t1, cpre);
                    if (ish) baseS[j & 1][brow] = fmaf(lamh, t2 - mu0 * t1 - (probit ? 0.f : run), cpre);
                }
                // (bar.arrive / bar.sync order the prior shared-memory accesses of the participants: no fence needed)
                __syncwarp();
                if (s + 1 < nb) gk_bar_arrive(2, 96);             // base row of node s + 2 ready (and: f_s consumed, see the solver)
                if (!(dbg & 4)) gk_bar_arrive(3, GK_SEQ5_HELP);                   // dPhi_s is in shared memory: the summers may start
                if (prof) { c1k = gk_clock(); pcA += c1k - ck; }
                if (!(dbg & 8)) {
                if (col && haveJ) Gd[j * PA + p] = t1;
                if (col && s + 1 < nb) Gd[(s + 1) * PA + p] += x1 * d;
                }
                if (prof) { const long long c2k = gk_clock(); pc1 += c2k - ck; ck = c2k; }
            }
            if (prof) { atomicAdd((unsigned long long *)P.prof + 8, (unsigned long long)pc0); atomicAdd((unsigned long long *)P.prof + 9, (unsigned long long)pc1);
                atomicAdd((unsigned long long *)P.prof + 11, (unsigned long long)pcA); atomicAdd((unsigned long long *)P.prof + 12, (unsigned long long)pcB); atomicAdd((unsigned long long *)P.prof + 13, (unsigned long long)pcC); }
        } else if (u < GK_SEQ5_HELP) {
            // ---- summers: thread (quad q of four columns, row class r) accumulates sum over its rows s' <= s of X[s', j] dPhi_s'
            const int v = u - 64, q = v % NQ, r = v / NQ;
            const bool active = r < SR;
            float wS[4], wA[4], w1[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int p = 4 * q + e;
                const bool col = p < PA;
                wS[e] = ((p < Pn && !probit) || (p >= Pn && col && probit)) ? 1.f : 0.f;
                wA[e] = (p >= Pn && col && !probit) ? 1.f : 0.f;
                w1[e] = (p < Pn && probit) ? 2.f : 0.f;
            }
            long long pc2 = 0, ck = 0;
            const bool prof = PROF && P.prof != nullptr && v == 0;
            // Rows are owned in groups of four: class r holds the dPhi quads of rows 4 (r + k SR) + e, k < NG, e < 4, in REGISTERS
            // (loaded once, in the phase the row is drawn; zero until then), so a phase costs each thread one or two 16-byte loads
            // of X and no reload of dPhi -- the shared-memory pipe, which the solver's and the finishers' handshakes also go
            // through, stays nearly idle.
            constexpr int NG = (NB / 4 + SR - 1) / SR;
            float4 dc[NG][4];
#pragma unroll
            for (int k = 0; k < NG; ++k)
#pragma unroll
                for (int e = 0; e < 4; ++e) dc[k][e] = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int s = 0; s < ((dbg & 4) ? 0 : nb); ++s) {
                __syncwarp();
                gk_bar_sync(3, GK_SEQ5_HELP);                     // dPhi_s stored by the finishers
                if (prof) ck = gk_clock_after(dph);
                if (active) {
                    const int gs = s >> 2;
                    if (gs % SR == r) {                           // this class owns row s: take its quad into the register cache
                        const float4 dn = *(const float4 *)(dph + (size_t)s * PAp + 4 * q);
                        const int kg = gs / SR, e = s & 3;
#pragma unroll
                        for (int k = 0; k < NG; ++k)
#pragma unroll
                            for (int ee = 0; ee < 4; ++ee)
                                if (k == kg && ee == e) dc[k][ee] = dn;
                    }
                    const int jt = s + 3;                         // target of the NEXT phase; its terms s' <= s are all known now
                    if (jt < nb && !(dbg & 1)) {
                        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
                        const float *srow = Sb + (size_t)jt * NB;     // symmetric in-block tiles: X[s', jt] = X[jt, s'] (contiguous in s')
                        const uint8_t *arow = Ab + (size_t)jt * NB;
#pragma unroll
                        for (int k = 0; k < NG; ++k) {
                            const int g = r + k * SR;
                            if (4 * g <= s) {                     // rows not drawn yet hold zeros in the cache
                                const float4 xs4 = *(const float4 *)(srow + 4 * g);
                                const uchar4 xa4 = *(const uchar4 *)(arow + 4 * g);
                                const float xsv[4] = {xs4.x, xs4.y, xs4.z, xs4.w};
                                const float xav[4] = {(float)xa4.x, (float)xa4.y, (float)xa4.z, (float)xa4.w};
#pragma unroll
                                for (int e = 0; e < 4; ++e) {
                                    const float4 d4 = dc[k][e];
                                    if (HOMO) {                   // all four columns of a quad take the same X
                                        const float x = fmaf(wS[0], xsv[e], fmaf(wA[0], xav[e], w1[0]));
                                        a0 = fmaf(x, d4.x, a0); a1 = fmaf(x, d4.y, a1); a2 = fmaf(x, d4.z, a2); a3 = fmaf(x, d4.w, a3);
                                    } else {
                                        a0 = fmaf(fmaf(wS[0], xsv[e], fmaf(wA[0], xav[e], w1[0])), d4.x, a0);
                                        a1 = fmaf(fmaf(wS[1], xsv[e], fmaf(wA[1], xav[e], w1[1])), d4.y, a1);
                                        a2 = fmaf(fmaf(wS[2], xsv[e], fmaf(wA[2], xav[e], w1[2])), d4.z, a2);
                                        a3 = fmaf(fmaf(wS[3], xsv[e], fmaf(wA[3], xav[e], w1[3])), d4.w, a3);
                                    }
                                }
                            }
                        }
                        *(float4 *)(part + ((size_t)((s + 1) & 1) * SR + r) * PAp + 4 * q) = make_float4(a0, a1, a2, a3);
                    }
                }
                __syncwarp();
                if (s + 1 < nb) gk_bar_arrive(4, GK_SEQ5_HELP);    // sums for the next phase stored (bar.arrive orders the prior stores)
                if (prof) { pc2 += gk_clock() - ck; }
            }
            if (prof) atomicAdd((unsigned long long *)P.prof + 10, (unsigned long long)pc2);
        }
    }
    __syncthreads();
    // ---- block exit: dPhi for k_eig_push, G rows, features, column sums of F
    typedef float ACC;
    for (int t = tid; t < nb * PA; t += GK_SEQ5_THREADS) {
        ((ACC *)P.dphi)[t] = dph[(t / PA) * PAp + (t % PA)];
        P.G[(size_t)n0 * PA + t] += (double)Gd[t];
    }
    double *diff = (double *)Gb;                          // [nb][D] new - old (Gb is dead)
    for (int t = tid; t < nb * D; t += GK_SEQ5_THREADS) {
        const double fv = (double)fnS[t], o = P.F[(size_t)n0 * D + t];
        diff[t] = fv - o;
        P.F[(size_t)n0 * D + t] = fv;
    }
    __syncthreads();
    if (tid < D) {
        double s = sF0[tid];
        for (int i = 0; i < nb; ++i) s += diff[i * D + tid];
        P.glob[GK_EIG_GLOB_SUMF + tid] = s;
    }
}
