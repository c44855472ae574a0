// gk_eig_seq4.cuh -- the serial draws of one node block from the running G (one CTA), included by gk_eig_fast.cuh.
//
// Exactly the recursion of eigenmodel.py:884-908 inside a block: node t is drawn from N(prec^-1 h, prec^-1) with
//   prec[a,b] = d_ab/sigma_F + lmbda_a lmbda_b g[ab],   h[a] = lmbda_a (g[P+a] - (sum_m f_m - f_t)[a] - mu_0 g[Q+a]),
//   g = G[n0+t] (all partners as they were at block entry, from HBM) + Gd[t] (changes of the block's earlier nodes).
// Warp 0 = solver, alone on its SM sub-partition (warps 4, 8, 12 idle).  Its per-node chain bounds the whole sweep,
// so it carries no fp64 when storage and factorisation are fp32 (compute type C = float): the block's rows of G are
// converted once at kernel entry (Gb), the "- (sum_m f_m - f_t)" term of h is folded into the A-linear column of Gb
// plus a running sum of the block's own changes, every lane owns up to three fixed columns with their constants in
// registers, and node t's change reaches row t+1 through a register "carry".  The other 12 warps = updaters: while the
// solver works on node t they add node t-1's change to rows >= t+1 of Gd (four rows in flight per thread) and prepare
// Phi(old f) of node t+1.  One barrier per node.
// On exit: G[n0+j] += Gd[j], F <- new features, sumF, and dphi[i] = Phi(new f_i) - Phi(old f_i) for k_eig_push.
#pragma once

// cycle counter the compiler may not move across memory operations or barriers (debug probes, env GK_EIG_PROF).
// NB: on B200 %clock64 advanced ~0.58 ticks per SM cycle in these kernels (ncu sm__cycles_elapsed vs these counters), so the
// probes give the SHARE of each phase, not SM cycles; each read also costs tens of cycles.
// programmatic dependent launch (sm_90+): no-ops unless the kernel was launched with the stream-serialization attribute
__device__ __forceinline__ void gk_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void gk_pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ long long gk_clock() { long long t; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory"); return t; }

// factor prec = L L^T (packed upper, in the factorisation type) and draw f = L^-T (L^-1 h + z): right-looking
// factorisation with the forward substitution fused in, column-oriented back substitution.
template <typename CH, int D>
__device__ __forceinline__ void gk_chol_draw4(const CH *__restrict__ Lam, const CH *__restrict__ hv, const CH *__restrict__ z,
                                              CH *__restrict__ fout) {
    constexpr int Q = gk_edim<D>::Q;
    CH a[Q], rinv[D], w[D], zz[D];
#pragma unroll
    for (int p = 0; p < Q; ++p) a[p] = Lam[p];
#pragma unroll
    for (int d = 0; d < D; ++d) { w[d] = hv[d]; zz[d] = z[d]; }
#pragma unroll
    for (int jj = 0; jj < D; ++jj) {
        CH r;
        if constexpr (sizeof(CH) == 4) r = rsqrtf(a[gk_qidx(D, jj, jj)]); else r = rsqrt(a[gk_qidx(D, jj, jj)]);
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
#pragma unroll
    for (int d = 0; d < D; ++d) w[d] += zz[d];
#pragma unroll
    for (int jj = D - 1; jj >= 0; --jj) {
        w[jj] *= rinv[jj];
#pragma unroll
        for (int kk = 0; kk < jj; ++kk) w[kk] -= a[gk_qidx(D, kk, jj)] * w[jj];
    }
#pragma unroll
    for (int d = 0; d < D; ++d) fout[d] = w[d];
}

struct gk_eig_seq4_params {
    int n0, nb, N;
    const void *Sblk; int ldS;           // S[n0+i][n0+j] = Sblk[i*ldS + j]
    const uint8_t *Ablk; int ldA;
    double *F; const double *z; const double *lmbda;
    double mu0, inv_sigma_F;
    double *G;                           // [npad][PA]
    void *dphi;                          // gk_eacc<T> [NB][PA]
    double *glob;                        // sumF at GK_EIG_GLOB_SUMF
    double *out_prec, *out_h;
    int probit;
    long long *prof;                     // optional cycle counters (debug)
};

template <typename T, typename CH> struct gk_ctype { typedef double type; };
template <> struct gk_ctype<float, float> { typedef float type; };

template <typename T, typename CH, int D>
size_t gk_eig_seq4_smem() {
    typedef typename gk_ctype<T, CH>::type C;
    constexpr int NB = gk_eig_nb<T>::NB, PA = gk_edim<D>::PA;
    return (size_t)2 * NB * PA * sizeof(C) + (size_t)NB * NB * sizeof(T) + (size_t)NB * D * (2 * sizeof(CH) + sizeof(C)) +
           (size_t)NB * NB + 64;
}

#define GK_SEQ4_SBK 16          // nodes per sub-block: eager in-block updates stay inside it, later rows are flushed per sub-block
#define GK_SEQ4_UPD_THREADS (GK_EIG_SEQ_THREADS - (GK_EIG_SEQ_THREADS / 128) * 32)     // warps with (warp & 3) != 0

template <typename T, typename CH, int D>
__global__ void __launch_bounds__(GK_EIG_SEQ_THREADS, 1) k_eig_seq4(gk_eig_seq4_params P) {
    typedef typename gk_ctype<T, CH>::type C;
    typedef typename gk_eacc<T>::type ACC;
    constexpr int NB = gk_eig_nb<T>::NB, Q = gk_edim<D>::Q, Pn = gk_edim<D>::P, PA = gk_edim<D>::PA;
    constexpr int RUP = GK_SEQ4_UPD_THREADS / PA;         // updater rows in flight
    constexpr int VT = 16 / sizeof(T);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    C *Gb = (C *)smem_raw;                                // [NB][PA]  G rows at block entry (A-linear column: minus sum_m f_m + f_t)
    C *Gd = Gb + (size_t)NB * PA;                         // [NB][PA]  changes made by the block's earlier nodes
    T *Sb = (T *)(Gd + (size_t)NB * PA);                  // [NB][NB]
    CH *zc = (CH *)(Sb + (size_t)NB * NB);                // [NB][D]
    CH *fnS = zc + (size_t)NB * D;                        // [NB][D]   new features
    C *foldC = (C *)(fnS + (size_t)NB * D);               // [NB][D]   features at block entry
    uint8_t *Ab = (uint8_t *)(foldC + (size_t)NB * D);    // [NB][NB]
    __shared__ CH LamC[gk_edim<D>::Q], hvC[D];
    __shared__ C dph[GK_SEQ4_SBK][gk_edim<D>::PA], phio[2][gk_edim<D>::PA];     // dph: changes of the current sub-block's nodes
    __shared__ double sF0[D];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nb = P.nb, n0 = P.n0;
    const bool probit = P.probit != 0;
    const long long kstart = P.prof ? gk_clock() : 0;

    // ---- block entry: in-block tiles of S / Asym (16-byte loads), G rows, features, normals
    //      (everything up to gk_pdl_wait() is independent of the kernels just before this one in the stream, so with a
    //       programmatic launch it runs while k_eig_push_cols is still pushing the previous block's change)
    for (int idx = tid; idx < nb * (NB / VT); idx += GK_EIG_SEQ_THREADS) {
        const int i = idx / (NB / VT), v = idx - i * (NB / VT);
        *(int4 *)(Sb + (size_t)i * NB + v * VT) = __ldg((const int4 *)((const T *)P.Sblk + (size_t)i * P.ldS + v * VT));
    }
    if (!probit)
        for (int idx = tid; idx < nb * (NB / 16); idx += GK_EIG_SEQ_THREADS) {
            const int i = idx / (NB / 16), v = idx - i * (NB / 16);
            *(int4 *)(Ab + (size_t)i * NB + v * 16) = __ldg((const int4 *)(P.Ablk + (size_t)i * P.ldA + v * 16));
        }
    for (int t = tid; t < nb * D; t += GK_EIG_SEQ_THREADS) {
        zc[t] = (CH)P.z[(size_t)n0 * D + t];
        foldC[t] = (C)P.F[(size_t)n0 * D + t];
    }
    gk_pdl_wait();                                        // G and the running column sums come from the previous kernels
    if (tid < D) sF0[tid] = P.glob[GK_EIG_GLOB_SUMF + tid];
    __syncthreads();
    for (int t = tid; t < nb * PA; t += GK_EIG_SEQ_THREADS) {
        const int i = t / PA, p = t - i * PA;
        double g = P.G[(size_t)n0 * PA + t];
        if (p >= Pn && !probit) g -= sF0[p - Pn] - P.F[(size_t)(n0 + i) * D + (p - Pn)];
        Gb[t] = (C)g; Gd[t] = C(0);
    }
    // operands of an in-block pair (i < j): xS multiplies the S columns, xA the "A" columns
    auto xS = [&](int i, int j) -> C { return probit ? C(2) : (C)Sb[i * NB + j]; };
    auto xA = [&](int i, int j) -> C { return probit ? (C)Sb[i * NB + j] : (C)Ab[i * NB + j]; };

    // ---- solver lane roles: slot 0 = quadratic column lane; slot 1 = quadratic column lane+32, or (lanes 16..16+D-1) the
    //      S-linear column of a = lane-16; slot 2 = the A-linear column of the same a
    const bool lin = lane >= 16 && lane < 16 + D, has0 = lane < Q, hasB = lane + 32 < Q, has1 = hasB || lin;
    const int la = lin ? lane - 16 : 0, c0 = lane, c1 = lin ? Q + la : lane + 32, c2 = Pn + la;
    int a0 = 0, b0 = 0, a1 = 0, b1 = 0;
    if (has0) gk_col_ab<D>(c0, a0, b0);
    if (hasB) gk_col_ab<D>(c1, a1, b1);
    C k00 = 0, k01 = 0, k10 = 0, k11 = 0, lamL = 0;
    if (warp == 0) {
        if (has0) { k01 = (C)(P.lmbda[a0] * P.lmbda[b0]); k00 = a0 == b0 ? (C)P.inv_sigma_F : C(0); }
        if (hasB) { k11 = (C)(P.lmbda[a1] * P.lmbda[b1]); k10 = a1 == b1 ? (C)P.inv_sigma_F : C(0); }
        if (lin) { lamL = (C)P.lmbda[la]; k11 = (C)(-P.lmbda[la] * P.mu0); a1 = b1 = la; }
    }
    C car0 = 0, car1 = 0, car2 = 0, runS = 0;             // node t-1's change to row t; running sum of the block's own f changes
    // ---- updater roles
    const bool is_upd = (warp & 3) != 0;
    const int u = is_upd ? (warp - 1 - (warp >> 2)) * 32 + lane : -1;
    const int up = u >= 0 ? u % PA : 0, ur = u >= 0 ? u / PA : RUP;
    int ua = 0, ub = 0;
    if (u >= 0 && u < PA) { gk_col_ab<D>(up, ua, ub); phio[0][up] = up < Q ? foldC[ua] * foldC[ub] : foldC[ua]; }
    __syncthreads();

    long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, ck0 = 0, ck1 = 0;
    const bool prof = P.prof != nullptr;
    const long long lstart = prof ? gk_clock() : 0;
    for (int t = 0; t <= nb; ++t) {
        if (warp == 0) {
            if (t < nb) {
                if (prof) ck0 = gk_clock();
                const int n = n0 + t, cur = t & 1, row = t * PA;
                // ---- conditional of node t (no fp64 when C == float); all loads first, then arithmetic, then stores
                const C gd0 = has0 ? Gd[row + c0] : C(0), gb0 = has0 ? Gb[row + c0] : C(0);
                const C gd1 = has1 ? Gd[row + c1] : C(0), gb1 = has1 ? Gb[row + c1] : C(0);
                const C gd2 = lin ? Gd[row + c2] : C(0), gb2 = lin ? Gb[row + c2] : C(0);
                const C e0 = gd0 + car0, e1 = gd1 + car1, e2 = gd2 + car2;
                const C y0 = k00 + k01 * (gb0 + e0);
                C y1 = k10 + k11 * (gb1 + e1);
                if (lin) y1 += lamL * (gb2 + e2 - (probit ? C(0) : runS));
                if (has0) { LamC[c0] = (CH)y0; Gd[row + c0] = e0; }
                if (has1) {
                    Gd[row + c1] = e1;
                    if (lin) { hvC[la] = (CH)y1; Gd[row + c2] = e2; } else LamC[c1] = (CH)y1;
                }
                if (P.out_prec) {
                    if (has0) { P.out_prec[((size_t)n * D + a0) * D + b0] = (double)y0; P.out_prec[((size_t)n * D + b0) * D + a0] = (double)y0; }
                    if (hasB) { P.out_prec[((size_t)n * D + a1) * D + b1] = (double)y1; P.out_prec[((size_t)n * D + b1) * D + a1] = (double)y1; }
                }
                if (P.out_h && lin) P.out_h[(size_t)n * D + la] = (double)y1;
                __syncwarp();
                // ---- factor and draw (one lane: the chain is latency-bound)
                if (lane == 0) gk_chol_draw4<CH, D>(LamC, hvC, zc + t * D, fnS + t * D);
                __syncwarp();
                // ---- the node's change, and its carry into row t+1 (loads first)
                const bool nx = t + 1 < nb;
                const CH *fn = fnS + t * D;
                const C xs = nx ? xS(t, t + 1) : C(0), xa = nx ? xA(t, t + 1) : C(0);
                const C f0a = (C)fn[a0], f0b = (C)fn[b0], f1a = (C)fn[a1], f1b = (C)fn[b1];
                const C ph0 = has0 ? phio[cur][c0] : C(0), ph1 = hasB ? phio[cur][c1] : C(0), fo = lin ? foldC[t * D + la] : C(0);
                const C d0 = f0a * f0b - ph0;
                const C d1 = lin ? f1a - fo : f1a * f1b - ph1;
                car0 = xs * d0; car1 = xs * d1;
                if (lin) { car2 = xa * d1; runS += d1; }
                if (has0) { dph[t % GK_SEQ4_SBK][c0] = d0; ((ACC *)P.dphi)[row + c0] = (ACC)d0; }
                if (has1) { dph[t % GK_SEQ4_SBK][c1] = d1; ((ACC *)P.dphi)[row + c1] = (ACC)d1; }
                if (lin) { dph[t % GK_SEQ4_SBK][c2] = d1; ((ACC *)P.dphi)[row + c2] = (ACC)d1; }
            }
        } else if (ur < RUP) {
            if (prof) ck0 = gk_clock();
            // Phi(old f) of node t+1 for the solver's next iteration
            if (u < PA && t + 1 < nb) phio[(t + 1) & 1][up] = up < Q ? foldC[(t + 1) * D + ua] * foldC[(t + 1) * D + ub] : foldC[(t + 1) * D + ua];
            if (t > 0) {
                // eager part: node s = t-1 reaches the later rows of its own sub-block (row s+1 is the solver's carry)
                const int s = t - 1, send = min(nb, (s / GK_SEQ4_SBK + 1) * GK_SEQ4_SBK);
                const C d = dph[s % GK_SEQ4_SBK][up];
                if (up < Pn) { for (int j = s + 2 + ur; j < send; j += RUP) Gd[(size_t)j * PA + up] += xS(s, j) * d; }
                else { for (int j = s + 2 + ur; j < send; j += RUP) Gd[(size_t)j * PA + up] += xA(s, j) * d; }
            }
        }
        if (prof) { ck1 = gk_clock(); pc[0] += ck1 - ck0; ck0 = ck1; }
        __syncthreads();
        if (prof) { ck1 = gk_clock(); pc[7] += ck1 - ck0; }
        // flush: the sub-block [t+1-SBK, t] is complete -> every later row takes its 16 changes at once (registers hold the
        // column's changes; one read-modify-write of Gd per 16 FMAs).  The pair (t, t+1) stays with the solver's carry.
        if (t < nb && (t + 1) % GK_SEQ4_SBK == 0 && t + 1 < nb) {
            if (prof) ck0 = gk_clock();
            constexpr int FR = GK_EIG_SEQ_THREADS / PA;
            const int fc = tid % PA, fr = tid / PA, s0 = t + 1 - GK_SEQ4_SBK;
            if (fr < FR) {
                C dd[GK_SEQ4_SBK];
#pragma unroll
                for (int k = 0; k < GK_SEQ4_SBK; ++k) dd[k] = dph[k][fc];
                for (int j = t + 1 + fr; j < nb; j += FR) {
                    C acc = C(0);
                    if (fc < Pn) {
#pragma unroll
                        for (int k = 0; k < GK_SEQ4_SBK; ++k) acc += xS(s0 + k, j) * dd[k];
                        if (j == t + 1) acc -= xS(t, j) * dd[GK_SEQ4_SBK - 1];
                    } else {
#pragma unroll
                        for (int k = 0; k < GK_SEQ4_SBK; ++k) acc += xA(s0 + k, j) * dd[k];
                        if (j == t + 1) acc -= xA(t, j) * dd[GK_SEQ4_SBK - 1];
                    }
                    Gd[(size_t)j * PA + fc] += acc;
                }
            }
            __syncthreads();
            if (prof) { ck1 = gk_clock(); pc[5] += ck1 - ck0; }
        }
    }
    if (prof) { pc[3] = gk_clock() - lstart; pc[4] = lstart - kstart; }
    if (prof && (tid == 0 || tid == 63)) for (int k = 0; k < 8; ++k) atomicAdd((unsigned long long *)P.prof + (tid ? 8 : 0) + k, (unsigned long long)pc[k]);
    // ---- block exit: G rows, features, column sums of F
    for (int t = tid; t < nb * PA; t += GK_EIG_SEQ_THREADS) P.G[(size_t)n0 * PA + t] += (double)Gd[t];
    double *diff = (double *)Gb;                          // [nb][D] new - old (Gb is dead)
    for (int t = tid; t < nb * D; t += GK_EIG_SEQ_THREADS) {
        const double f = (double)fnS[t], o = P.F[(size_t)n0 * D + t];
        diff[t] = f - o;
        P.F[(size_t)n0 * D + t] = f;
    }
    __syncthreads();
    if (tid < D) {
        double s = sF0[tid];
        for (int i = 0; i < nb; ++i) s += diff[i * D + tid];
        P.glob[GK_EIG_GLOB_SUMF + tid] = s;
    }
    if (prof && tid == 0) atomicAdd((unsigned long long *)P.prof + 6, (unsigned long long)(gk_clock() - kstart));
}
