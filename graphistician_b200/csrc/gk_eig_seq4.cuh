// gk_eig_seq4.cuh -- the serial draws of one node block from the running G (one CTA), included by gk_eig_fast.cuh.
//
// Exactly the recursion of eigenmodel.py:884-908 inside a block: node t is drawn from N(prec^-1 h, prec^-1) with
//   prec[a,b] = d_ab/sigma_F + lmbda_a lmbda_b g[ab],   h[a] = lmbda_a (g[P+a] - (sum_m f_m - f_t)[a] - mu_0 g[Q+a]),
//   g = G[n0+t] (all partners as they were at block entry, from HBM) + Gd[t] (changes of the block's earlier nodes).
// Warp 0 = solver.  Its per-node chain is the bound of the whole sweep, so it is kept minimal: every lane owns up to
// four fixed columns (two quadratic, one S-linear, one A-linear) with their constants in registers; the factorisation
// inputs are stored in the factorisation type; node t's change reaches row t+1 through a register "carry" instead of
// shared memory.  Warps 1..15 = updaters: while the solver works on node t they add node t-1's change to rows >= t+1 of
// Gd, in the accumulation type U (fp32 with fp32 storage: the change multiplies fp32 S entries anyway), four rows in
// flight per thread, and prepare Phi(old f) of node t+1.  One barrier per node.
// On exit: G[n0+j] += Gd[j] and dphi[i] = Phi(new f_i) - Phi(old f_i) go to global memory for k_eig_push.
#pragma once

// factor prec = L L^T (packed upper, in the factorisation type) and draw f = L^-T (L^-1 h + z): right-looking
// factorisation with the forward substitution fused in, column-oriented back substitution.
template <typename CH, typename U, int D>
__device__ __forceinline__ void gk_chol_draw4(const CH *__restrict__ Lam, const CH *__restrict__ hv, const CH *__restrict__ z,
                                              double *__restrict__ foutD, U *__restrict__ foutU) {
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
    for (int d = 0; d < D; ++d) { foutD[d] = (double)w[d]; foutU[d] = (U)w[d]; }
}

struct gk_eig_seq4_params {
    int n0, nb, N;
    const void *Sblk; int ldS;           // S[n0+i][n0+j] = Sblk[i*ldS + j]
    const uint8_t *Ablk; int ldA;
    double *F; const double *z; const double *lmbda;
    double mu0, inv_sigma_F;
    double *G;                           // [npad][PA]
    void *dphi;                          // U [NB][PA]
    double *glob;                        // sumF at GK_EIG_GLOB_SUMF
    double *out_prec, *out_h;
    int probit;
    long long *prof;                     // optional cycle counters (debug)
};

template <typename T, typename CH, int D>
size_t gk_eig_seq4_smem() {
    typedef typename gk_eacc<T>::type U;
    constexpr int NB = gk_eig_nb<T>::NB, PA = gk_edim<D>::PA;
    return (size_t)NB * PA * sizeof(U) + (size_t)NB * (NB + 1) * sizeof(T) + (size_t)NB * (NB + 4) + (size_t)NB * D * (8 + sizeof(CH)) + 64;
}

template <typename T, typename CH, int D>
__global__ void __launch_bounds__(GK_EIG_SEQ_THREADS, 1) k_eig_seq4(gk_eig_seq4_params P) {
    typedef typename gk_eacc<T>::type U;
    constexpr int NB = gk_eig_nb<T>::NB, Q = gk_edim<D>::Q, Pn = gk_edim<D>::P, PA = gk_edim<D>::PA;
    constexpr int NUPD = GK_EIG_SEQ_THREADS - 32, RUP = NUPD / PA;       // updater rows in flight
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *fold = (double *)smem_raw;                    // [NB][D]   features at block entry
    U *Gd = (U *)(fold + (size_t)NB * D);                 // [NB][PA]  in-block changes
    T *Sb = (T *)(Gd + (size_t)NB * PA);                  // [NB][NB+1]
    CH *zc = (CH *)(Sb + (size_t)NB * (NB + 1));          // [NB][D]
    uint8_t *Ab = (uint8_t *)(zc + (size_t)NB * D);       // [NB][NB+4]
    __shared__ CH LamC[gk_edim<D>::Q], hvC[D];
    __shared__ double fnewD[D];
    __shared__ U fnewU[D], dph[2][gk_edim<D>::PA], phio[2][gk_edim<D>::PA];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nb = P.nb, n0 = P.n0;
    const bool probit = P.probit != 0;

    for (int t = tid; t < nb * PA; t += GK_EIG_SEQ_THREADS) Gd[t] = U(0);
    for (int t = tid; t < nb * D; t += GK_EIG_SEQ_THREADS) {
        zc[t] = (CH)P.z[(size_t)n0 * D + t];
        fold[t] = P.F[(size_t)n0 * D + t];
    }
    for (int t = tid; t < nb * nb; t += GK_EIG_SEQ_THREADS) {
        int i = t / nb, j = t - i * nb;
        if (j > i) {
            Sb[i * (NB + 1) + j] = ((const T *)P.Sblk)[(size_t)i * P.ldS + j];
            if (!probit) Ab[i * (NB + 4) + j] = P.Ablk[(size_t)i * P.ldA + j];
        }
    }
    // operands of an in-block pair (i < j): xs multiplies the S columns, xa the "A" columns
    auto xS = [&](int i, int j) -> U { return probit ? U(2) : (U)Sb[i * (NB + 1) + j]; };
    auto xA = [&](int i, int j) -> U { return probit ? (U)Sb[i * (NB + 1) + j] : (U)Ab[i * (NB + 4) + j]; };

    // ---- solver lane roles: quadratic columns cA = lane, cB = lane + 32; lanes 16..16+D-1 also own the two linear columns
    const bool hasA = lane < Q, hasB = lane + 32 < Q, lin = lane >= 16 && lane < 16 + D;
    const int cA = lane, cB = lane + 32, la = lane - 16, cS = Q + la, cH = Pn + la;
    int aA = 0, bA = 0, aB = 0, bB = 0;
    if (hasA) gk_col_ab<D>(cA, aA, bA);
    if (hasB) gk_col_ab<D>(cB, aB, bB);
    double lpA = 0, lpB = 0, dgA = 0, dgB = 0, lamL = 0, sumF = 0;
    if (warp == 0) {
        if (hasA) { lpA = P.lmbda[aA] * P.lmbda[bA]; dgA = aA == bA ? P.inv_sigma_F : 0.0; }
        if (hasB) { lpB = P.lmbda[aB] * P.lmbda[bB]; dgB = aB == bB ? P.inv_sigma_F : 0.0; }
        if (lin) { lamL = P.lmbda[la]; sumF = P.glob[GK_EIG_GLOB_SUMF + la]; }
    }
    U carA = 0, carB = 0, carS = 0, carH = 0;             // node t-1's change to row t (registers)
    double gA = 0, gB = 0, gS = 0, gH = 0;                // prefetched G[n0+t] of the owned columns
    if (warp == 0 && nb > 0) {
        const double *g = P.G + (size_t)n0 * PA;
        if (hasA) gA = g[cA];
        if (hasB) gB = g[cB];
        if (lin) { gS = g[cS]; gH = g[cH]; }
    }
    __syncthreads();                                      // fold is complete
    // ---- updater roles
    const int u = tid - 32, up = u >= 0 ? u % PA : 0, ur = u >= 0 ? u / PA : RUP;
    int ua = 0, ub = 0;
    if (u >= 0 && u < PA) { gk_col_ab<D>(up, ua, ub); phio[0][up] = (U)(up < Q ? fold[ua] * fold[ub] : fold[ua]); }
    __syncthreads();

    long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, c0 = 0, c1 = 0;
    const bool prof = P.prof != nullptr;
    for (int t = 0; t <= nb; ++t) {
        if (warp == 0) {
            if (t < nb) {
                if (prof) c0 = clock64();
                const int n = n0 + t, cur = t & 1;
                // ---- conditional of node t
                if (hasA) {
                    const double v = dgA + lpA * (gA + (double)(Gd[t * PA + cA] + carA));
                    LamC[cA] = (CH)v;
                    if (P.out_prec) { P.out_prec[((size_t)n * D + aA) * D + bA] = v; P.out_prec[((size_t)n * D + bA) * D + aA] = v; }
                }
                if (hasB) {
                    const double v = dgB + lpB * (gB + (double)(Gd[t * PA + cB] + carB));
                    LamC[cB] = (CH)v;
                    if (P.out_prec) { P.out_prec[((size_t)n * D + aB) * D + bB] = v; P.out_prec[((size_t)n * D + bB) * D + aB] = v; }
                }
                double fo = 0.0;
                if (lin) {
                    fo = fold[t * D + la];
                    const double s = gS + (double)(Gd[t * PA + cS] + carS), h = gH + (double)(Gd[t * PA + cH] + carH);
                    const double v = lamL * (h - (probit ? 0.0 : (sumF - fo)) - P.mu0 * s);
                    hvC[la] = (CH)v;
                    if (P.out_h) P.out_h[(size_t)n * D + la] = v;
                }
                __syncwarp();
                if (prof) { c1 = clock64(); pc[0] += c1 - c0; c0 = c1; }
                // ---- factor and draw (one lane: the chain is latency-bound)
                if (lane == 0) gk_chol_draw4<CH, U, D>(LamC, hvC, zc + t * D, fnewD, fnewU);
                // meanwhile the other lanes fold the consumed carry into Gd (final write-back) and prefetch G[n0+t+1]
                else {
                    if (hasA) Gd[t * PA + cA] += carA;
                    if (hasB) Gd[t * PA + cB] += carB;
                    if (lin) { Gd[t * PA + cS] += carS; Gd[t * PA + cH] += carH; }
                }
                if (lane == 0) { if (hasA) Gd[t * PA + cA] += carA; if (hasB) Gd[t * PA + cB] += carB; }
                if (t + 1 < nb) {
                    const double *g = P.G + (size_t)(n + 1) * PA;
                    if (hasA) gA = g[cA];
                    if (hasB) gB = g[cB];
                    if (lin) { gS = g[cS]; gH = g[cH]; }
                }
                __syncwarp();
                if (prof) { c1 = clock64(); pc[1] += c1 - c0; c0 = c1; }
                // ---- the node's change, and its carry into row t+1
                const bool nx = t + 1 < nb;
                const U xs = nx ? xS(t, t + 1) : U(0), xa = nx ? xA(t, t + 1) : U(0);
                if (hasA) {
                    const U d = fnewU[aA] * fnewU[bA] - phio[cur][cA];
                    dph[cur][cA] = d; ((U *)P.dphi)[t * PA + cA] = d; carA = xs * d;
                }
                if (hasB) {
                    const U d = fnewU[aB] * fnewU[bB] - phio[cur][cB];
                    dph[cur][cB] = d; ((U *)P.dphi)[t * PA + cB] = d; carB = xs * d;
                }
                if (lin) {
                    const double f = fnewD[la];
                    P.F[(size_t)n * D + la] = f; sumF += f - fo;
                    const U d = fnewU[la] - (U)fo;
                    dph[cur][cS] = d; dph[cur][cH] = d;
                    ((U *)P.dphi)[t * PA + cS] = d; ((U *)P.dphi)[t * PA + cH] = d;
                    carS = xs * d; carH = xa * d;
                }
                if (prof) { c1 = clock64(); pc[2] += c1 - c0; c0 = c1; }
            }
        } else if (ur < RUP) {
            if (prof) c0 = clock64();
            // Phi(old f) of node t+1 for the solver's next iteration
            if (u < PA && t + 1 < nb) phio[(t + 1) & 1][up] = (U)(up < Q ? fold[(t + 1) * D + ua] * fold[(t + 1) * D + ub] : fold[(t + 1) * D + ua]);
            if (t > 0) {
                const int s = t - 1;
                const U d = dph[s & 1][up];
                int j = t + 1 + ur;
                if (up < Pn) {
                    for (; j + 3 * RUP < nb; j += 4 * RUP) {
                        const U x0 = xS(s, j), x1 = xS(s, j + RUP), x2 = xS(s, j + 2 * RUP), x3 = xS(s, j + 3 * RUP);
                        U *g = Gd + (size_t)j * PA + up;
                        const U g0 = g[0], g1 = g[RUP * PA], g2 = g[2 * RUP * PA], g3 = g[3 * RUP * PA];
                        g[0] = g0 + x0 * d; g[RUP * PA] = g1 + x1 * d; g[2 * RUP * PA] = g2 + x2 * d; g[3 * RUP * PA] = g3 + x3 * d;
                    }
                    for (; j < nb; j += RUP) Gd[(size_t)j * PA + up] += xS(s, j) * d;
                } else {
                    for (; j + 3 * RUP < nb; j += 4 * RUP) {
                        const U x0 = xA(s, j), x1 = xA(s, j + RUP), x2 = xA(s, j + 2 * RUP), x3 = xA(s, j + 3 * RUP);
                        U *g = Gd + (size_t)j * PA + up;
                        const U g0 = g[0], g1 = g[RUP * PA], g2 = g[2 * RUP * PA], g3 = g[3 * RUP * PA];
                        g[0] = g0 + x0 * d; g[RUP * PA] = g1 + x1 * d; g[2 * RUP * PA] = g2 + x2 * d; g[3 * RUP * PA] = g3 + x3 * d;
                    }
                    for (; j < nb; j += RUP) Gd[(size_t)j * PA + up] += xA(s, j) * d;
                }
            }
            if (prof) { c1 = clock64(); pc[6] += c1 - c0; }
        }
        if (prof) c0 = clock64();
        __syncthreads();
        if (prof) { c1 = clock64(); pc[7] += c1 - c0; }
    }
    if (prof && (tid == 0 || tid == 32)) for (int k = 0; k < 8; ++k) atomicAdd((unsigned long long *)P.prof + (tid ? 8 : 0) + k, (unsigned long long)pc[k]);
    if (warp == 0 && lin) P.glob[GK_EIG_GLOB_SUMF + la] = sumF;
    for (int t = tid; t < nb * PA; t += GK_EIG_SEQ_THREADS) P.G[(size_t)n0 * PA + t] += (double)Gd[t];
}
