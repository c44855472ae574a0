// gk_eig_mf.cuh -- device kernels of the mean-field logistic eigenmodel (K11), included by gk_eig.cu.
//
// Replaces, in upstream eigenmodel.py:
//   :1055-1064 _meanfieldupdate_Omega : E[omega_nm] = 1/2 mean_s tanh(z_s/2)/z_s, z_s ~ N(E[mu_nm], Var[mu_nm]), N_samples
//              activations PER ELEMENT (a 1000 x N x N array in the reference), with E[mu], E[mu^2] of :1177-1224 fused in
//   :1090-1124 _meanfieldupdate_F     : the serial per-node update (k_eig_mf_seq) on the running-G machinery of gk_eig_fast.cuh
//   :1066-1088, :1126-1157            : the N^2 sums of the mu0 / lmbda updates (tile partials here, k_eig_gstats_phi)
#pragma once
#include "gk_eig_fast.cuh"

#define GK_MFE_TILE 64

// Phi (double [N][PA]) holds per node: E[f f^T] packed (a <= b) | E[f] | E[f]  -- the layout the contraction kernels use.
// One CTA = one 64 x 64 tile with I <= J of the symmetric S = E_O + E_O^T; writes both orientations.
//   E[mu_nm]   = mu0 + sum_a lam_a f_na f_ma
//   E[mu_nm^2] = mu0sq + 2 mu0 (E[mu_nm] - mu0) + sum_{a<=b} wq_ab Phi_n[ab] Phi_m[ab]     wq = (2 - d_ab) E[lmbda lmbda^T]_ab
// part[0][tile] = sum of E_O over the tile's directed entries (diagonal entries once), part[1][tile] = sum over its
// unordered pairs of S_nm (E[mu_nm] - mu0)  (= sum over directed off-diagonal entries of E_O * E_FdotF, eigenmodel.py:1083).
template <typename T, int D>
__global__ void __launch_bounds__(256) k_eig_mf_omega(T *__restrict__ S, int ld, int N, const double *__restrict__ Phi,
                                                      const double *__restrict__ wq, const double *__restrict__ lam, double mu0,
                                                      double sig_mu0, int n_samples, uint64_t seed, double *__restrict__ part) {
    constexpr int TS = GK_MFE_TILE, Q = gk_edim<D>::Q, PA = gk_edim<D>::PA;
    __shared__ float qi[TS][Q + 1], qj[TS][Q + 1], fi[TS][D], fj[TS][D];
    __shared__ float tile[TS][TS + 1];
    __shared__ double red[2][8];
    const int tid = threadIdx.x, I = blockIdx.y * TS, J = blockIdx.x * TS;
    const int pidx = blockIdx.y * gridDim.x + blockIdx.x;
    if (J < I) { if (tid == 0) { part[pidx] = 0.0; part[gridDim.x * gridDim.y + pidx] = 0.0; } return; }
    const bool diag_tile = I == J;
    for (int t = tid; t < TS * Q; t += 256) {
        const int r = t / Q, p = t - r * Q;
        qi[r][p] = I + r < N ? (float)(Phi[(size_t)(I + r) * PA + p] * wq[p]) : 0.f;
        qj[r][p] = J + r < N ? (float)Phi[(size_t)(J + r) * PA + p] : 0.f;
    }
    for (int t = tid; t < TS * D; t += 256) {
        const int r = t / D, d = t - r * D;
        fi[r][d] = I + r < N ? (float)(Phi[(size_t)(I + r) * PA + Q + d] * lam[d]) : 0.f;
        fj[r][d] = J + r < N ? (float)Phi[(size_t)(J + r) * PA + Q + d] : 0.f;
    }
    __syncthreads();
    gk_xoshiro g; g.seed(seed, (uint32_t)tid, (uint32_t)I, (uint32_t)J);
    const int j = tid & (TS - 1), i0 = tid / TS;
    const float m0 = (float)mu0, m0sq = (float)(sig_mu0 + mu0 * mu0), inv_ns = 1.0f / (float)n_samples;
    double sEO = 0.0, sEOF = 0.0;
    for (int i = i0; i < TS; i += 256 / TS) {
        const int n = I + i, m = J + j;
        float out = 0.f;
        if (n < N && m < N && (!diag_tile || m >= n)) {
            float dot = 0.f, quad = 0.f;
#pragma unroll
            for (int d = 0; d < D; ++d) dot += fi[i][d] * fj[j][d];
#pragma unroll
            for (int p = 0; p < Q; ++p) quad += qi[i][p] * qj[j][p];
            const bool dg = n == m;
            const float Em = dg ? m0 : m0 + dot;
            const float var = dg ? (float)sig_mu0 : fmaxf(m0sq + 2.f * m0 * dot + quad - Em * Em, 0.f);
            const float sd = sqrtf(var);
            const int ns = dg ? n_samples : 2 * n_samples;       // both directed entries of an off-diagonal pair
            float acc = 0.f;
            for (int s = 0; s < ns; s += 2) {
                // Box-Muller pair
                const float u1 = g.unif(), u2 = g.unif();
                const float rr = sqrtf(-2.0f * __logf(u1));
                float sn, cs; __sincosf(2.0f * GK_PI_F * u2, &sn, &cs);
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    if (s + h < ns) {
                        const float z = Em + sd * rr * (h ? sn : cs), az = fabsf(z);
                        // tanh(z/2)/z = (1 - e^-|z|) / (|z| (1 + e^-|z|)); series near 0
                        const float e = __expf(-az);
                        acc += az < 1e-3f ? 0.5f - az * az * (1.0f / 24.0f) : __fdividef(1.0f - e, az * (1.0f + e));
                    }
                }
            }
            out = 0.5f * acc * inv_ns;                           // E_O[n,m] + E_O[m,n]  (diagonal: E_O[n,n])
            sEO += (double)out;
            if (!dg) sEOF += (double)(out * dot);
            if (dg) out = 0.f;
        }
        tile[i][j] = out;
    }
    __syncthreads();
    for (int i = i0; i < TS; i += 256 / TS) {
        const int n = I + i, m = J + j;
        if (n < N && m < ld) {
            const float v = (diag_tile && j < i) ? tile[j][i] : tile[i][j];
            S[(size_t)n * ld + m] = m < N ? (T)v : T(0);
        }
        if (!diag_tile) {
            const int n2 = J + i, m2 = I + j;
            if (n2 < N && m2 < ld) S[(size_t)n2 * ld + m2] = m2 < N ? (T)tile[j][i] : T(0);
        }
    }
    for (int o = 16; o; o >>= 1) { sEO += __shfl_xor_sync(0xffffffffu, sEO, o); sEOF += __shfl_xor_sync(0xffffffffu, sEOF, o); }
    if ((tid & 31) == 0) { red[0][tid >> 5] = sEO; red[1][tid >> 5] = sEOF; }
    __syncthreads();
    if (tid == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < 8; ++w) { a += red[0][w]; b += red[1][w]; }
        part[pidx] = a; part[gridDim.x * gridDim.y + pidx] = b;
    }
}

// Phi[n] = [Sigma_n[ab] + mu_na mu_nb (a <= b) | lin_n | lin_n],  lin = Flin (or mu when Flin == nullptr)
template <int D>
__global__ void k_eig_mf_phi(const double *__restrict__ mu, const double *__restrict__ Sig, const double *__restrict__ Flin, int N,
                             double *__restrict__ Phi) {
    constexpr int Q = gk_edim<D>::Q, Pn = gk_edim<D>::P, PA = gk_edim<D>::PA;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < (size_t)N * PA; i += (size_t)gridDim.x * blockDim.x) {
        const int p = (int)(i % PA); const size_t n = i / PA;
        int a, b; gk_col_ab<D>(p, a, b);
        double v;
        if (p < Q) v = Sig[(n * D + a) * D + b] + mu[n * D + a] * mu[n * D + b];
        else v = Flin ? Flin[n * D + a] : mu[n * D + a];
        (void)Pn;
        Phi[i] = v;
    }
}

struct gk_eig_mfseq_params {
    int n0, nb, N;
    const void *Sblk; int ldS;
    const uint8_t *Ablk; int ldA;
    double *G, *Phi, *mu, *Sig;          // [npad][PA], [N][PA], [N][D], [N][D][D]  (Phi / mu / Sig updated in place)
    const double *llT, *lam;             // E[lmbda lmbda^T] (D x D), E[lmbda] (D)
    double mu0, inv_sigma_F, stepsize;
    int track;                           // 1: the linear columns follow mu (h built from mf_mu_F); 0: they stay at Flin (reference)
    void *dphi;                          // gk_eacc<T> [NB][PA]
    double *glob;                        // running column sums of the linear features at GK_EIG_GLOB_SUMF
    double *out_prec, *out_h;
    int probit;                          // 1: S == 2 and the S buffer holds E[Z] + E[Z]^T (eigenmodel.py:587-650)
};

template <typename T, int D>
size_t gk_eig_mfseq_smem() {
    constexpr int NB = gk_eig_nb<T>::NB, PA = gk_edim<D>::PA;
    return (size_t)NB * PA * 8 + (size_t)NB * NB * sizeof(T) + (size_t)NB * NB + 64;
}

// Serial mean-field update of one node block (eigenmodel.py:1090-1124): node t's q(f_t) = N(Sigma h, Sigma),
// Sigma = blend(prec^-1), prec = I/sigma_F + E[ll^T] o sum_m S_tm E[f_m f_m^T] with the CURRENT q of every other node.
template <typename T, int D>
__global__ void __launch_bounds__(256, 1) k_eig_mf_seq(gk_eig_mfseq_params P) {
    typedef typename gk_eacc<T>::type ACC;
    constexpr int NB = gk_eig_nb<T>::NB, Q = gk_edim<D>::Q, Pn = gk_edim<D>::P, PA = gk_edim<D>::PA, VT = 16 / sizeof(T);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *Gd = (double *)smem_raw;                      // [NB][PA]
    T *Sb = (T *)(Gd + (size_t)NB * PA);                  // [NB][NB]
    uint8_t *Ab = (uint8_t *)(Sb + (size_t)NB * NB);      // [NB][NB]
    __shared__ double gs[gk_edim<D>::PA], dps[gk_edim<D>::PA], L[D][D], Sn[D][D], hv[D], mun[D], sumF[D];
    const int tid = threadIdx.x, nb = P.nb, n0 = P.n0;
    for (int idx = tid; idx < nb * (NB / VT); idx += 256) {
        const int i = idx / (NB / VT), v = idx - i * (NB / VT);
        *(int4 *)(Sb + (size_t)i * NB + v * VT) = __ldg((const int4 *)((const T *)P.Sblk + (size_t)i * P.ldS + v * VT));
    }
    const bool probit = P.probit != 0;
    if (!probit)
        for (int idx = tid; idx < nb * (NB / 16); idx += 256) {
            const int i = idx / (NB / 16), v = idx - i * (NB / 16);
            *(int4 *)(Ab + (size_t)i * NB + v * 16) = __ldg((const int4 *)(P.Ablk + (size_t)i * P.ldA + v * 16));
        }
    for (int t = tid; t < nb * PA; t += 256) Gd[t] = 0.0;
    if (tid < D) sumF[tid] = P.glob[GK_EIG_GLOB_SUMF + tid];
    int pa = 0, pb = 0;
    if (tid < PA) gk_col_ab<D>(tid, pa, pb);
    __syncthreads();
    for (int t = 0; t < nb; ++t) {
        const int n = n0 + t;
        if (tid < PA) gs[tid] = P.G[(size_t)n * PA + tid] + Gd[t * PA + tid];
        __syncthreads();
        // conditional: prec (full, symmetric) into L, h into hv
        if (tid < D * D) {
            const int a = tid / D, b = tid - a * D;
            const double v = (a == b ? P.inv_sigma_F : 0.0) + P.llT[a * D + b] * gs[gk_qidx(D, min(a, b), max(a, b))];
            L[a][b] = v;
            if (P.out_prec) P.out_prec[((size_t)n * D + a) * D + b] = v;
        } else if (tid >= 64 && tid < 64 + D) {
            const int a = tid - 64;
            const double lin = P.Phi[(size_t)n * PA + Q + a];
            const double v = P.lam[a] * (gs[Pn + a] - (probit ? 0.0 : sumF[a] - lin) - P.mu0 * gs[Q + a]);
            hv[a] = v;
            if (P.out_h) P.out_h[(size_t)n * D + a] = v;
        }
        __syncthreads();
        if (tid == 0) {                                   // in-place Cholesky (lower) of the D x D precision, fp64
            for (int j = 0; j < D; ++j) {
                double d = L[j][j];
                for (int k = 0; k < j; ++k) d -= L[j][k] * L[j][k];
                d = sqrt(d); L[j][j] = d;
                const double r = 1.0 / d;
                for (int i = j + 1; i < D; ++i) {
                    double s = L[i][j];
                    for (int k = 0; k < j; ++k) s -= L[i][k] * L[j][k];
                    L[i][j] = s * r;
                }
            }
        }
        __syncthreads();
        if (tid < D) {                                    // column tid of prec^-1, blended into Sigma
            double x[D];
            for (int i = 0; i < D; ++i) {
                double s = i == tid ? 1.0 : 0.0;
                for (int k = 0; k < i; ++k) s -= L[i][k] * x[k];
                x[i] = s / L[i][i];
            }
            for (int i = D - 1; i >= 0; --i) {
                double s = x[i];
                for (int k = i + 1; k < D; ++k) s -= L[k][i] * x[k];
                x[i] = s / L[i][i];
            }
            for (int i = 0; i < D; ++i) {
                const double o = P.Sig[((size_t)n * D + i) * D + tid];
                const double v = (1.0 - P.stepsize) * o + P.stepsize * x[i];
                Sn[i][tid] = v; P.Sig[((size_t)n * D + i) * D + tid] = v;
            }
        }
        __syncthreads();
        if (tid < D) {
            double s = 0.0;
            for (int k = 0; k < D; ++k) s += Sn[tid][k] * hv[k];
            const double v = (1.0 - P.stepsize) * P.mu[(size_t)n * D + tid] + P.stepsize * s;
            mun[tid] = v; P.mu[(size_t)n * D + tid] = v;
        }
        __syncthreads();
        if (tid < PA) {
            const double o = P.Phi[(size_t)n * PA + tid];
            double v = o;
            if (tid < Q) v = Sn[pa][pb] + mun[pa] * mun[pb];
            else if (P.track) v = mun[pa];
            const double d = v - o;
            P.Phi[(size_t)n * PA + tid] = v;
            dps[tid] = d; ((ACC *)P.dphi)[t * PA + tid] = (ACC)d;
            if (tid >= Pn) sumF[pa] += d;                 // one owner per a (the A-linear column)
        }
        __syncthreads();
        for (int e = tid; e < (nb - t - 1) * PA; e += 256) {
            const int jr = e / PA, p = e - jr * PA, j = t + 1 + jr;
            const double x = p < Pn ? (probit ? 2.0 : (double)Sb[t * NB + j]) : (probit ? (double)Sb[t * NB + j] : (double)Ab[t * NB + j]);
            Gd[j * PA + p] += x * dps[p];
        }
        __syncthreads();
    }
    for (int t = tid; t < nb * PA; t += 256) P.G[(size_t)n0 * PA + t] += Gd[t];
    if (tid < D) P.glob[GK_EIG_GLOB_SUMF + tid] = sumF[tid];
}

// pair statistics from an all-current G and Phi (one CTA per node range, fixed order):
//   Pq[ab] = 1/2 sum_n Phi_n[ab] G[n][ab],  hS[a] = 1/2 sum_n Phi_n[Q+a] G[n][Q+a],
//   hK[a]  = 1/2 sum_n Phi_n[P+a] (G[n][P+a] - (sumF[a] - Phi_n[P+a]))
template <int D>
__global__ void k_eig_gstats_phi(const double *__restrict__ G, const double *__restrict__ Phi, int N, const double *__restrict__ glob,
                                 double *__restrict__ part, int probit) {
    constexpr int Pn = gk_edim<D>::P, PA = gk_edim<D>::PA;
    __shared__ double red[8][gk_edim<D>::PA];
    const int tid = threadIdx.x, p = tid % 64, g = tid / 64;
    const int per = (N + gridDim.x - 1) / gridDim.x, lo = blockIdx.x * per, hi = min(N, lo + per);
    double s = 0.0;
    if (p < PA) {
        int pa, pb; gk_col_ab<D>(p, pa, pb);
        const double sF = glob[GK_EIG_GLOB_SUMF + pa];
        for (int n = lo + g; n < hi; n += 8) {
            const double ph = Phi[(size_t)n * PA + p], gv = G[(size_t)n * PA + p];
            s += p < Pn ? ph * gv : ph * (gv - (probit ? 0.0 : sF - ph));
        }
        red[g][p] = s;
    }
    __syncthreads();
    if (tid < PA) { double t = 0.0; for (int k = 0; k < 8; ++k) t += red[k][tid]; part[(size_t)blockIdx.x * PA + tid] = 0.5 * t; }
}
