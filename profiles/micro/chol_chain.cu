// Micro-benchmark for the serial eigenmodel chain (k_eig_seq*): cycles per node of a dependent chain
//   conditional (depends on the previous draw) -> 8x8 Cholesky + two triangular solves -> draw,
// for several ways of running the factorisation on one warp.  nvcc -arch=sm_100a -O3 chol_chain.cu -o chol_chain
#include <cstdio>
#include <cuda_runtime.h>
constexpr int D = 8, Q = 36;
__host__ __device__ constexpr int qidx(int a, int b) { return a * D - a * (a - 1) / 2 + (b - a); }
__device__ __forceinline__ float rsq_fast(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

template <bool FAST>
__device__ __forceinline__ void chol_draw(float (&a)[Q], float (&w)[D], const float (&zz)[D]) {
    float rinv[D];
#pragma unroll
    for (int jj = 0; jj < D; ++jj) {
        float r = FAST ? rsq_fast(a[qidx(jj, jj)]) : rsqrtf(a[qidx(jj, jj)]);
        rinv[jj] = r; w[jj] *= r;
#pragma unroll
        for (int ii = jj + 1; ii < D; ++ii) { a[qidx(jj, ii)] *= r; w[ii] -= a[qidx(jj, ii)] * w[jj]; }
#pragma unroll
        for (int kk = jj + 1; kk < D; ++kk)
#pragma unroll
            for (int ii = kk; ii < D; ++ii) a[qidx(kk, ii)] -= a[qidx(jj, ii)] * a[qidx(jj, kk)];
    }
#pragma unroll
    for (int d = 0; d < D; ++d) w[d] += zz[d];
#pragma unroll
    for (int jj = D - 1; jj >= 0; --jj) {
        w[jj] *= rinv[jj];
#pragma unroll
        for (int kk = 0; kk < jj; ++kk) w[kk] -= a[qidx(kk, jj)] * w[jj];
    }
}

// LDL^T form: pivots need a reciprocal (MUFU.RCP) only; the square roots for the noise are off the chain
__device__ __forceinline__ void ldl_draw(float (&a)[Q], float (&w)[D], const float (&zz)[D]) {
    float dinv[D], dr[D];
#pragma unroll
    for (int jj = 0; jj < D; ++jj) {
        float piv = a[qidx(jj, jj)], r;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(piv));
        dinv[jj] = r; dr[jj] = rsq_fast(piv);
        float lrow[D];
#pragma unroll
        for (int ii = jj + 1; ii < D; ++ii) { lrow[ii] = a[qidx(jj, ii)] * r; w[ii] -= lrow[ii] * w[jj]; }
#pragma unroll
        for (int kk = jj + 1; kk < D; ++kk)
#pragma unroll
            for (int ii = kk; ii < D; ++ii) a[qidx(kk, ii)] -= a[qidx(jj, ii)] * lrow[kk];
#pragma unroll
        for (int ii = jj + 1; ii < D; ++ii) a[qidx(jj, ii)] = lrow[ii];
    }
    // w = L^-1 h ; x = D^-1 w + D^-1/2 z ; f = L^-T x
#pragma unroll
    for (int d = 0; d < D; ++d) w[d] = w[d] * dinv[d] + zz[d] * dr[d];
#pragma unroll
    for (int jj = D - 1; jj >= 0; --jj)
#pragma unroll
        for (int kk = 0; kk < jj; ++kk) w[kk] -= a[qidx(kk, jj)] * w[jj];
}

// mode 0: lane 0 alone (operands through shared memory, as k_eig_seq4 does); 1: same with rsqrt.approx.ftz;
// 2: every lane factorises redundantly after a shared-memory all-gather (no broadcast of f afterwards); 3: LDL^T redundant
template <int MODE>
__global__ void k_chain(const float *Lam0, const float *h0, const float *z, float *out, long long *cyc, int iters) {
    __shared__ float LamS[Q + D], fS[D];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (warp != 0) return;
    float f[D];
#pragma unroll
    for (int d = 0; d < D; ++d) f[d] = 0.1f * d;
    // lane roles as in k_eig_seq4: lane < 32 -> column lane, lanes 0..3 also column lane+32; lanes 16..23 h[lane-16]
    int a0 = 0, b0 = 0; { int p = lane, a = 0; while (p >= D - a) { p -= D - a; ++a; } a0 = a; b0 = a + p; }
    int a1 = 0, b1 = 0; if (lane + 32 < Q) { int p = lane + 32, a = 0; while (p >= D - a) { p -= D - a; ++a; } a1 = a; b1 = a + p; }
    const float l0 = Lam0[lane], l1 = lane + 32 < Q ? Lam0[lane + 32] : 0.f, hh = lane >= 16 && lane < 24 ? h0[lane - 16] : 0.f;
    long long t0 = clock64();
    unsigned long long g0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0));
    for (int it = 0; it < iters; ++it) {
        float zz[D];
#pragma unroll
        for (int d = 0; d < D; ++d) zz[d] = z[(it & 127) * D + d];
        if (MODE <= 1) {
            // every lane builds its columns of the conditional from the previous draw (f lives in shared memory)
            const float y0 = l0 + 0.01f * fS[a0] * fS[b0] * (it > 0);
            LamS[lane] = y0;
            if (lane + 32 < Q) LamS[lane + 32] = l1 + 0.01f * fS[a1] * fS[b1] * (it > 0);
            if (lane >= 16 && lane < 24) LamS[Q + lane - 16] = hh + 0.01f * fS[lane - 16] * (it > 0);
            __syncwarp();
            if (lane == 0) {
                float a[Q], w[D];
#pragma unroll
                for (int p = 0; p < Q; ++p) a[p] = LamS[p];
#pragma unroll
                for (int d = 0; d < D; ++d) w[d] = LamS[Q + d];
                chol_draw<MODE == 1>(a, w, zz);
#pragma unroll
                for (int d = 0; d < D; ++d) fS[d] = w[d];
            }
            __syncwarp();
        } else {
            const float y0 = l0 + 0.01f * f[a0] * f[b0] * (it > 0);
            LamS[lane] = y0;
            if (lane + 32 < Q) LamS[lane + 32] = l1 + 0.01f * f[a1] * f[b1] * (it > 0);
            if (lane >= 16 && lane < 24) LamS[Q + lane - 16] = hh + 0.01f * f[lane - 16] * (it > 0);
            __syncwarp();
            float a[Q], w[D];
            const float4 *L4 = (const float4 *)LamS;
#pragma unroll
            for (int p = 0; p < Q / 4; ++p) { float4 v = L4[p]; a[4 * p] = v.x; a[4 * p + 1] = v.y; a[4 * p + 2] = v.z; a[4 * p + 3] = v.w; }
#pragma unroll
            for (int p = 0; p < D / 4; ++p) { float4 v = L4[Q / 4 + p]; w[4 * p] = v.x; w[4 * p + 1] = v.y; w[4 * p + 2] = v.z; w[4 * p + 3] = v.w; }
            if (MODE == 2) chol_draw<true>(a, w, zz); else ldl_draw(a, w, zz);
#pragma unroll
            for (int d = 0; d < D; ++d) f[d] = w[d];
            __syncwarp();
        }
    }
    unsigned long long g1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
    long long t1 = clock64();
    if (lane == 0) {
        cyc[0] = t1 - t0; cyc[1] = (long long)(g1 - g0);
        for (int d = 0; d < D; ++d) out[d] = MODE <= 1 ? fS[d] : f[d];
    }
}


// hand-ordered right-looking factorisation: the next pivot's rsqrt is issued before the rest of the trailing update
__device__ __forceinline__ void chol_draw_la(float (&a)[Q], float (&w)[D], const float (&zz)[D]) {
    float rinv[D];
    float r = rsq_fast(a[qidx(0, 0)]);
#pragma unroll
    for (int jj = 0; jj < D; ++jj) {
        rinv[jj] = r;
        float rn = 0.f;
        if (jj + 1 < D) {
            a[qidx(jj, jj + 1)] *= r;
            a[qidx(jj + 1, jj + 1)] -= a[qidx(jj, jj + 1)] * a[qidx(jj, jj + 1)];
            rn = rsq_fast(a[qidx(jj + 1, jj + 1)]);
        }
        w[jj] *= r;
#pragma unroll
        for (int ii = jj + 2; ii < D; ++ii) a[qidx(jj, ii)] *= r;
#pragma unroll
        for (int ii = jj + 1; ii < D; ++ii) w[ii] -= a[qidx(jj, ii)] * w[jj];
#pragma unroll
        for (int kk = jj + 1; kk < D; ++kk)
#pragma unroll
            for (int ii = kk; ii < D; ++ii)
                if (!(kk == jj + 1 && ii == jj + 1)) a[qidx(kk, ii)] -= a[qidx(jj, ii)] * a[qidx(jj, kk)];
        r = rn;
    }
#pragma unroll
    for (int d = 0; d < D; ++d) w[d] += zz[d];
#pragma unroll
    for (int jj = D - 1; jj >= 0; --jj) {
        w[jj] *= rinv[jj];
#pragma unroll
        for (int kk = 0; kk < jj; ++kk) w[kk] -= a[qidx(kk, jj)] * w[jj];
    }
}

// registers only: lane 0 keeps f and rebuilds the conditional itself (rank-1 change of a fixed base), no shared memory
template <int MODE>
__global__ void k_chain_reg(const float *Lam0, const float *h0, const float *z, float *out, long long *cyc, int iters) {
    if (threadIdx.x != 0) return;
    float base[Q], hb[D], f[D];
#pragma unroll
    for (int p = 0; p < Q; ++p) base[p] = Lam0[p];
#pragma unroll
    for (int d = 0; d < D; ++d) { hb[d] = h0[d]; f[d] = 0.f; }
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        float zz[D], a[Q], w[D], g[D];
#pragma unroll
        for (int d = 0; d < D; ++d) { zz[d] = z[(it & 127) * D + d]; g[d] = 0.01f * f[d]; }
#pragma unroll
        for (int aa = 0; aa < D; ++aa) {
#pragma unroll
            for (int bb = aa; bb < D; ++bb) a[qidx(aa, bb)] = base[qidx(aa, bb)] + g[aa] * f[bb];
            w[aa] = hb[aa] + g[aa];
        }
        if (MODE == 0) chol_draw<true>(a, w, zz); else chol_draw_la(a, w, zz);
#pragma unroll
        for (int d = 0; d < D; ++d) f[d] = w[d];
    }
    long long t1 = clock64();
    cyc[0] = t1 - t0;
    for (int d = 0; d < D; ++d) out[d] = f[d];
}

// latencies: dependent chains of one instruction type
__global__ void k_lat(float *out, long long *cyc, int iters) {
    float x = 1.0001f + threadIdx.x * 1e-6f;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) x = rsq_fast(x);
    long long t1 = clock64();
    float y = x;
    for (int i = 0; i < iters; ++i) y = y * 1.0001f + 0.5f;
    long long t2 = clock64();
    float s = y;
    for (int i = 0; i < iters; ++i) s = __shfl_sync(0xffffffffu, s, (threadIdx.x + 1) & 31);
    long long t3 = clock64();
    if (threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; }
    out[threadIdx.x] = s;
}


// issue rate of independent FFMAs from ONE warp (16 chains), with `active` lanes; and LDS latency (pointer chase)
__global__ void k_ilp(float *out, long long *cyc, int iters, int active) {
    __shared__ int chase[64];
    if (threadIdx.x < 64) chase[threadIdx.x] = (threadIdx.x * 17 + 1) & 63;
    __syncthreads();
    long long t0 = 0, t1 = 0, t2 = 0;
    float acc[16];
    if ((int)threadIdx.x < active) {
#pragma unroll
        for (int k = 0; k < 16; ++k) acc[k] = 1.0f + k + threadIdx.x;
        t0 = clock64();
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int k = 0; k < 16; ++k) acc[k] = acc[k] * 1.0001f + 0.5f;
        }
        t1 = clock64();
        int p = threadIdx.x;
        for (int i = 0; i < iters; ++i) p = chase[p];
        t2 = clock64();
        float s = p;
#pragma unroll
        for (int k = 0; k < 16; ++k) s += acc[k];
        out[threadIdx.x] = s;
        if (threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; }
    }
}

int main() {
    float hL[Q], hh[D], hz[128 * D];
    for (int a = 0; a < D; ++a) for (int b = a; b < D; ++b) hL[qidx(a, b)] = a == b ? 10.f + a : 0.3f / (1 + b - a);
    for (int d = 0; d < D; ++d) hh[d] = 0.5f - 0.1f * d;
    for (int i = 0; i < 128 * D; ++i) hz[i] = ((i * 7919) % 1000) / 500.f - 1.f;
    float *dL, *dh, *dz, *dout; long long *dc;
    cudaMalloc(&dL, sizeof hL); cudaMalloc(&dh, sizeof hh); cudaMalloc(&dz, sizeof hz); cudaMalloc(&dout, 256 * 4); cudaMalloc(&dc, 64);
    cudaMemcpy(dL, hL, sizeof hL, cudaMemcpyHostToDevice); cudaMemcpy(dh, hh, sizeof hh, cudaMemcpyHostToDevice);
    cudaMemcpy(dz, hz, sizeof hz, cudaMemcpyHostToDevice);
    const int iters = 4096;
    long long c[4]; float o[D];
    for (int mode = 0; mode < 4; ++mode) {
        for (int rep = 0; rep < 2; ++rep) {
            if (mode == 0) k_chain<0><<<1, 32>>>(dL, dh, dz, dout, dc, iters);
            if (mode == 1) k_chain<1><<<1, 32>>>(dL, dh, dz, dout, dc, iters);
            if (mode == 2) k_chain<2><<<1, 32>>>(dL, dh, dz, dout, dc, iters);
            if (mode == 3) k_chain<3><<<1, 32>>>(dL, dh, dz, dout, dc, iters);
            cudaDeviceSynchronize();
        }
        cudaMemcpy(c, dc, 16, cudaMemcpyDeviceToHost); cudaMemcpy(o, dout, sizeof o, cudaMemcpyDeviceToHost);
        printf("mode %d: %.1f clock64 ticks/node, %.1f ns/node  f = %.5f %.5f %.5f %.5f  [%s]\n", mode, (double)c[0] / iters, (double)c[1] / iters,
               o[0], o[1], o[2], o[7], cudaGetErrorString(cudaGetLastError()));
    }
    for (int mode = 0; mode < 2; ++mode) {
        for (int rep = 0; rep < 2; ++rep) {
            if (mode == 0) k_chain_reg<0><<<1, 32>>>(dL, dh, dz, dout, dc, iters); else k_chain_reg<1><<<1, 32>>>(dL, dh, dz, dout, dc, iters);
            cudaDeviceSynchronize();
        }
        cudaMemcpy(c, dc, 8, cudaMemcpyDeviceToHost); cudaMemcpy(o, dout, sizeof o, cudaMemcpyDeviceToHost);
        printf("registers-only chain, order %d: %.1f cycles/node  f = %.5f %.5f %.5f\n", mode, (double)c[0] / iters, o[0], o[1], o[7]);
    }
    k_lat<<<1, 32>>>(dout, dc, 4096); cudaDeviceSynchronize();
    cudaMemcpy(c, dc, 24, cudaMemcpyDeviceToHost);
    printf("latency (clock64 ticks): rsqrt.approx %.1f  ffma %.1f  shfl %.1f\n", c[0] / 4096.0, c[1] / 4096.0, c[2] / 4096.0);
    for (int active : {1, 16, 32}) {
        k_ilp<<<1, 64>>>(dout, dc, 4096, active); cudaDeviceSynchronize();
        cudaMemcpy(c, dc, 16, cudaMemcpyDeviceToHost);
        printf("active lanes %2d: independent FFMA %.2f cycles each; LDS dependent latency %.1f\n", active, c[0] / 4096.0 / 16, c[1] / 4096.0);
    }
    return 0;
}
