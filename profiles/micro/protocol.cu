// Skeleton of k_eig_seq5's handshakes: solver (registers-only 8x8 chain) + 2 finisher warps + 7 summer warps, named barriers
// 1..4 as in the kernel, with adjustable stand-in work.  Which handoff costs what?  nvcc -arch=sm_100a -O3 protocol.cu
#include <cstdio>
#include <cuda_runtime.h>
constexpr int D = 8, Q = 36;
__host__ __device__ constexpr int qidx(int a, int b) { return a * D - a * (a - 1) / 2 + (b - a); }
__device__ __forceinline__ float rsq_fast(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void bar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// K1: dependent shared-memory round trips of a finisher before arrive 3; K2: between arrive 3 and arrive 2 (one barrier 4 inside);
// W: iterations of the summers' streaming loop (3 loads + 4 FMA each); MID: solver waits in the middle (1) or at the end (0)
template <int K1, int K2, int W, int NSUM>
__global__ void k_proto(const float *Lam0, const float *h0, const float *z, float *out, long long *cyc, int iters) {
    __shared__ __align__(16) float buf[8192+2048];
    __shared__ __align__(16) float baseS[2][48];
    __shared__ float fS[8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 10240; i += blockDim.x) buf[i] = 1.0f / (1 + (i & 1023));
    if (threadIdx.x < 96) baseS[threadIdx.x / 48][threadIdx.x % 48] = (threadIdx.x % 48) < 36 ? Lam0[threadIdx.x % 48] : 0.3f;
    __syncthreads();
    const int nhelp = 64 + 32 * NSUM;
    if (warp == 0) {
        float base[48], hb[D], f[D];
#pragma unroll
        for (int p = 0; p < 48; ++p) base[p] = baseS[0][p];
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
            float rinv[D];
#pragma unroll
            for (int jj = 0; jj < D; ++jj) {
                float r = rsq_fast(a[qidx(jj, jj)]);
                rinv[jj] = r; w[jj] *= r;
#pragma unroll
                for (int ii = jj + 1; ii < D; ++ii) { a[qidx(jj, ii)] *= r; w[ii] -= a[qidx(jj, ii)] * w[jj]; }
#pragma unroll
                for (int kk = jj + 1; kk < D; ++kk)
#pragma unroll
                    for (int ii = kk; ii < D; ++ii) a[qidx(kk, ii)] -= a[qidx(jj, ii)] * a[qidx(jj, kk)];
            }
            __syncwarp();
            if (it >= 1) bar_sync(2, 96);
            {
                const unsigned baddr = (unsigned)__cvta_generic_to_shared(baseS[(it + 1) & 1]);
#pragma unroll
                for (int q = 0; q < 12; ++q)
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(base[4 * q]), "=f"(base[4 * q + 1]), "=f"(base[4 * q + 2]), "=f"(base[4 * q + 3]) : "r"(baddr + 16 * q) : "memory");
            }
#pragma unroll
            for (int d = 0; d < D; ++d) w[d] += zz[d];
#pragma unroll
            for (int jj = D - 1; jj >= 0; --jj) {
                w[jj] *= rinv[jj];
#pragma unroll
                for (int kk = 0; kk < jj; ++kk) w[kk] -= a[qidx(kk, jj)] * w[jj];
            }
#pragma unroll
            for (int d = 0; d < D; ++d) f[d] = w[d];
            if (lane == 0) {
#pragma unroll
                for (int d = 0; d < D; ++d) fS[d] = f[d];
            }
            __syncwarp();
            bar_arrive(1, 32 + nhelp);
        }
        long long t1 = clock64();
        if (lane == 0) { cyc[0] = t1 - t0; for (int d = 0; d < D; ++d) out[d] = f[d]; }
    } else if (warp == 1 || warp == 2) {
        const int p = (warp - 1) * 32 + lane;
        float acc = 0.f;
        for (int it = 0; it < iters; ++it) {
            __syncwarp();
            bar_sync(1, 32 + nhelp);
            float v = fS[p & 7];
            int idx = p;
#pragma unroll
            for (int k = 0; k < K1; ++k) { v += buf[idx & 8191]; idx = idx * 3 + (int)v + 64; }     // dependent round trips
            buf[8192 + p] = v;
            __syncwarp();
            bar_arrive(3, nhelp);
#pragma unroll
            for (int k = 0; k < K2; ++k) { v += buf[idx & 8191]; idx = idx * 3 + (int)v + 64; }
            buf[8300 + p] = v;
            __syncwarp();
            bar_sync(4, 64);
            if (p < 44) baseS[it & 1][p] = (p < 36 ? Lam0[p] : 0.3f) + 1e-9f * (v + buf[8300 + ((p + 1) & 63)]);
            __syncwarp();
            if (it + 1 < iters) bar_arrive(2, 96);
            acc += v;
        }
        out[64 + threadIdx.x] = acc;
    } else if ((warp & 3) != 0 && warp < 3 + NSUM + (NSUM + 2) / 3 + 1) {
        float a0 = 0, a1 = 0, a2 = 0, a3 = 0;
        int nmine = 0;
        for (int it = 0; it < iters; ++it) {
            __syncwarp();
            bar_sync(1, 32 + nhelp);
            bar_sync(3, nhelp);
#pragma unroll 4
            for (int k = 0; k < W; ++k) {
                const float x = buf[(k * 17 + it) & 8191] + (float)((const unsigned char *)buf)[(k * 13 + it) & 8191];
                const float4 d4 = *(const float4 *)(buf + ((k * 512 + threadIdx.x * 4) & 8188));
                a0 = fmaf(x, d4.x, a0); a1 = fmaf(x, d4.y, a1); a2 = fmaf(x, d4.z, a2); a3 = fmaf(x, d4.w, a3);
            }
            *(float4 *)(buf + 8704 + threadIdx.x * 4) = make_float4(a0, a1, a2, a3);
            ++nmine;
        }
        out[512 + threadIdx.x] = a0 + a1 + a2 + a3 + nmine;
    }
}

int main() {
    float hL[48], hh[D], hz[128 * D];
    for (int a = 0; a < D; ++a) for (int b = a; b < D; ++b) hL[qidx(a, b)] = a == b ? 10.f + a : 0.3f / (1 + b - a);
    for (int i = 36; i < 48; ++i) hL[i] = 0.3f;
    for (int d = 0; d < D; ++d) hh[d] = 0.5f - 0.1f * d;
    for (int i = 0; i < 128 * D; ++i) hz[i] = ((i * 7919) % 1000) / 500.f - 1.f;
    float *dL, *dh, *dz, *dout; long long *dc;
    cudaMalloc(&dL, sizeof hL); cudaMalloc(&dh, sizeof hh); cudaMalloc(&dz, sizeof hz); cudaMalloc(&dout, 4096 * 4); cudaMalloc(&dc, 64);
    cudaMemcpy(dL, hL, sizeof hL, cudaMemcpyHostToDevice); cudaMemcpy(dh, hh, sizeof hh, cudaMemcpyHostToDevice);
    cudaMemcpy(dz, hz, sizeof hz, cudaMemcpyHostToDevice);
    const int iters = 4096; long long c[2];
#define RUN(NAME, ...) for (int rep = 0; rep < 2; ++rep) { k_proto<__VA_ARGS__><<<1, 384>>>(dL, dh, dz, dout, dc, iters); cudaDeviceSynchronize(); } \
    cudaMemcpy(c, dc, 8, cudaMemcpyDeviceToHost); printf("%-64s %.0f cycles/node [%s]\n", NAME, (double)c[0] / iters, cudaGetErrorString(cudaGetLastError()));
    RUN("handshakes only (K1=0,K2=0,W=0), 7 summer warps", 0, 0, 0, 7)
    RUN("finisher 2+2 round trips", 2, 2, 0, 7)
    RUN("finisher 4+6 round trips", 4, 6, 0, 7)
    RUN("finisher 4+6, summers 4 iterations", 4, 6, 4, 7)
    RUN("finisher 4+6, summers 8 iterations", 4, 6, 8, 7)
    RUN("finisher 4+6, summers 16 iterations", 4, 6, 16, 7)
    RUN("finisher 2+2, summers 8 iterations", 2, 2, 8, 7)
    return 0;
}
