// What slows a latency-bound solver warp down when other warps share its SM?  One warp runs the registers-only 8x8
// factor-and-draw chain of chol_chain.cu (~460 cycles/node alone); the variants add, one at a time: a block fence per node,
// a named-barrier handshake per node with partner warps, partner warps hammering shared memory, partner warps on the
// solver's own sub-partition.  nvcc -arch=sm_100a -O3 solver_env.cu -o solver_env
#include <cstdio>
#include <cuda_runtime.h>
constexpr int D = 8, Q = 36;
__host__ __device__ constexpr int qidx(int a, int b) { return a * D - a * (a - 1) / 2 + (b - a); }
__device__ __forceinline__ float rsq_fast(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void bar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }

__device__ __forceinline__ void chol_draw(float (&a)[Q], float (&w)[D], const float (&zz)[D]) {
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
#pragma unroll
    for (int d = 0; d < D; ++d) w[d] += zz[d];
#pragma unroll
    for (int jj = D - 1; jj >= 0; --jj) {
        w[jj] *= rinv[jj];
#pragma unroll
        for (int kk = 0; kk < jj; ++kk) w[kk] -= a[qidx(kk, jj)] * w[jj];
    }
}

// FENCE: __threadfence_block per node; HAND: bar.arrive(1) + bar.sync(2) per node with the partner warps;
// LOADW: partner warps run a shared-memory streaming loop; SAMEQ: partner warps include warps 4, 8 (solver's sub-partition)
template <int FENCE, int HAND, int LOADW, int SAMEQ, int BASE>
__global__ void k_env(const float *Lam0, const float *h0, const float *z, float *out, long long *cyc, int iters) {
    __shared__ __align__(16) float buf[8192];
    __shared__ __align__(16) float baseS[2][48];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) buf[i] = 1.0f / (1 + i);
    if (threadIdx.x < 96) baseS[threadIdx.x / 48][threadIdx.x % 48] = threadIdx.x < 36 ? Lam0[threadIdx.x % 36] : 0.3f;
    __syncthreads();
    const bool partner = warp != 0 && (SAMEQ || (warp & 3) != 0);
    const int npart = SAMEQ ? (blockDim.x / 32 - 1) : (blockDim.x / 32 - 1 - (blockDim.x / 32 - 1) / 4);
    if (warp == 0) {
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
            if (BASE) {
#pragma unroll
                for (int q = 0; q < 9; ++q) { float4 v = ((const float4 *)baseS[it & 1])[q]; base[4 * q] = v.x; base[4 * q + 1] = v.y; base[4 * q + 2] = v.z; base[4 * q + 3] = v.w; }
            }
#pragma unroll
            for (int aa = 0; aa < D; ++aa) {
#pragma unroll
                for (int bb = aa; bb < D; ++bb) a[qidx(aa, bb)] = base[qidx(aa, bb)] + g[aa] * f[bb];
                w[aa] = hb[aa] + g[aa];
            }
            chol_draw(a, w, zz);
#pragma unroll
            for (int d = 0; d < D; ++d) f[d] = w[d];
            if (lane == 0) { buf[(it & 63) * 8] = f[0]; }
            if (FENCE) __threadfence_block();
            if (HAND) { __syncwarp(); bar_arrive(1, 32 + 32 * npart); bar_sync(2, 32 + 32); }
        }
        long long t1 = clock64();
        if (lane == 0) { cyc[0] = t1 - t0; for (int d = 0; d < D; ++d) out[d] = f[d]; }
    } else if (partner) {
        float acc = 0.f;
        for (int it = 0; it < iters; ++it) {
            if (HAND) { bar_sync(1, 32 + 32 * npart); }
            if (LOADW) {
                for (int k = 0; k < LOADW; ++k) acc += buf[(lane * 4 + k * 128 + warp * 32 + it) & 8191] * buf[(k * 64 + it) & 8191];
            }
            if (HAND && warp == 1) { __threadfence_block(); __syncwarp(); bar_arrive(2, 32 + 32); }
        }
        out[64 + threadIdx.x] = acc;
    }
}

int main() {
    float hL[Q], hh[D], hz[128 * D];
    for (int a = 0; a < D; ++a) for (int b = a; b < D; ++b) hL[qidx(a, b)] = a == b ? 10.f + a : 0.3f / (1 + b - a);
    for (int d = 0; d < D; ++d) hh[d] = 0.5f - 0.1f * d;
    for (int i = 0; i < 128 * D; ++i) hz[i] = ((i * 7919) % 1000) / 500.f - 1.f;
    float *dL, *dh, *dz, *dout; long long *dc;
    cudaMalloc(&dL, sizeof hL); cudaMalloc(&dh, sizeof hh); cudaMalloc(&dz, sizeof hz); cudaMalloc(&dout, 4096 * 4); cudaMalloc(&dc, 64);
    cudaMemcpy(dL, hL, sizeof hL, cudaMemcpyHostToDevice); cudaMemcpy(dh, hh, sizeof hh, cudaMemcpyHostToDevice);
    cudaMemcpy(dz, hz, sizeof hz, cudaMemcpyHostToDevice);
    const int iters = 4096; long long c[2];
#define RUN(NAME, ...) for (int rep = 0; rep < 2; ++rep) { k_env<__VA_ARGS__><<<1, 384>>>(dL, dh, dz, dout, dc, iters); cudaDeviceSynchronize(); } \
    cudaMemcpy(c, dc, 8, cudaMemcpyDeviceToHost); printf("%-70s %.0f cycles/node [%s]\n", NAME, (double)c[0] / iters, cudaGetErrorString(cudaGetLastError()));
    RUN("alone (other warps exit)", 0, 0, 0, 0, 0)
    RUN("+ base row from shared memory (9 LDS.128)", 0, 0, 0, 0, 1)
    RUN("+ __threadfence_block per node", 1, 0, 0, 0, 1)
    RUN("+ handshake (arrive 1 / sync 2) with 9 idle partner warps", 1, 1, 0, 0, 1)
    RUN("+ partners stream shared memory, 16 loads each per node", 1, 1, 16, 0, 1)
    RUN("+ partners stream shared memory, 64 loads each per node", 1, 1, 64, 0, 1)
    RUN("same, partners also on the solver's sub-partition (11 warps)", 1, 1, 64, 1, 1)
    RUN("no handshake, partners stream 64 loads per node", 1, 0, 64, 0, 1)
    return 0;
}
