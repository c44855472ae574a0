// gk_eig_tc.cu -- the Omega . (F (x) F) contraction of the logistic eigenmodel on the 5th-generation tensor cores.
//
// eigenmodel.py:884-901 builds, for every node n, prec_n = I/sigma_F + sum_m (w_nm + w_mn) (f_m f_m^T) o (lmbda lmbda^T) and
// h_n from sum_m [(k_nm + k_mn) - mu_0 (w_nm + w_mn)] f_m.  With S = Omega + Omega^T (symmetric) and Phi(f) = [f_a f_b | f_a]
// the S part is ONE GEMM per sweep:
//     G[n][p] = sum_m S[n][m] Phi[m][p]        M = n (N nodes), K = m (N partners), N = p (44 monomials, padded to 64)
// A = S read row-major IS the K-major A operand (no transpose: S is symmetric), B = Phi^T (64 x N, written once per sweep,
// rounded to tf32), D = 128 x 64 fp32 in tensor memory.  Per CTA: one 128-node strip x one chunk of partners.
//   warp 0      TMA producer: cp.async.bulk.tensor tiles of S (128 x 32 fp32, SWIZZLE_128B), Phi^T (64 x 32), Asym (32 x 128 u8)
//               and a bulk copy of 32 feature rows into a 3-stage mbarrier ring
//   warp 1      allocates 128 TMEM columns (two accumulators); one elected lane issues tcgen05.mma.cta_group::1.kind::tf32 (M128 N64 K8, four per stage)
//               and tcgen05.commit's the stage back to the producer
//   warps 2-5   the Asym part (8 columns of G, 8 FMAs per byte of A) on the CUDA cores from the same stage, then the epilogue:
//               tcgen05.ld of the accumulator (each warp its 32 TMEM lanes) -> partial[chunk][n][p]
// Bound: HBM (5 bytes per (n, m) pair against 88 tensor flops + 16 CUDA-core flops).  fp64-storage, probit and mean-field
// contractions keep the FMA kernel (k_eig_gfull).
#include <cuda.h>
#include <cudaTypedefs.h>
#include "gk_internal.cuh"

#define TC_STAGES 2
#define TC_KB 32                                   // partners per stage (32 tf32 = one 128-byte swizzle row)
#define TC_THREADS 192
#define TC_GROUP 8                                 // stages per tensor-memory accumulation (256 partners)
#define TC_S_BYTES (128 * TC_KB * 4)               // 16384
#define TC_P_BYTES (64 * TC_KB * 4)                // 8192
#define TC_A_BYTES (TC_KB * 128)                   // 4096
#define TC_F_BYTES (TC_KB * 8 * 4)                 // 1024
#define TC_STAGE_BYTES (2 * TC_S_BYTES + TC_P_BYTES + TC_A_BYTES + TC_F_BYTES)       // S | S_lo | Phi^T | Asym | F = 46080 = 45 * 1024
#define TC_OFF_LO TC_S_BYTES
#define TC_OFF_P (2 * TC_S_BYTES)
#define TC_OFF_A (2 * TC_S_BYTES + TC_P_BYTES)
#define TC_OFF_F (2 * TC_S_BYTES + TC_P_BYTES + TC_A_BYTES)
#define TC_TX_BYTES (TC_S_BYTES + TC_P_BYTES + TC_A_BYTES + TC_F_BYTES)
#define TC_PA 52
#define TC_PS 44                                   // S columns (36 quadratic + 8 linear)

__device__ __forceinline__ uint32_t tc_smem(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tc_mbar_init(uint64_t *b, int count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tc_smem(b)), "r"(count)); }
__device__ __forceinline__ void tc_mbar_expect_tx(uint64_t *b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc_smem(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tc_mbar_arrive(uint64_t *b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tc_smem(b)) : "memory"); }
__device__ __forceinline__ void tc_mbar_wait(uint64_t *b, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}"
                 ::"r"(tc_smem(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tc_tma_2d(void *dst, const CUtensorMap *map, int x, int y, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(tc_smem(dst)), "l"(map), "r"(tc_smem(bar)), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void tc_bulk_1d(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(tc_smem(dst)), "l"(src), "r"(bytes), "r"(tc_smem(bar)) : "memory");
}
// shared-memory matrix descriptor: K-major, SWIZZLE_128B, 8-row groups 1024 bytes apart (cute::UMMA::SmemDescriptor, sm100 version 1)
__device__ __forceinline__ uint64_t tc_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor: D fp32, A and B tf32, both K-major, N = 64, M = 128 (cute::UMMA::InstrDescriptor)
#define TC_IDESC ((1u << 4) | (2u << 7) | (2u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24))
__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(TC_IDESC), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(tc_smem(bar)) : "memory");
}
#define TC_LD16(r, taddr)                                                                                              \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];" \
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),  \
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])              \
                 : "r"(taddr))

__host__ __device__ inline int tc_qidx(int a, int b) { return a * 8 - a * (a - 1) / 2 + (b - a); }

// Phi^T (64 x npad, tf32-rounded fp32; rows 44..63 and columns >= N zero) and the features as fp32 rows (npad x 8)
__global__ void k_tc_prep(const double *__restrict__ F, int N, int npad, float *__restrict__ PhiT, float *__restrict__ Ff) {
    for (int m = blockIdx.x * blockDim.x + threadIdx.x; m < npad; m += gridDim.x * blockDim.x) {
        float f[8];
#pragma unroll
        for (int d = 0; d < 8; ++d) f[d] = m < N ? (float)F[(size_t)m * 8 + d] : 0.f;
#pragma unroll
        for (int d = 0; d < 8; ++d) Ff[(size_t)m * 8 + d] = f[d];
#pragma unroll
        for (int a = 0; a < 8; ++a) {
#pragma unroll
            for (int b = a; b < 8; ++b) {
                float v = f[a] * f[b]; uint32_t t;
                asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(t) : "f"(v));
                PhiT[(size_t)tc_qidx(a, b) * npad + m] = __uint_as_float(t);
            }
            uint32_t t;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(t) : "f"(f[a]));
            PhiT[(size_t)(36 + a) * npad + m] = __uint_as_float(t);
        }
        for (int p = TC_PS; p < 64; ++p) PhiT[(size_t)p * npad + m] = 0.f;
    }
}

struct tc_params {
    int chunk_rows, npad;
    const float *Ff;
    float *partial;                     // [nchunks][npad][TC_PA]
};

__global__ void __launch_bounds__(TC_THREADS, 2)
k_eig_gfull_tc(const __grid_constant__ CUtensorMap mapS, const __grid_constant__ CUtensorMap mapP,
               const __grid_constant__ CUtensorMap mapA, tc_params P) {
    extern __shared__ __align__(1024) unsigned char tc_smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[TC_STAGES], empty_bar[TC_STAGES], lo_bar[TC_STAGES], accF_bar[2], accE_bar[2];
    __shared__ uint32_t tmem_base_s;
    unsigned char *stages = (unsigned char *)(((uintptr_t)tc_smem_raw + 1023) & ~(uintptr_t)1023);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n0 = blockIdx.x * 128, m0 = blockIdx.y * P.chunk_rows, nk = P.chunk_rows / TC_KB;

    if (tid == 0) {
        for (int s = 0; s < TC_STAGES; ++s) { tc_mbar_init(&full_bar[s], 1); tc_mbar_init(&empty_bar[s], 5); tc_mbar_init(&lo_bar[s], 4); }
        for (int b = 0; b < 2; ++b) { tc_mbar_init(&accF_bar[b], 1); tc_mbar_init(&accE_bar[b], 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {                                   // 128 TMEM columns: two 128 x 64 fp32 accumulators
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(tc_smem(&tmem_base_s)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base_s;

    if (warp == 0) {
        // ---------------- TMA producer ----------------
        if (lane == 0) {
            for (int kb = 0; kb < nk; ++kb) {
                const int s = kb % TC_STAGES;
                tc_mbar_wait(&empty_bar[s], ((kb / TC_STAGES) & 1) ^ 1);
                unsigned char *st = stages + (size_t)s * TC_STAGE_BYTES;
                const int m = m0 + kb * TC_KB;
                tc_mbar_expect_tx(&full_bar[s], TC_TX_BYTES);
                tc_tma_2d(st, &mapS, m, n0, &full_bar[s]);                                  // S[n0 .. +128][m .. +32]
                tc_tma_2d(st + TC_OFF_P, &mapP, m, 0, &full_bar[s]);                      // Phi^T[0 .. 64][m .. +32]
                tc_tma_2d(st + TC_OFF_A, &mapA, n0, m, &full_bar[s]);        // Asym[m .. +32][n0 .. +128]
                tc_bulk_1d(st + TC_OFF_F, P.Ff + (size_t)m * 8, TC_F_BYTES, &full_bar[s]);
            }
        }
    } else if (warp == 1) {
        // ---------------- MMA issuer ----------------
        // The tensor-core accumulator drops low-order bits of every aligned product (measured: a relative bias of 2e-3 after
        // 2048 positive terms, 4e-4 -- the tf32 operand truncation -- after 128), so no accumulation in tensor memory is longer
        // than TC_GROUP stages (256 partners): two 64-column accumulators alternate, the workers fold a finished one into fp32
        // registers with ordinary round-to-nearest adds while the other fills.
        for (int kb = 0; kb < nk; ++kb) {
            const int s = kb % TC_STAGES, g = kb / TC_GROUP, buf = g & 1;
            if (kb % TC_GROUP == 0 && g >= 2) {                                             // the workers have drained this accumulator
                tc_mbar_wait(&accE_bar[buf], ((g >> 1) + 1) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            }
            tc_mbar_wait(&full_bar[s], (kb / TC_STAGES) & 1);
            tc_mbar_wait(&lo_bar[s], (kb / TC_STAGES) & 1);                                 // the workers have written S_lo of this stage
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (lane == 0) {
                const uint32_t sa = tc_smem(stages + (size_t)s * TC_STAGE_BYTES);
                const uint64_t adesc = tc_desc(sa), ldesc = tc_desc(sa + TC_OFF_LO), bdesc = tc_desc(sa + TC_OFF_P);
#pragma unroll
                for (int k = 0; k < TC_KB / 8; ++k) {                                       // K = 8 tf32 = 32 bytes inside the swizzle row
                    tc_mma_tf32(tmem_d + 64 * buf, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), ((kb % TC_GROUP) | k) != 0);
                    tc_mma_tf32(tmem_d + 64 * buf, ldesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), 1u);
                }
                tc_commit(&empty_bar[s]);                                                   // smem slot free once these MMAs have read it
                if (kb % TC_GROUP == TC_GROUP - 1 || kb == nk - 1) tc_commit(&accF_bar[buf]);   // this group's accumulator is complete
            }
            __syncwarp();
        }
    } else {
        // ---------------- Asym columns on the CUDA cores; fold of the finished accumulators; epilogue ----------------
        const int q = warp & 3, n = q * 32 + lane;                                          // node n0 + n == TMEM lane n
        float acc[8], gs[48];
#pragma unroll
        for (int d = 0; d < 8; ++d) acc[d] = 0.f;
#pragma unroll
        for (int j = 0; j < 48; ++j) gs[j] = 0.f;
        const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16);
        for (int kb = 0; kb < nk; ++kb) {
            const int s = kb % TC_STAGES;
            tc_mbar_wait(&full_bar[s], (kb / TC_STAGES) & 1);
            unsigned char *st = stages + (size_t)s * TC_STAGE_BYTES;
            {   // S = S_hi + S_lo with S_hi = the 19 bits the tf32 multiplier reads: S_lo goes through a second MMA, so the operand
                // truncation (a one-sided 3e-4 relative error, amplified when the features shrink during a sweep) disappears.
                // Elementwise at identical offsets, so the 128-byte swizzle of the tile carries over.
                const uint4 *src = (const uint4 *)st; uint4 *dst = (uint4 *)(st + TC_OFF_LO);
                const int wt = q * 32 + lane;
#pragma unroll
                for (int i = 0; i < TC_S_BYTES / 16 / 128; ++i) {
                    const uint4 v = src[wt + 128 * i];
                    uint4 o;
                    o.x = __float_as_uint(__uint_as_float(v.x) - __uint_as_float(v.x & 0xffffe000u));
                    o.y = __float_as_uint(__uint_as_float(v.y) - __uint_as_float(v.y & 0xffffe000u));
                    o.z = __float_as_uint(__uint_as_float(v.z) - __uint_as_float(v.z & 0xffffe000u));
                    o.w = __float_as_uint(__uint_as_float(v.w) - __uint_as_float(v.w & 0xffffe000u));
                    dst[wt + 128 * i] = o;
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");              // generic-proxy stores -> visible to the MMA's reads
                __syncwarp();
                if (lane == 0) tc_mbar_arrive(&lo_bar[s]);
            }
            const uint8_t *At = st + TC_OFF_A;
            const float4 *Fs = (const float4 *)(st + TC_OFF_F);
#pragma unroll 8
            for (int m = 0; m < TC_KB; ++m) {
                const float a = (float)At[m * 128 + n];
                const float4 f0 = Fs[2 * m], f1 = Fs[2 * m + 1];
                acc[0] = fmaf(a, f0.x, acc[0]); acc[1] = fmaf(a, f0.y, acc[1]); acc[2] = fmaf(a, f0.z, acc[2]); acc[3] = fmaf(a, f0.w, acc[3]);
                acc[4] = fmaf(a, f1.x, acc[4]); acc[5] = fmaf(a, f1.y, acc[5]); acc[6] = fmaf(a, f1.z, acc[6]); acc[7] = fmaf(a, f1.w, acc[7]);
            }
            __syncwarp();
            if (lane == 0) tc_mbar_arrive(&empty_bar[s]);
            if (kb % TC_GROUP == TC_GROUP - 1 || kb == nk - 1) {
                const int g = kb / TC_GROUP, buf = g & 1;
                tc_mbar_wait(&accF_bar[buf], (g >> 1) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                uint32_t r[16];
#pragma unroll
                for (int c = 0; c < 3; ++c) {                                               // columns 0..47 hold the 44 S columns
                    TC_LD16(r, taddr + 64 * buf + 16 * c);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                    for (int j = 0; j < 16; ++j) gs[16 * c + j] += __uint_as_float(r[j]);
                }
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) tc_mbar_arrive(&accE_bar[buf]);
            }
        }
        if (n0 + n < P.npad) {
            float *out = P.partial + ((size_t)blockIdx.y * P.npad + (size_t)(n0 + n)) * TC_PA;
#pragma unroll
            for (int j = 0; j < TC_PS; ++j) out[j] = gs[j];
#pragma unroll
            for (int d = 0; d < 8; ++d) out[TC_PS + d] = acc[d];
        }
    }
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(tmem_d) : "memory");
}

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled_v12000 tc_encode_fn() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = (PFN_cuTensorMapEncodeTiled_v12000)p;
    }
    return fn;
}

static bool tc_map_2d(CUtensorMap *map, CUtensorMapDataType dt, int esize, const void *base, uint64_t inner, uint64_t outer,
                      uint64_t row_bytes, uint32_t box_inner, uint32_t box_outer, CUtensorMapSwizzle sw) {
    auto fn = tc_encode_fn();
    if (!fn) return false;
    cuuint64_t dims[2] = {inner, outer}, strides[1] = {row_bytes};
    cuuint32_t box[2] = {box_inner, box_outer}, estr[2] = {1, 1};
    (void)esize;
    return fn(map, dt, 2, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// G partials of the whole network on the tensor cores.  S: nrows x ld fp32 (symmetric, zero diagonal), Asym: nrows x ld u8,
// F: N x 8 doubles.  work: device scratch of gk_tc_work_bytes(npad) bytes.  partial: [nchunks][npad][52] floats.
size_t gk_tc_work_bytes(int npad) { return (size_t)npad * 64 * 4 + (size_t)npad * 8 * 4 + 256; }

int gk_tc_gfull(gk_ctx *ctx, const float *S, const uint8_t *Asym, int ld, int N, const double *F, void *work, float *partial,
                int nchunks, int chunk_rows) {
    const int npad = ld;
    if (chunk_rows % TC_KB || npad % 128) return gk_fail(ctx, GK_EINVAL, "gk_tc_gfull: chunk_rows %% 32 and ld %% 128 must be 0");
    float *PhiT = (float *)work, *Ff = PhiT + (size_t)npad * 64;
    CUtensorMap mS, mP, mA;
    if (!tc_map_2d(&mS, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, S, (uint64_t)ld, (uint64_t)N, (uint64_t)ld * 4, TC_KB, 128, CU_TENSOR_MAP_SWIZZLE_128B) ||
        !tc_map_2d(&mP, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, PhiT, (uint64_t)npad, 64, (uint64_t)npad * 4, TC_KB, 64, CU_TENSOR_MAP_SWIZZLE_128B) ||
        !tc_map_2d(&mA, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, Asym, (uint64_t)ld, (uint64_t)N, (uint64_t)ld, 128, TC_KB, CU_TENSOR_MAP_SWIZZLE_NONE))
        return gk_fail(ctx, GK_ECUDA, "gk_tc_gfull: cuTensorMapEncodeTiled failed");
    k_tc_prep<<<(npad + 255) / 256, 256, 0, ctx->stream>>>(F, N, npad, PhiT, Ff);
    GK_CHECK_LAUNCH(ctx);
    const size_t smem = (size_t)TC_STAGES * TC_STAGE_BYTES + 1024;
    if (gk_attr_needed(ctx, (const void *)k_eig_gfull_tc))
        GK_CUDA(ctx, cudaFuncSetAttribute(k_eig_gfull_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    tc_params P;
    P.chunk_rows = chunk_rows; P.npad = npad; P.Ff = Ff; P.partial = partial;
    k_eig_gfull_tc<<<dim3(npad / 128, nchunks), TC_THREADS, smem, ctx->stream>>>(mS, mP, mA, P);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}
