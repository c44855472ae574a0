// gk_internal.cuh -- shared host/device declarations for libgk.so (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>
#include "../../include/gk.h"

#define GK_PAD 128            // columns of every N x N device array are padded to a multiple of this
#define GK_NFAMILY 8
#define GK_ATTR_STRIP 0                 // attr_done slots: 0..5 k_strip_stats<T,B>, 8.. eigen kernels, 32.. mean field

struct gk_timing {
    double ms = 0.0; int64_t launches = 0; double bytes = 0.0;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending; std::vector<double> pending_bytes;
};

struct gk_comm;
struct gk_ctx {
    gk_comm *comm = nullptr;                  // the engine's own NCCL communicator (gk_comm.cu), or none
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    int num_sms = 148;
    int64_t launches = 0;
    bool timing = false;
    gk_timing fam[GK_NFAMILY];
    std::string err;
    // scratch (grown on demand, freed with the ctx)
    void *scratch[12] = {nullptr}; size_t scratch_bytes[12] = {0};
    // K1 statistics of the labels the last unsharded gk_sbm_sweep ended on (gk_sbm_sweep_stats)
    struct { const void *net = nullptr; int K = 0, S = 0; bool valid = false; std::vector<double> cell; } last_cells;
    // state of an open sharded sweep
    struct gk_sweep *sweep = nullptr;
    // cudaFuncSetAttribute is per device: remembered per context, never in function-local statics
    bool attr_done[64] = {false};
    std::vector<const void *> attr_funcs;     // kernels whose attributes this context has set
    bool sbm_prof = false;                    // env GK_SBM_PROF, read once at context creation
    int sbm_overlap_mode = 3;                 // gk_ctx_set_option(GK_CTX_OPT_SBM_OVERLAP): 0 never, 1 block b+1 beside block b after a sweep with few movers, 2 the same always, 3 (default) = 1 + the next sweep's strip passes speculated in the background
    bool sbm_overlap = true;                  // env GK_SBM_NO_OVERLAP clears it: strip pass of block b+1 beside the serial kernel of block b
    cudaStream_t aux = nullptr; cudaEvent_t ev_fork = nullptr, ev_join = nullptr;   // second (low-priority) stream of that overlap, created on first use
    long long last_sweep_moves = -1; const void *last_sweep_net = nullptr;   // moves of the last unsharded gk_sbm_sweep (and on which network)
    // whole-sweep speculation of gk_sbm_sweep: strip passes of the next sweep running on `aux` from the labels in spec_labels
    struct { bool active = false, self = false; const void *net = nullptr; int K = 0, bn = 0, N = 0, nblk = 0;
             std::vector<int32_t> labels; std::vector<cudaEvent_t> ev; } spec;
    bool fork_after_lp = false;               // launch_sample records ev_fork between k_lp_from_stats and the serial kernel
};

struct gk_net {
    int N = 0, B = 0, dtype = GK_F32, row0 = 0, nrows = 0;
    int ld = 0;                       // padded column count (multiple of GK_PAD)
    uint8_t *A = nullptr, *At = nullptr;   // nrows x ld
    void *W = nullptr, *Wt = nullptr;      // nrows x ld x B of dtype
    uint8_t *diagA = nullptr;              // N   (A[n,n]; replicated on every rank)
    double *diagW = nullptr;               // N x B
    bool have_diag = false;
    size_t bytes = 0;
    // lazily built symmetric forms for the eigenmodel / latent-distance passes
    uint8_t *Asym = nullptr;               // nrows x ld : A + A^T
};

int gk_fail(gk_ctx *ctx, int code, const char *fmt, ...);
// gk_comm.cu: in-stream collectives on the context's communicator (no-ops for a world of one)
void gk_comm_release(gk_ctx *ctx);
int gk_comm_allreduce_f64(gk_ctx *ctx, double *buf_dev, size_t n);
int gk_comm_allreduce_i64(gk_ctx *ctx, long long *buf_dev, size_t n);
int gk_comm_broadcast(gk_ctx *ctx, void *buf_dev, size_t bytes, int root);
void gk_comm_row_range(int N, int rank, int world, int *r0, int *r1);
extern "C" int gk_comm_rank(gk_ctx *ctx);
extern "C" int gk_comm_world(gk_ctx *ctx);
// true the first time this context sees kernel `f` (cudaFuncSetAttribute is per device, so the memo is per context)
static inline bool gk_attr_needed(gk_ctx *ctx, const void *f) {
    for (const void *g : ctx->attr_funcs) if (g == f) return false;
    ctx->attr_funcs.push_back(f);
    return true;
}
int gk_net_refresh_asym(gk_ctx *ctx, gk_net *net);  // recomputes an existing net->Asym after a re-upload; gk_eig.cu
int gk_net_ensure_asym(gk_ctx *ctx, gk_net *net);   // builds net->Asym = A + A^T (zero diagonal) once; gk_eig.cu
void *gk_scratch(gk_ctx *ctx, int slot, size_t bytes);
void gk_spec_cancel(gk_ctx *ctx);                  // waits for and discards background strip passes (gk_sbm.cu speculation)
// gk_eig_tc.cu: the full Omega . (F (x) F) contraction on tcgen05 tensor cores fed by TMA (fp32 storage, D = 8)
size_t gk_tc_work_bytes(int npad);
int gk_tc_gfull(gk_ctx *ctx, const float *S, const uint8_t *Asym, int ld, int N, const double *F, void *work, float *partial,
                int nchunks, int chunk_rows);   // returns nullptr (and sets err) on failure

#define GK_CUDA(ctx, call)                                                                     \
    do {                                                                                       \
        cudaError_t _e = (call);                                                               \
        if (_e != cudaSuccess)                                                                 \
            return gk_fail((ctx), GK_ECUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call,         \
                           cudaGetErrorString(_e));                                            \
    } while (0)

#define GK_CHECK_LAUNCH(ctx)                                                                   \
    do {                                                                                       \
        (ctx)->launches++;                                                                     \
        GK_CUDA((ctx), cudaGetLastError());                                                    \
    } while (0)

// timing brackets around a kernel family (no-ops unless gk_ctx_kernel_timing(ctx, 1))
void gk_time_begin(gk_ctx *ctx, int family, double bytes);
void gk_time_end(gk_ctx *ctx, int family);

static inline int gk_pad(int n) { return (n + GK_PAD - 1) / GK_PAD * GK_PAD; }
static inline size_t gk_esize(int dtype) { return dtype == GK_F64 ? 8 : 4; }
