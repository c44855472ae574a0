// gk_core.cu -- context, error handling, timing, network upload and device-side transposes.
#include <stdlib.h>
#include <stdarg.h>
#include <string.h>
#include "gk_internal.cuh"

int gk_fail(gk_ctx *ctx, int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (ctx) ctx->err = buf;
    return code;
}

void gk_spec_cancel(gk_ctx *ctx) {
    if (!ctx || !ctx->spec.active) return;
    if (ctx->aux) cudaStreamSynchronize(ctx->aux);
    ctx->spec.active = false;
}

void *gk_scratch(gk_ctx *ctx, int slot, size_t bytes) {
    // slots 2, 3, 10, 11 may be in use by the background strip passes of a speculated sweep: any other user ends the speculation
    if (ctx->spec.active && !ctx->spec.self && (slot == 2 || slot == 3 || slot == 10 || slot == 11)) gk_spec_cancel(ctx);
    if (ctx->scratch_bytes[slot] >= bytes && ctx->scratch[slot]) return ctx->scratch[slot];
    if (ctx->scratch[slot]) { cudaStreamSynchronize(ctx->stream); cudaFree(ctx->scratch[slot]); ctx->scratch[slot] = nullptr; }
    size_t want = bytes + bytes / 4 + 256;
    if (cudaMalloc(&ctx->scratch[slot], want) != cudaSuccess) {
        ctx->scratch_bytes[slot] = 0; ctx->scratch[slot] = nullptr;
        gk_fail(ctx, GK_ENOMEM, "scratch slot %d: cudaMalloc(%zu) failed", slot, want);
        return nullptr;
    }
    ctx->scratch_bytes[slot] = want;
    return ctx->scratch[slot];
}

void gk_time_begin(gk_ctx *ctx, int family, double bytes) {
    if (!ctx->timing) return;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a, ctx->stream);
    ctx->fam[family].pending.push_back({a, b});
    ctx->fam[family].pending_bytes.push_back(bytes);
}
void gk_time_end(gk_ctx *ctx, int family) {
    if (!ctx->timing) return;
    cudaEventRecord(ctx->fam[family].pending.back().second, ctx->stream);
}

extern "C" int gk_ctx_create(int device, gk_ctx **out) {
    if (!out) return GK_EINVAL;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return GK_ECUDA;   // no CPU fallback: fail loudly
    if (device < 0 || device >= ndev) return GK_EINVAL;
    if (cudaSetDevice(device) != cudaSuccess) return GK_ECUDA;
    gk_ctx *ctx = new gk_ctx();
    ctx->device = device;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) ctx->num_sms = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return GK_ECUDA; }
    ctx->own_stream = true;
    ctx->sbm_prof = getenv("GK_SBM_PROF") != nullptr;
    ctx->sbm_overlap = getenv("GK_SBM_NO_OVERLAP") == nullptr;
    if (!ctx->sbm_overlap) ctx->sbm_overlap_mode = 0;
    if (const char *mo = getenv("GK_SBM_OVERLAP_MODE")) { const int v = atoi(mo); if (v >= 0 && v <= 3) ctx->sbm_overlap_mode = v; }
    *out = ctx;
    return GK_OK;
}

extern "C" void gk_ctx_destroy(gk_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    gk_comm_release(ctx);
    if (ctx->aux) { cudaStreamSynchronize(ctx->aux); cudaStreamDestroy(ctx->aux); }
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    for (cudaEvent_t e : ctx->spec.ev) cudaEventDestroy(e);
    for (int i = 0; i < 12; ++i) if (ctx->scratch[i]) cudaFree(ctx->scratch[i]);
    for (auto &f : ctx->fam) for (auto &p : f.pending) { cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" int gk_ctx_set_option(gk_ctx *ctx, int option, int value) {
    if (!ctx) return GK_EINVAL;
    if (option == GK_CTX_OPT_SBM_OVERLAP && value >= 0 && value <= 3) { if (value != 3) gk_spec_cancel(ctx); ctx->sbm_overlap_mode = value; return GK_OK; }
    return gk_fail(ctx, GK_EINVAL, "gk_ctx_set_option: unknown option %d or value %d", option, value);
}

extern "C" const char *gk_last_error(gk_ctx *ctx) { return ctx ? ctx->err.c_str() : "null ctx"; }

extern "C" int gk_ctx_set_stream(gk_ctx *ctx, void *s) {
    if (!ctx) return GK_EINVAL;
    cudaStreamSynchronize(ctx->stream);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    ctx->stream = (cudaStream_t)s; ctx->own_stream = false;
    return GK_OK;
}

extern "C" int gk_ctx_synchronize(gk_ctx *ctx) {
    if (!ctx) return GK_EINVAL;
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GK_OK;
}

extern "C" int64_t gk_ctx_launch_count(gk_ctx *ctx) { return ctx ? ctx->launches : -1; }

extern "C" int gk_ctx_kernel_timing(gk_ctx *ctx, int enable) {
    if (!ctx) return GK_EINVAL;
    ctx->timing = enable != 0;
    return GK_OK;
}

extern "C" int gk_ctx_kernel_timing_read(gk_ctx *ctx, int family, double *ms, int64_t *launches, double *bytes) {
    if (!ctx || family < 0 || family >= GK_NFAMILY) return GK_EINVAL;
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->aux) GK_CUDA(ctx, cudaStreamSynchronize(ctx->aux));      // kernels of the overlapped / speculated strip passes
    gk_timing &f = ctx->fam[family];
    for (size_t i = 0; i < f.pending.size(); ++i) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, f.pending[i].first, f.pending[i].second) == cudaSuccess) {
            f.ms += t; f.launches++; f.bytes += f.pending_bytes[i];
        }
        cudaEventDestroy(f.pending[i].first); cudaEventDestroy(f.pending[i].second);
    }
    f.pending.clear(); f.pending_bytes.clear();
    if (ms) *ms = f.ms;
    if (launches) *launches = f.launches;
    if (bytes) *bytes = f.bytes;
    f.ms = 0; f.launches = 0; f.bytes = 0;
    return GK_OK;
}

// ---------------------------------------------------------------------------------------------
// device helpers: dtype conversion and tiled transposes (run once per upload)
// ---------------------------------------------------------------------------------------------
template <typename S, typename D>
__global__ void k_convert_rows(const S *__restrict__ src, D *__restrict__ dst, int nrows, size_t row_elems,
                               size_t dst_pitch_elems) {
    size_t total = (size_t)nrows * row_elems;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        size_t r = i / row_elems, cidx = i - r * row_elems;
        dst[r * dst_pitch_elems + cidx] = (D)src[i];
    }
}

// out[j][i] = in[i][j] for an n x n matrix of B-vectors with padded pitch ld (elements)
template <typename T>
__global__ void k_transpose(const T *__restrict__ in, T *__restrict__ out, int n, int ld, int B) {
    __shared__ T tile[32][33];
    int bx = blockIdx.x * 32, by = blockIdx.y * 32;
    for (int b = 0; b < B; ++b) {
        for (int r = threadIdx.y; r < 32; r += blockDim.y) {
            int i = by + r, j = bx + threadIdx.x;
            tile[r][threadIdx.x] = (i < n && j < n) ? in[((size_t)i * ld + j) * B + b] : T(0);
        }
        __syncthreads();
        for (int r = threadIdx.y; r < 32; r += blockDim.y) {
            int j = bx + r, i = by + threadIdx.x;
            if (i < n && j < n) out[((size_t)j * ld + i) * B + b] = tile[threadIdx.x][r];
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------
// network
// ---------------------------------------------------------------------------------------------
extern "C" int gk_network_create(gk_ctx *ctx, int N, int B, int dev_dtype, int row0, int nrows, gk_net **out) {
    if (!ctx || !out) return GK_EINVAL;
    *out = nullptr;
    if (N <= 0 || B < 0 || row0 < 0 || nrows <= 0 || row0 + nrows > N) return gk_fail(ctx, GK_EINVAL, "bad network shape");
    if (dev_dtype != GK_F32 && dev_dtype != GK_F64) return gk_fail(ctx, GK_EINVAL, "bad dtype");
    if (B > 4) return gk_fail(ctx, GK_EUNSUPPORTED, "B=%d: this build compiles weight dimensions B in {0,...,4}", B);
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    gk_net *net = new gk_net();
    net->N = N; net->B = B; net->dtype = dev_dtype; net->row0 = row0; net->nrows = nrows; net->ld = gk_pad(N);
    size_t abytes = (size_t)nrows * net->ld, wbytes = abytes * B * gk_esize(dev_dtype);
    cudaError_t e = cudaMalloc(&net->A, abytes);
    if (e == cudaSuccess) e = cudaMalloc(&net->At, abytes);
    if (e == cudaSuccess && B) e = cudaMalloc(&net->W, wbytes);
    if (e == cudaSuccess && B) e = cudaMalloc(&net->Wt, wbytes);
    if (e == cudaSuccess) e = cudaMalloc(&net->diagA, (size_t)N);
    if (e == cudaSuccess) e = cudaMalloc(&net->diagW, sizeof(double) * N * (B ? B : 1));
    if (e != cudaSuccess) { gk_network_destroy(ctx, net); return gk_fail(ctx, GK_ENOMEM, "network alloc: %s", cudaGetErrorString(e)); }
    net->bytes = 2 * abytes + 2 * wbytes + N + sizeof(double) * N * (B ? B : 1);
    cudaMemsetAsync(net->diagA, 0, N, ctx->stream);
    cudaMemsetAsync(net->diagW, 0, sizeof(double) * N * (B ? B : 1), ctx->stream);
    *out = net;
    return GK_OK;
}

extern "C" int gk_network_destroy(gk_ctx *ctx, gk_net *net) {
    if (ctx && ctx->last_cells.net == net) { ctx->last_cells.valid = false; ctx->last_cells.net = nullptr; }
    if (ctx && ctx->spec.net == (const void *)net) { gk_spec_cancel(ctx); ctx->spec.net = nullptr; }
    if (!net) return GK_OK;
    if (ctx) { cudaSetDevice(ctx->device); cudaStreamSynchronize(ctx->stream); }
    cudaFree(net->A); cudaFree(net->At); cudaFree(net->W); cudaFree(net->Wt);
    cudaFree(net->diagA); cudaFree(net->diagW); cudaFree(net->Asym);
    delete net;
    return GK_OK;
}

extern "C" size_t gk_network_device_bytes(gk_net *net) { return net ? net->bytes : 0; }

static int upload_w(gk_ctx *ctx, gk_net *net, void *dst, const void *src, int w_dtype) {
    size_t row_elems = (size_t)net->N * net->B, pitch_elems = (size_t)net->ld * net->B;
    size_t des = gk_esize(net->dtype), ses = gk_esize(w_dtype);
    GK_CUDA(ctx, cudaMemsetAsync(dst, 0, (size_t)net->nrows * pitch_elems * des, ctx->stream));
    if (w_dtype == net->dtype) {
        GK_CUDA(ctx, cudaMemcpy2DAsync(dst, pitch_elems * des, src, row_elems * ses, row_elems * ses, net->nrows,
                                       cudaMemcpyHostToDevice, ctx->stream));
        return GK_OK;
    }
    size_t sbytes = (size_t)net->nrows * row_elems * ses;
    void *tmp = gk_scratch(ctx, 0, sbytes);
    if (!tmp) return GK_ENOMEM;
    GK_CUDA(ctx, cudaMemcpyAsync(tmp, src, sbytes, cudaMemcpyHostToDevice, ctx->stream));
    int grid = ctx->num_sms * 8;
    if (w_dtype == GK_F64) k_convert_rows<double, float><<<grid, 256, 0, ctx->stream>>>((const double *)tmp, (float *)dst, net->nrows, row_elems, pitch_elems);
    else k_convert_rows<float, double><<<grid, 256, 0, ctx->stream>>>((const float *)tmp, (double *)dst, net->nrows, row_elems, pitch_elems);
    GK_CHECK_LAUNCH(ctx);
    return GK_OK;
}

extern "C" int gk_network_upload(gk_ctx *ctx, gk_net *net, const uint8_t *A_rows, const void *W_rows, int w_dtype,
                                 const uint8_t *At_rows, const void *Wt_rows) {
    if (ctx && ctx->last_cells.net == net) ctx->last_cells.valid = false;      // statistics of the previous data
    if (ctx && ctx->last_sweep_net == (const void *)net) ctx->last_sweep_moves = -1;
    if (ctx && ctx->spec.net == (const void *)net) gk_spec_cancel(ctx);
    if (!ctx || !net || !A_rows) return GK_EINVAL;
    if (net->B && !W_rows) return gk_fail(ctx, GK_EINVAL, "W required when B > 0");
    bool whole = net->row0 == 0 && net->nrows == net->N;
    if (!At_rows && !whole) return gk_fail(ctx, GK_EINVAL, "row-sharded upload needs the transposed shard (At_rows/Wt_rows)");
    GK_CUDA(ctx, cudaSetDevice(ctx->device));
    int N = net->N, ld = net->ld, rc;
    size_t abytes = (size_t)net->nrows * ld;
    GK_CUDA(ctx, cudaMemsetAsync(net->A, 0, abytes, ctx->stream));
    GK_CUDA(ctx, cudaMemsetAsync(net->At, 0, abytes, ctx->stream));
    GK_CUDA(ctx, cudaMemcpy2DAsync(net->A, ld, A_rows, N, N, net->nrows, cudaMemcpyHostToDevice, ctx->stream));
    if (net->B && (rc = upload_w(ctx, net, net->W, W_rows, w_dtype)) != GK_OK) return rc;
    if (At_rows) {
        GK_CUDA(ctx, cudaMemcpy2DAsync(net->At, ld, At_rows, N, N, net->nrows, cudaMemcpyHostToDevice, ctx->stream));
        if (net->B) {
            if (!Wt_rows) return gk_fail(ctx, GK_EINVAL, "Wt_rows required with At_rows");
            if ((rc = upload_w(ctx, net, net->Wt, Wt_rows, w_dtype)) != GK_OK) return rc;
        }
    } else {
        dim3 blk(32, 8), grd((N + 31) / 32, (N + 31) / 32);
        k_transpose<uint8_t><<<grd, blk, 0, ctx->stream>>>(net->A, net->At, N, ld, 1);
        GK_CHECK_LAUNCH(ctx);
        if (net->B) {
            GK_CUDA(ctx, cudaMemsetAsync(net->Wt, 0, abytes * net->B * gk_esize(net->dtype), ctx->stream));
            if (net->dtype == GK_F32) k_transpose<float><<<grd, blk, 0, ctx->stream>>>((const float *)net->W, (float *)net->Wt, N, ld, net->B);
            else k_transpose<double><<<grd, blk, 0, ctx->stream>>>((const double *)net->W, (double *)net->Wt, N, ld, net->B);
            GK_CHECK_LAUNCH(ctx);
        }
    }
    if (net->Asym) {                                     // a re-upload into a live handle: refresh A + A^T in place
        int rc = gk_net_refresh_asym(ctx, net);
        if (rc) return rc;
    }
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));   // host buffers may be reused by the caller
    return GK_OK;
}

extern "C" int gk_network_set_diag(gk_ctx *ctx, gk_net *net, const uint8_t *diagA, const double *diagW) {
    if (!ctx || !net || !diagA) return GK_EINVAL;
    GK_CUDA(ctx, cudaMemcpyAsync(net->diagA, diagA, net->N, cudaMemcpyHostToDevice, ctx->stream));
    if (net->B) {
        if (!diagW) return gk_fail(ctx, GK_EINVAL, "diagW required when B > 0");
        GK_CUDA(ctx, cudaMemcpyAsync(net->diagW, diagW, sizeof(double) * net->N * net->B, cudaMemcpyHostToDevice, ctx->stream));
    }
    GK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    net->have_diag = true;
    return GK_OK;
}
