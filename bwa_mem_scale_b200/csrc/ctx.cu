// Context, memory and timing entry points of the C ABI (include/bwagpu.h).
#include "ctx.cuh"
#include <stdlib.h>

static char g_err[512] = "";

extern "C" int bwagpu_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" int bwagpu_create(int device, bwagpu_ctx **out)
{
    if (!out) return BWAGPU_ERR_ARG;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        snprintf(g_err, sizeof g_err, "no CUDA device available (%s); libbwagpu has no CPU path",
                 e == cudaSuccess ? "0 devices" : cudaGetErrorString(e));
        cudaGetLastError();
        return BWAGPU_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= n) {
        snprintf(g_err, sizeof g_err, "device %d out of range [0,%d)", device, n);
        return BWAGPU_ERR_ARG;
    }
    bwagpu_ctx *ctx = (bwagpu_ctx *)calloc(1, sizeof(bwagpu_ctx));
    if (!ctx) return BWAGPU_ERR_NOMEM;
    ctx->device = device;
    ctx->err[0] = 0;
#define CK(call)                                                                         \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess) {                                                        \
            snprintf(g_err, sizeof g_err, "%s -> %s", #call, cudaGetErrorString(e__));   \
            free(ctx);                                                                   \
            return BWAGPU_ERR_CUDA;                                                      \
        }                                                                                \
    } while (0)
    CK(cudaSetDevice(device));
    CK(cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device));
    if (getenv("BWAGPU_L2_FETCH")) {                           // experiments: L2 fetch granularity hint (32, 64 or 128 bytes)
        const int v = atoi(getenv("BWAGPU_L2_FETCH"));
        if (v == 32 || v == 64 || v == 128) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)v);
    }
    CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&ctx->ev_start));
    CK(cudaEventCreate(&ctx->ev_stop));
    CK(cudaEventCreate(&ctx->kev0));
    CK(cudaEventCreate(&ctx->kev1));
    CK(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&ctx->ev_block, cudaEventBlockingSync | cudaEventDisableTiming));
    ctx->blocking_sync = getenv("BWAGPU_BLOCKING_SYNC") && atoi(getenv("BWAGPU_BLOCKING_SYNC")) > 0;
    CK(cudaEventCreateWithFlags(&ctx->ev_main_done, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&ctx->ev_h2d_done, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&ctx->ev_d2h_done, cudaEventDisableTiming));
    ctx->d2h_pending_slot = -1;
#undef CK
    *out = ctx;
    return BWAGPU_OK;
}

extern "C" void bwagpu_destroy(bwagpu_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    bwagpu_stream_sync(ctx);
    for (int i = 0; i < BWAGPU_NSCRATCH; ++i)
        if (ctx->scratch[i]) cudaFree(ctx->scratch[i]);
    if (ctx->index_borrowed)
        ctx->d_cp_occ = ctx->d_sa_ms = ctx->d_sa_ls = ctx->d_ref_string = ctx->d_fma_all = ctx->d_fma_last = ctx->d_emf_loc =
            ctx->d_emf_seed = ctx->d_ert_kmer = ctx->d_ert_mlt = nullptr;
    if (ctx->d_cp_occ) cudaFree(ctx->d_cp_occ);
    if (ctx->d_sa_ms) cudaFree(ctx->d_sa_ms);
    if (ctx->d_sa_ls) cudaFree(ctx->d_sa_ls);
    if (ctx->d_ref_string) { cudaFree(ctx->d_ref_string); cudaFree(ctx->d_contig_off); cudaFree(ctx->d_contig_len); cudaFree(ctx->d_contig_alt); }
    if (ctx->d_fma_all) cudaFree(ctx->d_fma_all);
    if (ctx->d_fma_last) cudaFree(ctx->d_fma_last);
    if (ctx->d_emf_loc) cudaFree(ctx->d_emf_loc);
    if (ctx->d_emf_seed) cudaFree(ctx->d_emf_seed);
    if (ctx->d_ert_kmer) cudaFree(ctx->d_ert_kmer);
    if (ctx->d_ert_mlt) cudaFree(ctx->d_ert_mlt);
    cudaEventDestroy(ctx->ev_block);
    cudaEventDestroy(ctx->ev_start);
    cudaEventDestroy(ctx->ev_stop);
    cudaEventDestroy(ctx->kev0);
    cudaEventDestroy(ctx->kev1);
    cudaStreamSynchronize(ctx->copy_stream);
    cudaEventDestroy(ctx->ev_main_done); cudaEventDestroy(ctx->ev_h2d_done); cudaEventDestroy(ctx->ev_d2h_done);
    cudaStreamDestroy(ctx->copy_stream);
    cudaStreamDestroy(ctx->stream);
    free(ctx);
}

// A second context on the same GPU — its own stream, scratch and error state — that works on the index replica held by
// `owner`: what a host with several worker threads per GPU uses, so that one thread's copies and host-side waits overlap
// another's kernels without a second copy of the index in HBM.  The borrower must be destroyed before the owner and can
// not load, build or unload index parts itself.  Options that live with the index (FMA on/off) are copied as they are now.
extern "C" int bwagpu_ctx_share_index(bwagpu_ctx *ctx, const bwagpu_ctx *owner)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!owner || owner == ctx) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "ctx_share_index: owner is NULL or the context itself");
    if (owner->device != ctx->device) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "ctx_share_index: contexts are on different GPUs (%d, %d)", ctx->device, owner->device);
    if (owner->index_borrowed) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "ctx_share_index: the owner itself borrows its index");
    if (!ctx->index_borrowed && (ctx->d_cp_occ || ctx->d_ref_string || ctx->d_fma_all || ctx->d_fma_last || ctx->d_emf_seed || ctx->d_ert_kmer))
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "ctx_share_index: this context already holds index parts of its own");
    ctx->index_borrowed = 1;
    ctx->has_fm = owner->has_fm;
    ctx->d_cp_occ = owner->d_cp_occ; ctx->d_sa_ms = owner->d_sa_ms; ctx->d_sa_ls = owner->d_sa_ls;
    for (int i = 0; i < 5; ++i) ctx->fm_count[i] = owner->fm_count[i];
    ctx->fm_sentinel = owner->fm_sentinel; ctx->fm_ref_seq_len = owner->fm_ref_seq_len;
    ctx->d_fma_all = owner->d_fma_all; ctx->d_fma_last = owner->d_fma_last; ctx->fma_enabled = owner->fma_enabled;
    ctx->d_ref_string = owner->d_ref_string; ctx->d_contig_off = owner->d_contig_off; ctx->d_contig_len = owner->d_contig_len;
    ctx->d_contig_alt = owner->d_contig_alt; ctx->l_pac = owner->l_pac; ctx->n_contigs = owner->n_contigs;
    ctx->d_emf_loc = owner->d_emf_loc; ctx->d_emf_seed = owner->d_emf_seed; ctx->emf = owner->emf;
    ctx->d_ert_kmer = owner->d_ert_kmer; ctx->d_ert_mlt = owner->d_ert_mlt; ctx->ert_mlt_bytes = owner->ert_mlt_bytes;
    return BWAGPU_OK;
}

extern "C" const char *bwagpu_last_error(const bwagpu_ctx *ctx) { return ctx ? ctx->err : g_err; }

extern "C" int64_t bwagpu_launch_count(const bwagpu_ctx *ctx) { return ctx ? ctx->launches : 0; }

extern "C" int bwagpu_set_profiling(bwagpu_ctx *ctx, int on)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    ctx->profiling = on ? 1 : 0;
    for (int i = 0; i < BWAGPU_KERN_N; ++i) { ctx->kern_ms[i] = 0; ctx->kern_n[i] = 0; }
    return BWAGPU_OK;
}

extern "C" int bwagpu_kernel_time(const bwagpu_ctx *ctx, int which, double *ms, int64_t *launches)
{
    if (!ctx || which < 0 || which >= BWAGPU_KERN_N) return BWAGPU_ERR_ARG;
    if (ms) *ms = ctx->kern_ms[which];
    if (launches) *launches = ctx->kern_n[which];
    return BWAGPU_OK;
}

extern "C" int bwagpu_synchronize(bwagpu_ctx *ctx)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}

extern "C" int bwagpu_timer_start(bwagpu_ctx *ctx)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, cudaEventRecord(ctx->ev_start, ctx->stream));
    return BWAGPU_OK;
}

extern "C" int bwagpu_timer_stop(bwagpu_ctx *ctx, float *elapsed_ms)
{
    if (!ctx || !elapsed_ms) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, cudaEventRecord(ctx->ev_stop, ctx->stream));
    BWAGPU_CUDA(ctx, cudaEventSynchronize(ctx->ev_stop));
    BWAGPU_CUDA(ctx, cudaEventElapsedTime(elapsed_ms, ctx->ev_start, ctx->ev_stop));
    return BWAGPU_OK;
}

extern "C" int bwagpu_malloc(bwagpu_ctx *ctx, int64_t bytes, void **dptr)
{
    if (!ctx || !dptr || bytes < 0) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, cudaMalloc(dptr, (size_t)(bytes ? bytes : 1)));
    return BWAGPU_OK;
}

extern "C" int bwagpu_free(bwagpu_ctx *ctx, void *dptr)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    BWAGPU_CUDA(ctx, cudaFree(dptr));
    return BWAGPU_OK;
}

extern "C" int bwagpu_h2d(bwagpu_ctx *ctx, void *dst, const void *src, int64_t bytes)
{
    if (!ctx || bytes < 0) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}

extern "C" int bwagpu_d2h(bwagpu_ctx *ctx, void *dst, const void *src, int64_t bytes)
{
    if (!ctx || bytes < 0) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}

extern "C" int bwagpu_host_alloc(bwagpu_ctx *ctx, int64_t bytes, void **hptr)
{
    if (!ctx || !hptr || bytes < 0) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, cudaHostAlloc(hptr, (size_t)(bytes ? bytes : 1), cudaHostAllocPortable));
    return BWAGPU_OK;
}

extern "C" int bwagpu_host_free(bwagpu_ctx *ctx, void *hptr)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (hptr) BWAGPU_CUDA(ctx, cudaFreeHost(hptr));
    return BWAGPU_OK;
}

extern "C" int bwagpu_h2d_async(bwagpu_ctx *ctx, void *dst, const void *src, int64_t bytes)
{
    if (!ctx || bytes < 0) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyHostToDevice, ctx->stream));
    return BWAGPU_OK;
}

extern "C" int bwagpu_d2h_async(bwagpu_ctx *ctx, void *dst, const void *src, int64_t bytes)
{
    if (!ctx || bytes < 0) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return BWAGPU_OK;
}

// ---- copies on the second stream ----
extern "C" int bwagpu_h2d_overlap(bwagpu_ctx *ctx, void *dst, const void *src, int64_t bytes)
{
    if (!ctx || bytes < 0) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    // after everything already enqueued on the main stream (the batch that last used dst), before what comes next
    BWAGPU_CUDA(ctx, cudaEventRecord(ctx->ev_main_done, ctx->stream));
    BWAGPU_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_main_done, 0));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyHostToDevice, ctx->copy_stream));
    BWAGPU_CUDA(ctx, cudaEventRecord(ctx->ev_h2d_done, ctx->copy_stream));
    ctx->h2d_pending = 1;
    return BWAGPU_OK;
}

extern "C" int bwagpu_h2d_fence(bwagpu_ctx *ctx)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (ctx->h2d_pending) {
        BWAGPU_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_h2d_done, 0));
        ctx->h2d_pending = 0;
    }
    return BWAGPU_OK;
}

extern "C" int bwagpu_d2h_overlap(bwagpu_ctx *ctx, void *dst, const void *src, int64_t bytes)
{
    if (!ctx || bytes < 0) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ctx->d2h_pending_slot >= 0) {          // one result buffer in flight at a time
        BWAGPU_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_d2h_done, 0));
        ctx->d2h_pending_slot = -1;
    }
    BWAGPU_CUDA(ctx, cudaEventRecord(ctx->ev_main_done, ctx->stream));
    BWAGPU_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_main_done, 0));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToHost, ctx->copy_stream));
    BWAGPU_CUDA(ctx, cudaEventRecord(ctx->ev_d2h_done, ctx->copy_stream));
    for (int i = 0; i < BWAGPU_NSCRATCH; ++i) {
        const char *b = (const char *)ctx->scratch[i];
        if (b && (const char *)src >= b && (const char *)src < b + ctx->scratch_bytes[i]) { ctx->d2h_pending_slot = i; break; }
    }
    return BWAGPU_OK;
}

extern "C" int bwagpu_copy_synchronize(bwagpu_ctx *ctx)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, cudaStreamSynchronize(ctx->copy_stream));
    return BWAGPU_OK;
}

extern "C" int bwagpu_memset(bwagpu_ctx *ctx, void *dst, int value, int64_t bytes)
{
    if (!ctx || bytes < 0) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(dst, value, (size_t)bytes, ctx->stream));
    return BWAGPU_OK;
}

int bwagpu_scratch(bwagpu_ctx *ctx, int slot, int64_t bytes, void **out)
{
    if (slot == ctx->d2h_pending_slot) {      // an overlapped D2H still reads this buffer: order the stream after it
        BWAGPU_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_d2h_done, 0));
        ctx->d2h_pending_slot = -1;
    }
    if (bytes > ctx->scratch_bytes[slot]) {
        if (ctx->scratch[slot]) {
            BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
            BWAGPU_CUDA(ctx, cudaFree(ctx->scratch[slot]));
            ctx->scratch[slot] = nullptr;
            ctx->scratch_bytes[slot] = 0;
        }
        int64_t want = bytes + bytes / 4 + 4096;
        BWAGPU_CUDA(ctx, cudaMalloc(&ctx->scratch[slot], (size_t)want));
        ctx->scratch_bytes[slot] = want;
    }
    *out = ctx->scratch[slot];
    return BWAGPU_OK;
}
