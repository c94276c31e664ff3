// bwagpu_bsw_get_scores_noalloc (include/bwagpu.h): the banded-SW extension seam for callers that
// must not allocate — code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc.  No counterpart in the reference.
//
// Same per-pair arithmetic as bwagpu_bsw_get_scores_device (bsw_lane_pair / bsw_scalar_pair of
// bsw_thread.cuh: one SIMD lane of the reference's getScores16 / getScores8, or scalarBandedSWA),
// but the per-thread DP column state lives in ONE file-scope device array of fixed size instead of
// the context's grow-only cudaMalloc'ed scratch.  Every GPU that loads the library holds the array
// until the process exits.  All calls on a GPU share it: two calls that overlap on that GPU (other
// streams, other contexts) race on it.  Nothing in it has to start at zero — every pair initialises
// the columns it later reads (first-row set-up of bsw_lane_pair / bsw_scalar_pair) — so the contents
// left by the previous call are harmless and no clearing pass is needed.
#include <mutex>
#include "ctx.cuh"
namespace {   // private copy of the per-pair device code: bsw.cu owns the externally visible one
#include "bsw_thread.cuh"
}

#define NOALLOC_TPB 128
#define NOALLOC_WORDS ((int64_t)BWAGPU_NOALLOC_SCRATCH_BYTES / 4)
#define NOALLOC_MAX_DEV 64

static __device__ __align__(256) uint32_t g_noalloc_scratch[NOALLOC_WORDS];

// g_noalloc_scratch is, in host code, only a stand-in for the symbol: the device address is asked
// for once per device (with that device current) and kept here.
static uint32_t *g_noalloc_ptr[NOALLOC_MAX_DEV];
static std::mutex g_noalloc_mu;

template <int KIND>
__global__ void __launch_bounds__(NOALLOC_TPB)
bsw_noalloc_kernel(bwagpu_bsw_params prm, bwagpu_seqpair *__restrict__ pairs,
                   const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer, int n, int max_len2,
                   int w, uint32_t *__restrict__ work)
{
    const int64_t T = (int64_t)gridDim.x * blockDim.x;
    const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t *col = work + t0;                        // word j of thread t at work[j * T + t]
    for (int64_t i = t0; i < n; i += T) {
        const bwagpu_seqpair sp = pairs[i];
        BswOut o;
        if (sp.len2 > max_len2 || sp.len2 < 0 || sp.len1 < 0) {
            o.score = o.qle = o.tle = o.gtle = o.gscore = o.max_off = -1;   // outside the stated bound: not computed
        } else if (KIND == BWAGPU_BSW_SCALAR) {
            o = bsw_scalar_pair(prm, qer + sp.idq, sp.len2, ref + sp.idr, sp.len1, w, sp.h0, col, T);
        } else {
            o = bsw_lane_pair<KIND>(prm, qer + sp.idq, sp.len2, ref + sp.idr, sp.len1, w, sp.h0, col, T);
        }
        bwagpu_seqpair *d = &pairs[i];
        d->score = o.score; d->tle = o.tle; d->gtle = o.gtle; d->qle = o.qle;
        d->gscore = o.gscore; d->max_off = o.max_off;
    }
}

static int noalloc_scratch_ptr(bwagpu_ctx *ctx, uint32_t **out)
{
    if (ctx->device < 0 || ctx->device >= NOALLOC_MAX_DEV)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "bsw_noalloc: device %d outside [0,%d)", ctx->device, NOALLOC_MAX_DEV);
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[ctx->device]) {
        void *p = nullptr;
        BWAGPU_CUDA(ctx, cudaGetSymbolAddress(&p, g_noalloc_scratch));    // ctx->device is current (caller)
        if (!p) BWAGPU_FAIL(ctx, BWAGPU_ERR_CUDA, "bsw_noalloc: cudaGetSymbolAddress returned NULL");
        g_noalloc_ptr[ctx->device] = (uint32_t *)p;
    }
    *out = g_noalloc_ptr[ctx->device];
    return BWAGPU_OK;
}

extern "C" int bwagpu_bsw_get_scores_noalloc(bwagpu_ctx *ctx, const bwagpu_bsw_params *prm, int kind,
                                             bwagpu_seqpair *d_pairArray, const uint8_t *d_seqBufRef,
                                             const uint8_t *d_seqBufQer, int32_t numPairs, int32_t max_len2,
                                             int32_t w)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!prm) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "bsw_noalloc: params is NULL");
    if (kind != BWAGPU_BSW_SCALAR && kind != BWAGPU_BSW_8BIT && kind != BWAGPU_BSW_16BIT)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "bsw_noalloc: kind must be 0, 8 or 16 (got %d)", kind);
    if (numPairs < 0 || w < 0 || max_len2 < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "bsw_noalloc: negative numPairs, max_len2 or w");
    if (prm->e_del <= 0 || prm->e_ins <= 0)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "bsw_noalloc: gap extension penalties must be positive");
    if (max_len2 > BWAGPU_NOALLOC_MAX_LEN2)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE,
                    "bsw_noalloc: max_len2 = %d exceeds the fixed scratch (%d MiB static array, queries up to %d bases); "
                    "use bwagpu_bsw_get_scores_device for longer queries",
                    max_len2, (int)(BWAGPU_NOALLOC_SCRATCH_BYTES >> 20), BWAGPU_NOALLOC_MAX_LEN2);
    if (numPairs == 0) return BWAGPU_OK;
    if (!d_pairArray || !d_seqBufRef || !d_seqBufQer)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "bsw_noalloc: NULL buffer with numPairs=%d", numPairs);
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    uint32_t *work;
    int rc = noalloc_scratch_ptr(ctx, &work);
    if (rc) return rc;
    // threads = what the array holds at this query length (never more than the pairs need)
    const int64_t cols = (int64_t)(kind == BWAGPU_BSW_SCALAR ? 2 : 1) * (max_len2 + 4);
    int64_t blocks = NOALLOC_WORDS / cols / NOALLOC_TPB;
    const int64_t need = ((int64_t)numPairs + NOALLOC_TPB - 1) / NOALLOC_TPB;
    if (blocks > need) blocks = need;
    if (blocks > (int64_t)ctx->sm_count * 8) blocks = (int64_t)ctx->sm_count * 8;
    if (blocks < 1) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "bsw_noalloc: scratch too small for max_len2 = %d", max_len2);
    switch (kind) {
    case BWAGPU_BSW_16BIT:
        bsw_noalloc_kernel<16><<<(int)blocks, NOALLOC_TPB, 0, ctx->stream>>>(*prm, d_pairArray, d_seqBufRef, d_seqBufQer,
                                                                           numPairs, max_len2, w, work);
        break;
    case BWAGPU_BSW_8BIT:
        bsw_noalloc_kernel<8><<<(int)blocks, NOALLOC_TPB, 0, ctx->stream>>>(*prm, d_pairArray, d_seqBufRef, d_seqBufQer,
                                                                          numPairs, max_len2, w, work);
        break;
    default:
        bsw_noalloc_kernel<0><<<(int)blocks, NOALLOC_TPB, 0, ctx->stream>>>(*prm, d_pairArray, d_seqBufRef, d_seqBufQer,
                                                                          numPairs, max_len2, w, work);
    }
    ctx->launches++;
    BWAGPU_CUDA(ctx, cudaGetLastError());
    return BWAGPU_OK;
}
