// S2 extension seam of the C ABI: bwagpu_bsw_get_scores{,_device}  (include/bwagpu.h).
// Replaces BandedPairWiseSW::getScores16 / getScores8 / scalarBandedSWAWrapper
// (reference src/bandedSWA.h:130-172).
#include <cub/device/device_radix_sort.cuh>
#include <stdlib.h>
#include "ctx.cuh"
#include "bsw_thread.cuh"
#include "bsw_fast.cuh"

// DP cells computed (sum over rows of hi - lo), accumulated per warp into the context's counter: the GCUPS numerator
__device__ __forceinline__ void bsw_add_cells(unsigned long long *cells, int mine)
{
    const unsigned m = __activemask();
    const unsigned tot = __reduce_add_sync(m, (unsigned)mine);
    if ((threadIdx.x & 31) == (__ffs(m) - 1) && tot) atomicAdd(cells, (unsigned long long)tot);
}

// Where a pair's sequences are.  Arena mode (the C ABI's SeqPair contract): bytes ascend from seqBufQer + idq and
// seqBufRef + idr.  Direct mode (internal, chain2aln): `qer` is the chunk's read buffer and `ref` the resident reference
// text; idq / idr hold the low 32 bits of the first byte's position, regid bit 0 says "descending" (a left extension
// walks the read and the text backwards) and regid >> 8 carries the text position's bits above 32.
struct BswView { const uint8_t *q, *t; int dir; };
__device__ __forceinline__ BswView bsw_pair_view(const bwagpu_seqpair &sp, const uint8_t *ref, const uint8_t *qer, int direct)
{
    BswView v;
    if (!direct) { v.q = qer + sp.idq; v.t = ref + sp.idr; v.dir = 1; }
    else {
        v.q = qer + (uint32_t)sp.idq;
        v.t = ref + (int64_t)(((uint64_t)((uint32_t)sp.regid >> 8) << 32) | (uint32_t)sp.idr);
        v.dir = (sp.regid & 1) ? -1 : 1;
    }
    return v;
}

#define BSW_TPB 128
#define BSW_FAST_TPB 64          // threads per block of the shared-memory kernel
#define BSW_FAST_MAX_LEN2 512    // longest query whose column state is kept in shared memory

// Fast path: block b works on the 64 pairs order[b*64 ..], which the sort has made similar in
// shape (longest first), so the lanes of a warp run rows of similar width for a similar number
// of rows.  Dynamic shared memory: (max_len2 + 3) column words per thread.
template <int KIND>
__global__ void __launch_bounds__(BSW_FAST_TPB)
bsw_fast_kernel(bwagpu_bsw_params prm, bwagpu_seqpair *__restrict__ pairs,
                const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer,
                const uint32_t *__restrict__ order, int first, int n, int w, int cols,
                unsigned long long *__restrict__ cells, int direct)
{
    extern __shared__ uint32_t colmem[];
    const int slot = first + blockIdx.x * BSW_FAST_TPB + threadIdx.x;
    int my_cells = 0;
    if (slot < n) {
        const uint32_t i = order[slot];
        const bwagpu_seqpair sp = pairs[i];
        const BswView v = bsw_pair_view(sp, ref, qer, direct);
        const BswOut o = bsw_lane_fast<KIND, BSW_FAST_TPB>(prm, v.q, sp.len2, v.t, sp.len1, w,
                                                            sp.h0, colmem + threadIdx.x,
                                                            colmem + cols * BSW_FAST_TPB + threadIdx.x, v.dir);
        bwagpu_seqpair *d = &pairs[i];
        d->score = o.score; d->tle = o.tle; d->gtle = o.gtle; d->qle = o.qle;
        d->gscore = o.gscore; d->max_off = o.max_off;
        my_cells = o.cells;
    }
    bsw_add_cells(cells, my_cells);
}

// generic path over order[0 .. n_long): pairs too long for the shared-memory kernel
template <int KIND>
__global__ void __launch_bounds__(BSW_TPB)
bsw_thread_order_kernel(bwagpu_bsw_params prm, bwagpu_seqpair *__restrict__ pairs,
                        const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer,
                        const uint32_t *__restrict__ order, int n, int w, uint32_t *__restrict__ work,
                        unsigned long long *__restrict__ cells, int direct)
{
    const int64_t T = (int64_t)gridDim.x * blockDim.x;
    const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t *col = work + t0;
    unsigned long long mine = 0;
    for (int64_t s = t0; s < n; s += T) {
        const uint32_t i = order[s];
        const bwagpu_seqpair sp = pairs[i];
        const BswView v = bsw_pair_view(sp, ref, qer, direct);
        const BswOut o = bsw_lane_pair<KIND>(prm, v.q, sp.len2, v.t, sp.len1, w, sp.h0, col, T, v.dir);
        bwagpu_seqpair *d = &pairs[i];
        d->score = o.score; d->tle = o.tle; d->gtle = o.gtle; d->qle = o.qle;
        d->gscore = o.gscore; d->max_off = o.max_off;
        mine += o.cells;
    }
    if (mine) atomicAdd(cells, mine);
}

// sort key: query length, then target length (descending order puts the longest pairs first)
__global__ void bsw_key_kernel(const bwagpu_seqpair *__restrict__ pairs, int n, uint32_t *__restrict__ keys,
                               uint32_t *__restrict__ idx, int *__restrict__ n_long)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int l2 = min(pairs[i].len2, 0xFFFF), l1 = min(pairs[i].len1, 0xFFFF);
    keys[i] = ((uint32_t)l2 << 16) | (uint32_t)l1;
    idx[i] = (uint32_t)i;
    if (l2 > BSW_FAST_MAX_LEN2) atomicAdd(n_long, 1);
}

// Persistent grid: each thread walks pairs tid, tid+T, ... and owns one column of the
// interleaved scratch (word j of thread t lives at work[j*T + t]).
template <int KIND>
__global__ void __launch_bounds__(BSW_TPB)
bsw_thread_kernel(bwagpu_bsw_params prm, bwagpu_seqpair *__restrict__ pairs,
                  const uint8_t *__restrict__ ref, const uint8_t *__restrict__ qer, int n, int w,
                  uint32_t *__restrict__ work, unsigned long long *__restrict__ cells, int direct)
{
    const int64_t T = (int64_t)gridDim.x * blockDim.x;
    const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t *col = work + t0;
    unsigned long long mine = 0;
    for (int64_t i = t0; i < n; i += T) {
        bwagpu_seqpair sp = pairs[i];
        const BswView v = bsw_pair_view(sp, ref, qer, direct);
        BswOut o;
        if (KIND == BWAGPU_BSW_SCALAR)
            o = bsw_scalar_pair(prm, v.q, sp.len2, v.t, sp.len1, w, sp.h0, col, T, v.dir);
        else
            o = bsw_lane_pair<KIND>(prm, v.q, sp.len2, v.t, sp.len1, w, sp.h0, col, T, v.dir);
        bwagpu_seqpair *d = &pairs[i];
        d->score = o.score; d->tle = o.tle; d->gtle = o.gtle; d->qle = o.qle;
        d->gscore = o.gscore; d->max_off = o.max_off;
        mine += o.cells;
    }
    if (mine) atomicAdd(cells, mine);
}

// Shared memory per thread follows the query length, so the sorted order is cut into length tiers
// and each tier is launched with the column count it needs: short tiers get several times the
// resident warps of the longest one.  out[t] = first slot of the descending order whose len2 is
// <= BSW_TIERS[t].
// BSW_FINE_TIERS = 1: a tier every 16 query bases instead of four tiers: a launch's shared memory per thread follows its
// longest query, so finer tiers waste less of it and more blocks fit per SM in the middle of the length range.
#ifndef BSW_FINE_TIERS
#define BSW_FINE_TIERS 1
#endif
#if BSW_FINE_TIERS
#define BSW_NTIER 8
#define BSW_TIER_LIST {136, 120, 104, 88, 72, 56, 40, 24}
#else
#define BSW_NTIER 4
#define BSW_TIER_LIST {112, 80, 56, 40}
#endif
__constant__ int BSW_TIERS[BSW_NTIER] = BSW_TIER_LIST;
static const int BSW_TIERS_H[BSW_NTIER] = BSW_TIER_LIST;
__global__ void bsw_tier_kernel(const uint32_t *__restrict__ keys_desc, int n, int *__restrict__ out)
{
    const int t = threadIdx.x;
    if (t >= BSW_NTIER) return;
    const uint32_t thr = (uint32_t)BSW_TIERS[t];
    int lo = 0, hi = n;                       // first index with (key >> 16) <= thr
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if ((keys_desc[mid] >> 16) <= thr) hi = mid; else lo = mid + 1;
    }
    out[t] = lo;
}

// max len2 over the batch (device reduction; result read back by the launcher)
__global__ void bsw_maxlen_kernel(const bwagpu_seqpair *__restrict__ pairs, int n, int *out)
{
    int m1 = 0, m2 = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        m1 = max(m1, pairs[i].len1);
        m2 = max(m2, pairs[i].len2);
    }
    m1 = __reduce_max_sync(0xffffffffu, m1);
    m2 = __reduce_max_sync(0xffffffffu, m2);
    if ((threadIdx.x & 31) == 0) { atomicMax(&out[0], m1); atomicMax(&out[1], m2); }
}

// the context's DP-cell counter (device, zeroed when first allocated; bwagpu_bsw_cells reads / resets it)
static int bsw_cells_ptr(bwagpu_ctx *ctx, unsigned long long **out)
{
    const bool fresh = ctx->scratch[SCR_BSW_CELLS] == nullptr;
    int rc = bwagpu_scratch(ctx, SCR_BSW_CELLS, 64, (void **)out);
    if (rc) return rc;
    if (fresh) BWAGPU_CUDA(ctx, cudaMemsetAsync(*out, 0, 64, ctx->stream));
    return BWAGPU_OK;
}

extern "C" int bwagpu_bsw_cells(bwagpu_ctx *ctx, int reset, int64_t *cells)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned long long *d;
    int rc = bsw_cells_ptr(ctx, &d);
    if (rc) return rc;
    if (cells) {
        unsigned long long h = 0;
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(&h, d, 8, cudaMemcpyDeviceToHost, ctx->stream));
        BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
        *cells = (int64_t)h;
    }
    if (reset) BWAGPU_CUDA(ctx, cudaMemsetAsync(d, 0, 8, ctx->stream));
    return BWAGPU_OK;
}

static int bsw_check(bwagpu_ctx *ctx, const bwagpu_bsw_params *prm, int kind, int32_t n, int32_t w)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!prm) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "bsw: params is NULL");
    if (kind != BWAGPU_BSW_SCALAR && kind != BWAGPU_BSW_8BIT && kind != BWAGPU_BSW_16BIT)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "bsw: kind must be 0, 8 or 16 (got %d)", kind);
    if (n < 0 || w < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "bsw: negative numPairs or w");
    if (prm->e_del <= 0 || prm->e_ins <= 0)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "bsw: gap extension penalties must be positive");
    return BWAGPU_OK;
}

// generic kernel over all pairs in input order (scalar kind, or BWAGPU_BSW_PATH=thread)
static int bsw_launch_thread(bwagpu_ctx *ctx, const bwagpu_bsw_params *prm, int kind,
                             bwagpu_seqpair *d_pairs, const uint8_t *d_ref, const uint8_t *d_qer,
                             int32_t n, int32_t w, int max_len2)
{
    int blocks = ctx->sm_count * 8;
    int need = (n + BSW_TPB - 1) / BSW_TPB;
    if (blocks > need) blocks = need;
    const int64_t T = (int64_t)blocks * BSW_TPB;
    const int64_t words = (kind == BWAGPU_BSW_SCALAR ? 2 : 1) * (int64_t)(max_len2 + 4) * T;
    void *work;
    int rc = bwagpu_scratch(ctx, SCR_BSW_WORK, words * 4, &work);
    if (rc) return rc;
    unsigned long long *cells;
    if ((rc = bsw_cells_ptr(ctx, &cells))) return rc;
    BWAGPU_KERNEL_BEGIN(ctx);
    switch (kind) {
    case BWAGPU_BSW_16BIT:
        bsw_thread_kernel<16><<<blocks, BSW_TPB, 0, ctx->stream>>>(*prm, d_pairs, d_ref, d_qer, n,
                                                                  w, (uint32_t *)work, cells, ctx->bsw_direct);
        break;
    case BWAGPU_BSW_8BIT:
        bsw_thread_kernel<8><<<blocks, BSW_TPB, 0, ctx->stream>>>(*prm, d_pairs, d_ref, d_qer, n,
                                                                 w, (uint32_t *)work, cells, ctx->bsw_direct);
        break;
    default:
        bsw_thread_kernel<0><<<blocks, BSW_TPB, 0, ctx->stream>>>(*prm, d_pairs, d_ref, d_qer, n,
                                                                 w, (uint32_t *)work, cells, ctx->bsw_direct);
    }
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_BSW);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    return BWAGPU_OK;
}

template <int KIND>
static int bsw_launch_sorted(bwagpu_ctx *ctx, const bwagpu_bsw_params *prm, bwagpu_seqpair *d_pairs,
                             const uint8_t *d_ref, const uint8_t *d_qer, int32_t n, int32_t w)
{
    // 1. order the pairs by (len2, len1) descending; count those too long for shared memory
    uint32_t *keys, *order; int *d_m;
    int rc;
    if ((rc = bwagpu_scratch(ctx, SCR_BSW_KEYS, (int64_t)n * 8, (void **)&keys))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_BSW_ORDER, (int64_t)n * 8, (void **)&order))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_MISC0, 64, (void **)&d_m))) return rc;
    unsigned long long *cells;
    if ((rc = bsw_cells_ptr(ctx, &cells))) return rc;
    uint32_t *keys_out = keys + n, *idx_in = order + n;
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_m, 0, 16, ctx->stream));
    bsw_key_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(d_pairs, n, keys, idx_in, d_m + 2);
    bsw_maxlen_kernel<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(d_pairs, n, d_m);
    ctx->launches += 2;
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairsDescending((void *)nullptr, bytes, keys, keys_out, idx_in, order, n, 0, 32, ctx->stream);
    void *tmp;
    if ((rc = bwagpu_scratch(ctx, SCR_CUB, (int64_t)bytes + 256, &tmp))) return rc;
    BWAGPU_CUDA(ctx, cub::DeviceRadixSort::SortPairsDescending(tmp, bytes, keys, keys_out, idx_in, order, n, 0, 32, ctx->stream));
    ctx->launches += 5;      // CUB onesweep: histogram + 4 digit passes
    bsw_tier_kernel<<<1, 32, 0, ctx->stream>>>(keys_out, n, d_m + 4);
    ctx->launches += 1;
    int h_m[16];
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(h_m, d_m, 64, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    const int max_len2 = h_m[1], n_long = h_m[2];
    // 2. the few long pairs (front of the order) take the generic global-scratch kernel
    if (n_long > 0) {
        int blocks = ctx->sm_count * 8;
        const int need = (n_long + BSW_TPB - 1) / BSW_TPB;
        if (blocks > need) blocks = need;
        const int64_t T = (int64_t)blocks * BSW_TPB;
        void *work;
        if ((rc = bwagpu_scratch(ctx, SCR_BSW_WORK, (int64_t)(max_len2 + 4) * T * 4, &work))) return rc;
        BWAGPU_KERNEL_BEGIN(ctx);
        bsw_thread_order_kernel<KIND><<<blocks, BSW_TPB, 0, ctx->stream>>>(*prm, d_pairs, d_ref, d_qer, order, n_long, w,
                                                                         (uint32_t *)work, cells, ctx->bsw_direct);
        BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_BSW_GENERIC);
        BWAGPU_CUDA(ctx, cudaGetLastError());
    }
    // 3. everything else: shared-memory kernel, one launch per length tier
    if (n - n_long > 0) {
        static const bool one_tier = getenv("BWAGPU_BSW_TIERS") && !strcmp(getenv("BWAGPU_BSW_TIERS"), "0");
        const int cap2 = max_len2 < BSW_FAST_MAX_LEN2 ? max_len2 : BSW_FAST_MAX_LEN2;
        auto smem_of = [](int len2) { const int cols = len2 + 3, qwords = (len2 + 7) / 8 + 1; return (size_t)(cols + qwords) * BSW_FAST_TPB * 4; };
        static BwagpuSmemBound smem_bound;
        BWAGPU_CUDA(ctx, bwagpu_raise_dyn_smem(smem_bound, ctx->device, bsw_fast_kernel<KIND>, (int)smem_of(cap2)));
        int lo = n_long;
        for (int t = 0; t <= BSW_NTIER; ++t) {
            int hi = t < BSW_NTIER ? h_m[4 + t] : n;           // slots [lo, hi) hold len2 in (TIERS[t], TIERS[t-1]]
            if (one_tier) hi = n;
            if (hi < lo) hi = lo;
            if (hi > lo) {
                int len2 = t == 0 ? cap2 : BSW_TIERS_H[t - 1];
                if (len2 > cap2) len2 = cap2;
                BWAGPU_KERNEL_BEGIN(ctx);
                bsw_fast_kernel<KIND><<<(hi - lo + BSW_FAST_TPB - 1) / BSW_FAST_TPB, BSW_FAST_TPB, smem_of(len2), ctx->stream>>>(
                    *prm, d_pairs, d_ref, d_qer, order, lo, hi, w, len2 + 3, cells, ctx->bsw_direct);
                BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_BSW);
                BWAGPU_CUDA(ctx, cudaGetLastError());
            }
            lo = hi;
            if (one_tier) break;
        }
    }
    return BWAGPU_OK;
}

static int bsw_launch(bwagpu_ctx *ctx, const bwagpu_bsw_params *prm, int kind,
                      bwagpu_seqpair *d_pairs, const uint8_t *d_ref, const uint8_t *d_qer,
                      int32_t n, int32_t w, int max_len2)
{
    if (n == 0) return BWAGPU_OK;
    static const bool force_thread = getenv("BWAGPU_BSW_PATH") && !strcmp(getenv("BWAGPU_BSW_PATH"), "thread");
    if (kind == BWAGPU_BSW_SCALAR || force_thread)
        return bsw_launch_thread(ctx, prm, kind, d_pairs, d_ref, d_qer, n, w, max_len2);
    if (kind == BWAGPU_BSW_16BIT) return bsw_launch_sorted<16>(ctx, prm, d_pairs, d_ref, d_qer, n, w);
    return bsw_launch_sorted<8>(ctx, prm, d_pairs, d_ref, d_qer, n, w);
}

static int bsw_device_maxlen(bwagpu_ctx *ctx, const bwagpu_seqpair *d_pairs, int32_t n, int out[2])
{
    int *d_m;
    int rc = bwagpu_scratch(ctx, SCR_MISC0, 64, (void **)&d_m);
    if (rc) return rc;
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_m, 0, 8, ctx->stream));
    int blocks = ctx->sm_count * 4;
    bsw_maxlen_kernel<<<blocks, 256, 0, ctx->stream>>>(d_pairs, n, d_m);
    ctx->launches++;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(out, d_m, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}

// chain2aln's jobs: sequences read straight from the chunk's read buffer and the resident reference text (BswView)
int bwagpu_bsw_scores_direct(bwagpu_ctx *ctx, const bwagpu_bsw_params *prm, int kind, bwagpu_seqpair *d_pairs,
                             const uint8_t *d_ref_string, const uint8_t *d_reads, int32_t numPairs, int32_t w)
{
    int rc = bsw_check(ctx, prm, kind, numPairs, w);
    if (rc) return rc;
    if (numPairs == 0) return BWAGPU_OK;
    int ml[2];
    rc = bsw_device_maxlen(ctx, d_pairs, numPairs, ml);
    if (rc) return rc;
    ctx->bsw_direct = 1;
    rc = bsw_launch(ctx, prm, kind, d_pairs, d_ref_string, d_reads, numPairs, w, ml[1]);
    ctx->bsw_direct = 0;
    return rc;
}

extern "C" int bwagpu_bsw_get_scores_device(bwagpu_ctx *ctx, const bwagpu_bsw_params *prm,
                                            int kind, bwagpu_seqpair *d_pairArray,
                                            const uint8_t *d_seqBufRef, int64_t ref_bytes,
                                            const uint8_t *d_seqBufQer, int64_t qer_bytes,
                                            int32_t numPairs, int32_t w)
{
    int rc = bsw_check(ctx, prm, kind, numPairs, w);
    if (rc) return rc;
    (void)ref_bytes; (void)qer_bytes;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    if (numPairs == 0) return BWAGPU_OK;
    int ml[2];
    rc = bsw_device_maxlen(ctx, d_pairArray, numPairs, ml);
    if (rc) return rc;
    return bsw_launch(ctx, prm, kind, d_pairArray, d_seqBufRef, d_seqBufQer, numPairs, w, ml[1]);
}

extern "C" int bwagpu_bsw_get_scores(bwagpu_ctx *ctx, const bwagpu_bsw_params *prm, int kind,
                                     bwagpu_seqpair *pairArray, const uint8_t *seqBufRef,
                                     const uint8_t *seqBufQer, int32_t numPairs, int32_t w)
{
    int rc = bsw_check(ctx, prm, kind, numPairs, w);
    if (rc) return rc;
    if (numPairs == 0) return BWAGPU_OK;
    if (!pairArray || !seqBufRef || !seqBufQer)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "bsw: NULL buffer with numPairs=%d", numPairs);
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    // extents of the two byte arenas actually referenced by this batch
    int64_t ref_bytes = 0, qer_bytes = 0;
    int max_len2 = 0;
    for (int32_t i = 0; i < numPairs; ++i) {
        const bwagpu_seqpair &sp = pairArray[i];
        if (sp.len1 < 0 || sp.len2 < 0 || sp.idr < 0 || sp.idq < 0)
            BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "bsw: pair %d has a negative length/offset", i);
        int64_t r = (int64_t)sp.idr + sp.len1, q = (int64_t)sp.idq + sp.len2;
        if (r > ref_bytes) ref_bytes = r;
        if (q > qer_bytes) qer_bytes = q;
        if (sp.len2 > max_len2) max_len2 = sp.len2;
    }
    bwagpu_seqpair *d_pairs; uint8_t *d_ref, *d_qer;
    if ((rc = bwagpu_scratch(ctx, SCR_BSW_PAIRS, (int64_t)numPairs * sizeof(bwagpu_seqpair),
                             (void **)&d_pairs))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_BSW_REF, ref_bytes + 16, (void **)&d_ref))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_BSW_QER, qer_bytes + 16, (void **)&d_qer))) return rc;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_pairs, pairArray, (size_t)numPairs * sizeof(bwagpu_seqpair),
                                     cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_ref, seqBufRef, (size_t)ref_bytes, cudaMemcpyHostToDevice,
                                     ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_qer, seqBufQer, (size_t)qer_bytes, cudaMemcpyHostToDevice,
                                     ctx->stream));
    rc = bsw_launch(ctx, prm, kind, d_pairs, d_ref, d_qer, numPairs, w, max_len2);
    if (rc) return rc;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(pairArray, d_pairs, (size_t)numPairs * sizeof(bwagpu_seqpair),
                                     cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}
