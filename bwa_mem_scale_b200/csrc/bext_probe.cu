// bwagpu_probe_bext (include/bwagpu.h): an A/B of the two ways to run FMI_search::backwardExt (reference
// src/FMI_search.cpp:2029-2056) on the GPU, isolated from the SMEM state machine, on the resident index:
//   mode 0  one walk per THREAD: the thread fetches both 64-byte CP_OCC records itself (2 x 256-bit loads each) and does
//           all eight popcounts — what seed_collect_kernel does;
//   mode 1  one walk per 8-LANE GROUP: lane b (0-3) owns base b of the record at sp, lane 4 + b base b of the record at
//           ep; each lane loads its 8-byte count and 8-byte one-hot word (the group's loads coalesce into the record's two
//           32-byte sectors), popcounts once, and __shfl_sync assembles (k, l, s) — the design BASELINE.json's north_star
//           describes.
// Every walk extends backwards along the reference text itself from a pseudo-random position (so the interval never
// empties: after ~16 bases it is the single suffix that occurs there, then every step is one dependent random touch of the
// table, like the deep part of a real read's walk), `steps` bases long.  Reports CP_OCC records fetched per second; the
// two modes produce the same final intervals (checksum), which the caller compares.
#include "ctx.cuh"
#include "fmi_core.cuh"

static FmView probe_view(const bwagpu_ctx *ctx)
{
    FmView f;
    f.count_tab = nullptr;
    f.cp_occ = (const int64_t *)ctx->d_cp_occ; f.sa_ms_byte = (const int8_t *)ctx->d_sa_ms; f.sa_ls_word = (const uint32_t *)ctx->d_sa_ls;
    for (int i = 0; i < 5; ++i) f.count[i] = ctx->fm_count[i];
    f.sentinel_index = ctx->fm_sentinel; f.ref_seq_len = ctx->fm_ref_seq_len; f.fma_all = nullptr; f.fma_last = nullptr;
    return f;
}

__device__ __forceinline__ int64_t probe_start(int64_t w, int64_t n_text, int steps)
{
    uint64_t h = (uint64_t)w * 0x9E3779B97F4A7C15ull + 0x1234567ull;
    h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
    return (int64_t)(h % (uint64_t)(n_text - steps - 2)) + steps + 1;
}

__global__ void __launch_bounds__(128) probe_thread_kernel(FmView f, const uint8_t *__restrict__ text, int64_t n_text, int64_t n_walks,
                                                           int steps, unsigned long long *__restrict__ out)
{
    unsigned long long recs = 0, sum = 0;
    for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < n_walks; w += (int64_t)gridDim.x * blockDim.x) {
        const int64_t p = probe_start(w, n_text, steps);
        int a = text[p] & 3;
        Biv v; v.k = fm_count(f, a); v.l = fm_count(f, 3 - a); v.s = fm_count(f, a + 1) - v.k;
        for (int t = 1; t <= steps; ++t) {
            a = text[p - t] & 3;
            recs += ((v.k >> 6) == ((v.k + v.s) >> 6)) ? 1 : 2;
            v = fm_bext(f, v, a);
        }
        sum += (unsigned long long)v.k + (unsigned long long)v.s * 3 + (unsigned long long)v.l * 7;
    }
    for (int d = 16; d; d >>= 1) { recs += __shfl_down_sync(0xffffffffu, recs, d); sum += __shfl_down_sync(0xffffffffu, sum, d); }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&out[0], recs); atomicAdd(&out[1], sum); }
}

__device__ __forceinline__ int64_t shfl64(int64_t v, int src, int width)
{
    const int lo = __shfl_sync(0xffffffffu, (int)(uint32_t)v, src, width);
    const int hi = __shfl_sync(0xffffffffu, (int)(uint32_t)((uint64_t)v >> 32), src, width);
    return (int64_t)(((uint64_t)(uint32_t)hi << 32) | (uint32_t)lo);
}

__global__ void __launch_bounds__(256) probe_group_kernel(FmView f, const uint8_t *__restrict__ text, int64_t n_text, int64_t n_walks,
                                                          int steps, unsigned long long *__restrict__ out)
{
    const int g = threadIdx.x & 7;                 // lane inside the group: base = g & 3, record = g >> 2 (0: sp, 1: ep)
    const int b = g & 3;
    const int64_t n_groups = ((int64_t)gridDim.x * blockDim.x) >> 3;
    unsigned long long recs = 0, sum = 0;
    for (int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3; w < n_walks; w += n_groups) {
        const int64_t p = probe_start(w, n_text, steps);
        int a = text[p] & 3;
        int64_t k = fm_count(f, a), l = fm_count(f, 3 - a), s = fm_count(f, a + 1) - k;
        for (int t = 1; t <= steps; ++t) {
            a = text[p - t] & 3;
            const int64_t pos = g < 4 ? k : k + s;                                  // sp for lanes 0-3, ep for lanes 4-7
            const int64_t *rec = f.cp_occ + (pos >> 6) * 8;
            const int64_t cnt = __ldg(rec + FM_W_CNT(b));
            const uint64_t bits = (uint64_t)__ldg(rec + FM_W_BITS(b));
            const int64_t occ = cnt + __popcll(bits & fm_prefix_mask((int)(pos & 63)));     // occ(pos, b)
            const int64_t other = shfl64(occ, g ^ 4, 8);
            const int64_t sz = g < 4 ? other - occ : occ - other;                   // s[b] on every lane
            // l[b] = l + sentinel + sum of s[c] for c > b: suffix sums over the four bases
            const int64_t s1 = shfl64(sz, (g & 4) | ((b + 1) & 3), 8), s2 = shfl64(sz, (g & 4) | ((b + 2) & 3), 8),
                          s3 = shfl64(sz, (g & 4) | ((b + 3) & 3), 8);
            const int64_t above = (b < 3 ? s1 : 0) + (b < 2 ? s2 : 0) + (b < 1 ? s3 : 0);
            const int64_t sent = (k <= f.sentinel_index && k + s > f.sentinel_index) ? 1 : 0;
            const int64_t nk = fm_count(f, b) + (g < 4 ? occ : other), nl = l + sent + above;
            if (g == 0) recs += ((k >> 6) == ((k + s) >> 6)) ? 1 : 2;
            k = shfl64(nk, a, 8); l = shfl64(nl, a, 8); s = shfl64(sz, a, 8);       // broadcast base a's result
        }
        if (g == 0) sum += (unsigned long long)k + (unsigned long long)s * 3 + (unsigned long long)l * 7;
    }
    for (int d = 16; d; d >>= 1) { recs += __shfl_down_sync(0xffffffffu, recs, d); sum += __shfl_down_sync(0xffffffffu, sum, d); }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&out[0], recs); atomicAdd(&out[1], sum); }
}

// mode 2: one walk per 4-LANE GROUP: lane b loads count and one-hot word of base b from BOTH records (four 8-byte loads;
// the group's lanes read one 32-byte sector per load), so s[b] needs no exchange
__global__ void __launch_bounds__(256) probe_group4_kernel(FmView f, const uint8_t *__restrict__ text, int64_t n_text, int64_t n_walks,
                                                           int steps, unsigned long long *__restrict__ out)
{
    const int b = threadIdx.x & 3;
    const int64_t n_groups = ((int64_t)gridDim.x * blockDim.x) >> 2;
    unsigned long long recs = 0, sum = 0;
    for (int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 2; w < n_walks; w += n_groups) {
        const int64_t p = probe_start(w, n_text, steps);
        int a = text[p] & 3;
        int64_t k = fm_count(f, a), l = fm_count(f, 3 - a), s = fm_count(f, a + 1) - k;
        for (int t = 1; t <= steps; ++t) {
            a = text[p - t] & 3;
            const int64_t ep = k + s;
            const int64_t *r0 = f.cp_occ + (k >> 6) * 8, *r1 = f.cp_occ + (ep >> 6) * 8;
            const int64_t c0 = __ldg(r0 + FM_W_CNT(b)), c1 = __ldg(r1 + FM_W_CNT(b));
            const uint64_t o0 = (uint64_t)__ldg(r0 + FM_W_BITS(b)), o1 = (uint64_t)__ldg(r1 + FM_W_BITS(b));
            const int64_t osp = c0 + __popcll(o0 & fm_prefix_mask((int)(k & 63)));
            const int64_t sz = c1 + __popcll(o1 & fm_prefix_mask((int)(ep & 63))) - osp;
            const int64_t s1 = shfl64(sz, (b + 1) & 3, 4), s2 = shfl64(sz, (b + 2) & 3, 4), s3 = shfl64(sz, (b + 3) & 3, 4);
            const int64_t above = (b < 3 ? s1 : 0) + (b < 2 ? s2 : 0) + (b < 1 ? s3 : 0);
            const int64_t sent = (k <= f.sentinel_index && ep > f.sentinel_index) ? 1 : 0;
            const int64_t nk = fm_count(f, b) + osp, nl = l + sent + above;
            if (b == 0) recs += ((k >> 6) == (ep >> 6)) ? 1 : 2;
            k = shfl64(nk, a, 4); l = shfl64(nl, a, 4); s = shfl64(sz, a, 4);
        }
        if (b == 0) sum += (unsigned long long)k + (unsigned long long)s * 3 + (unsigned long long)l * 7;
    }
    for (int d = 16; d; d >>= 1) { recs += __shfl_down_sync(0xffffffffu, recs, d); sum += __shfl_down_sync(0xffffffffu, sum, d); }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&out[0], recs); atomicAdd(&out[1], sum); }
}

// mode 3: one walk per PAIR of lanes: lane 0 owns the record at sp, lane 1 the record at ep (two 256-bit loads each)
__global__ void __launch_bounds__(256) probe_group2_kernel(FmView f, const uint8_t *__restrict__ text, int64_t n_text, int64_t n_walks,
                                                           int steps, unsigned long long *__restrict__ out)
{
    const int g = threadIdx.x & 1;
    const int64_t n_groups = ((int64_t)gridDim.x * blockDim.x) >> 1;
    unsigned long long recs = 0, sum = 0;
    for (int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 1; w < n_walks; w += n_groups) {
        const int64_t p = probe_start(w, n_text, steps);
        int a = text[p] & 3;
        int64_t k = fm_count(f, a), l = fm_count(f, 3 - a), s = fm_count(f, a + 1) - k;
        for (int t = 1; t <= steps; ++t) {
            a = text[p - t] & 3;
            const int64_t pos = g ? k + s : k;
            FmRec r;
            fm_load_rec(f, pos >> 6, r, false);
            const uint64_t m = fm_prefix_mask((int)(pos & 63));
            int64_t occ[4], oth[4];
#pragma unroll
            for (int b = 0; b < 4; ++b) { occ[b] = r.cnt[b] + __popcll(r.bits[b] & m); oth[b] = shfl64(occ[b], g ^ 1, 2); }
            int64_t sz[4];
#pragma unroll
            for (int b = 0; b < 4; ++b) sz[b] = g ? occ[b] - oth[b] : oth[b] - occ[b];
            const int64_t sent = (k <= f.sentinel_index && k + s > f.sentinel_index) ? 1 : 0;
            const int64_t l3 = l + sent, l2 = l3 + sz[3], l1 = l2 + sz[2], l0 = l1 + sz[1];
            if (g == 0) recs += ((k >> 6) == ((k + s) >> 6)) ? 1 : 2;
            const int64_t osp = a == 0 ? (g ? oth[0] : occ[0]) : a == 1 ? (g ? oth[1] : occ[1]) : a == 2 ? (g ? oth[2] : occ[2]) : (g ? oth[3] : occ[3]);
            k = fm_count(f, a) + osp;
            s = a == 0 ? sz[0] : a == 1 ? sz[1] : a == 2 ? sz[2] : sz[3];
            l = a == 0 ? l0 : a == 1 ? l1 : a == 2 ? l2 : l3;
        }
        if (g == 0) sum += (unsigned long long)k + (unsigned long long)s * 3 + (unsigned long long)l * 7;
    }
    for (int d = 16; d; d >>= 1) { recs += __shfl_down_sync(0xffffffffu, recs, d); sum += __shfl_down_sync(0xffffffffu, sum, d); }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&out[0], recs); atomicAdd(&out[1], sum); }
}

extern "C" int bwagpu_probe_bext(bwagpu_ctx *ctx, int mode, int64_t n_walks, int32_t steps, double *records_per_s,
                                 uint64_t *checksum, float *ms_out)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!ctx->has_fm || !ctx->d_ref_string) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "probe_bext: needs the FM-index and the reference text");
    if (mode < 0 || mode > 3 || n_walks <= 0 || steps <= 0 || 2 * ctx->l_pac < steps + 8) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "probe_bext: bad argument");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned long long *d_out;
    int rc = bwagpu_scratch(ctx, SCR_MISC1, 64, (void **)&d_out);
    if (rc) return rc;
    const FmView f = probe_view(ctx);
    float best = 1e30f; unsigned long long h[2] = {0, 0};
    for (int rep = 0; rep < 3; ++rep) {
        BWAGPU_CUDA(ctx, cudaMemsetAsync(d_out, 0, 16, ctx->stream));
        BWAGPU_CUDA(ctx, cudaEventRecord(ctx->kev0, ctx->stream));
        if (mode == 0) {
            int per_sm = 0;
            BWAGPU_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, probe_thread_kernel, 128, 0));
            probe_thread_kernel<<<ctx->sm_count * per_sm, 128, 0, ctx->stream>>>(f, (const uint8_t *)ctx->d_ref_string, 2 * ctx->l_pac, n_walks, steps, d_out);
        } else if (mode == 1) {
            int per_sm = 0;
            BWAGPU_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, probe_group_kernel, 256, 0));
            probe_group_kernel<<<ctx->sm_count * per_sm, 256, 0, ctx->stream>>>(f, (const uint8_t *)ctx->d_ref_string, 2 * ctx->l_pac, n_walks, steps, d_out);
        } else if (mode == 2) {
            int per_sm = 0;
            BWAGPU_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, probe_group4_kernel, 256, 0));
            probe_group4_kernel<<<ctx->sm_count * per_sm, 256, 0, ctx->stream>>>(f, (const uint8_t *)ctx->d_ref_string, 2 * ctx->l_pac, n_walks, steps, d_out);
        } else {
            int per_sm = 0;
            BWAGPU_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, probe_group2_kernel, 256, 0));
            probe_group2_kernel<<<ctx->sm_count * per_sm, 256, 0, ctx->stream>>>(f, (const uint8_t *)ctx->d_ref_string, 2 * ctx->l_pac, n_walks, steps, d_out);
        }
        BWAGPU_CUDA(ctx, cudaEventRecord(ctx->kev1, ctx->stream));
        BWAGPU_CUDA(ctx, cudaEventSynchronize(ctx->kev1));
        float ms = 0;
        BWAGPU_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->kev0, ctx->kev1));
        BWAGPU_CUDA(ctx, cudaMemcpy(h, d_out, 16, cudaMemcpyDeviceToHost));
        if (ms < best) best = ms;
    }
    BWAGPU_CUDA(ctx, cudaGetLastError());
    ctx->launches += 3;
    if (records_per_s) *records_per_s = (double)h[0] / (best * 1e-3);
    if (checksum) *checksum = h[1];
    if (ms_out) *ms_out = best;
    return BWAGPU_OK;
}
