// FM-index construction in HBM: bwagpu_index_build_fm / bwagpu_index_get_fm  (include/bwagpu.h).
//
// What the reference's `index` sub-command computes for the search side (FMI_search::build_index: suffix array of
// forward + reverse-complement text, BWT, CP_OCC check-points, SA sampled every 8th row; reference
// src/FMI_search.cpp:611-849, layout :64-68 of the header) with the same bytes, built from the text replica of
// bwagpu_index_load_ref.  It exists so that the human-scale configuration (2 x 3.1 G symbols) can be indexed on the
// GPU box in seconds — the reference's own SA-IS builder needs the better part of an hour there — and is the
// only way `bench.py` gets a 6.2 GB CP_OCC table.  It is NOT on the per-batch hot path.
//
// Method (no whole-text suffix array is ever held):
//   * the text is packed 2 bits per base, MSB first;
//   * suffixes are partitioned by their first P bases into 4^P buckets, processed one after the other;
//   * a bucket's suffixes are radix-sorted (CUB) on a 63-bit key = their first 21 symbols in a 3-bit code
//     (0 = past the end of the text, 1..4 = A..T), so the end of text sorts first without special cases;
//   * runs of equal keys (repeats) are compacted and refined round by round with CUB's segmented sort on the
//     next 21 symbols, until every run is resolved;
//   * each sorted bucket is streamed out as BWT symbols and SA samples at its global row offset;
//   * CP_OCC is two passes over the BWT (per-record one-hot words + counts, exclusive scan, counts).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_segmented_sort.cuh>
#include <stdlib.h>
#include <vector>
#include "ctx.cuh"

#define FMB_SYMS 21                 // symbols per 63-bit key

struct FmbText { const uint64_t *pk; int64_t n; };          // n symbols; pk zero-padded by two words

__device__ __forceinline__ uint64_t fmb_win32(const FmbText &t, int64_t i)       // bases i .. i+31, zero (= A) padded
{
    const int64_t w = i >> 5;
    const int s = (int)(i & 31) * 2;
    const uint64_t a = t.pk[w], b = t.pk[w + 1];
    return s ? (a << s) | (b >> (64 - s)) : a;
}

// FMB_SYMS symbols starting at i + h in the 3-bit code (0 once past the end)
__device__ __forceinline__ uint64_t fmb_key(const FmbText &t, int64_t i, int64_t h)
{
    const int64_t p = i + h;
    if (p >= t.n) return 0;
    const int64_t left = t.n - p;
    const int avail = left < FMB_SYMS ? (int)left : FMB_SYMS;
    const uint64_t x = fmb_win32(t, p);
    uint64_t key = 0;
#pragma unroll
    for (int k = 0; k < FMB_SYMS; ++k) {
        const uint64_t sym = (x >> (62 - 2 * k)) & 3;
        key = (key << 3) | (k < avail ? sym + 1 : 0);
    }
    return key;
}

__global__ void fmb_pack_kernel(const uint8_t *__restrict__ text, int64_t n, uint64_t *__restrict__ pk, int64_t words)
{
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= words) return;
    uint64_t v = 0;
    const int64_t b0 = w * 32;
    for (int k = 0; k < 32; ++k) {
        const int64_t i = b0 + k;
        const uint64_t c = i < n ? (uint64_t)(text[i] & 3) : 0;
        v |= c << (62 - 2 * k);
    }
    pk[w] = v;
}

// suffixes 0 .. n (n = the empty suffix) counted by their P-base prefix (missing bases read as A)
__global__ void fmb_hist_kernel(FmbText t, int P, unsigned long long *__restrict__ hist)
{
    extern __shared__ unsigned int sh[];
    const int nb = 1 << (2 * P);
    for (int k = threadIdx.x; k < nb; k += blockDim.x) sh[k] = 0;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= t.n; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned b = P ? (unsigned)(fmb_win32(t, i) >> (64 - 2 * P)) : 0u;
        atomicAdd(&sh[b], 1u);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < nb; k += blockDim.x) if (sh[k]) atomicAdd(&hist[k], (unsigned long long)sh[k]);
}

__global__ void fmb_gather_kernel(FmbText t, int P, unsigned bucket, uint64_t *__restrict__ keys, int64_t *__restrict__ vals,
                                  unsigned int *__restrict__ counter)
{
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x; i0 <= t.n; i0 += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = i0 + threadIdx.x;
        bool mine = false;
        if (i <= t.n) mine = (P ? (unsigned)(fmb_win32(t, i) >> (64 - 2 * P)) : 0u) == bucket;
        const unsigned m = __ballot_sync(0xffffffffu, mine);
        if (m) {
            const int lane = threadIdx.x & 31;
            unsigned base = 0;
            if (lane == __ffs(m) - 1) base = atomicAdd(counter, (unsigned)__popc(m));
            base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
            if (mine) {
                const unsigned slot = base + __popc(m & ((1u << lane) - 1));
                vals[slot] = i;
                keys[slot] = fmb_key(t, i, 0);
            }
        }
    }
}

__global__ void fmb_keygen_kernel(FmbText t, const int64_t *__restrict__ V, int m, int64_t h, uint64_t *__restrict__ K)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < m) K[j] = fmb_key(t, V[j], h);
}

// tied[j]: K[j] equals a neighbour inside its segment; head[j]: first element of such a run.  Entry m of both is 0
// so that the exclusive scans end with the totals.
__global__ void fmb_tie_kernel(const uint64_t *__restrict__ K, const uint8_t *__restrict__ segstart, int m,
                               int32_t *__restrict__ tied, int32_t *__restrict__ head)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j > m) return;
    if (j == m) { tied[j] = 0; head[j] = 0; return; }
    const uint64_t k = K[j];
    const bool eq_prev = j > 0 && K[j - 1] == k && !(segstart && segstart[j]);
    const bool eq_next = j + 1 < m && K[j + 1] == k && !(segstart && segstart[j + 1]);
    tied[j] = eq_prev || eq_next;
    head[j] = !eq_prev && eq_next;
}

__global__ void fmb_compact_kernel(const int64_t *__restrict__ V, const int32_t *__restrict__ S, const int32_t *__restrict__ tied,
                                   const int32_t *__restrict__ head, const int32_t *__restrict__ cpos,
                                   const int32_t *__restrict__ spos, int m, int64_t *__restrict__ V2, int32_t *__restrict__ S2,
                                   uint8_t *__restrict__ segstart2, int32_t *__restrict__ seg_off2)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j > m) return;
    if (j == m) { seg_off2[spos[m]] = cpos[m]; return; }
    if (!tied[j]) return;
    const int c = cpos[j];
    V2[c] = V[j];
    S2[c] = S ? S[j] : j;
    segstart2[c] = (uint8_t)head[j];
    if (head[j]) seg_off2[spos[j]] = c;
}

__global__ void fmb_scatter_kernel(const int64_t *__restrict__ V, const int32_t *__restrict__ S, int m, int64_t *__restrict__ vals)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < m) vals[S[j]] = V[j];
}

// a sorted bucket -> BWT symbols and SA samples at its rows
__global__ void fmb_emit_kernel(const int64_t *__restrict__ vals, int64_t cnt, int64_t row0, const uint8_t *__restrict__ text,
                                uint8_t *__restrict__ bwt, int8_t *__restrict__ ms, uint32_t *__restrict__ ls,
                                long long *__restrict__ sentinel)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= cnt) return;
    const int64_t pos = vals[j], row = row0 + j;
    if (pos == 0) { bwt[row] = 4; *sentinel = row; }
    else bwt[row] = text[pos - 1] & 3;
    if ((row & 7) == 0) { ms[row >> 3] = (int8_t)(pos >> 32); ls[row >> 3] = (uint32_t)pos; }
}

// CP_OCC pass 1: one thread per 64-symbol record: the four one-hot words (MSB = first symbol) and the per-record counts
__global__ void fmb_cp1_kernel(const uint8_t *__restrict__ bwt, int64_t n_rows, int64_t n_cp, int64_t *__restrict__ cp,
                               int64_t *__restrict__ cnt)
{
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_cp) return;
    uint64_t oh[4] = {0, 0, 0, 0};
    const int64_t b0 = c * 64;
    if (b0 + 64 <= n_rows) {
        const uint4 *p = (const uint4 *)(bwt + b0);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint4 v = p[q];
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                const unsigned s = (w[k >> 2] >> (8 * (k & 3))) & 0xFF;
                if (s < 4) oh[s] |= 1ull << (63 - (q * 16 + k));
            }
        }
    } else {
        for (int k = 0; k < 64 && b0 + k < n_rows; ++k) {
            const unsigned s = bwt[b0 + k];
            if (s < 4) oh[s] |= 1ull << (63 - k);
        }
    }
    int64_t *rec = cp + c * 8;
    for (int b = 0; b < 4; ++b) { rec[4 + b] = (int64_t)oh[b]; cnt[b * (n_cp + 1) + c] = __popcll(oh[b]); }
    if (c == 0) for (int b = 0; b < 4; ++b) cnt[b * (n_cp + 1) + n_cp] = 0;
}

__global__ void fmb_cp2_kernel(const int64_t *__restrict__ pre, int64_t n_cp, int64_t *__restrict__ cp)
{
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_cp) return;
    for (int b = 0; b < 4; ++b) cp[c * 8 + b] = pre[b * (n_cp + 1) + c];
}

namespace {
struct Tmp {                      // cudaMalloc'ed working set of one build, freed on every exit path
    std::vector<void *> p;
    ~Tmp() { for (void *q : p) cudaFree(q); }
    template <typename T> cudaError_t get(T **out, int64_t n)
    {
        void *q = nullptr;
        cudaError_t e = cudaMalloc(&q, (size_t)(n > 0 ? n : 1) * sizeof(T));
        if (e == cudaSuccess) p.push_back(q);
        *out = (T *)q;
        return e;
    }
};
}  // namespace

extern "C" int bwagpu_index_build_fm(bwagpu_ctx *ctx)
{
    BWAGPU_OWNS_INDEX(ctx);
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!ctx->d_ref_string) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_build_fm: no reference text loaded (bwagpu_index_load_ref)");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    cudaStream_t st = ctx->stream;
    const int64_t n = 2 * ctx->l_pac, n_rows = n + 1;
    if (n_rows >= (1LL << 38)) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "index_build_fm: text longer than 2^38 symbols is not supported");
    const uint8_t *text = (const uint8_t *)ctx->d_ref_string;
    // drop a previous replica
    if (ctx->d_cp_occ) { cudaFree(ctx->d_cp_occ); cudaFree(ctx->d_sa_ms); cudaFree(ctx->d_sa_ls); }
    ctx->d_cp_occ = ctx->d_sa_ms = ctx->d_sa_ls = nullptr; ctx->has_fm = false;
    if (ctx->d_fma_all) { cudaFree(ctx->d_fma_all); ctx->d_fma_all = nullptr; }
    if (ctx->d_fma_last) { cudaFree(ctx->d_fma_last); ctx->d_fma_last = nullptr; }
    ctx->fma_enabled = 0;
    const int64_t n_cp = (n_rows >> 6) + 1, n_sa = (n_rows >> 3) + 1;
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_cp_occ, (size_t)n_cp * 64));
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_sa_ms, (size_t)n_sa));
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_sa_ls, (size_t)n_sa * 4));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(ctx->d_sa_ms, 0, (size_t)n_sa, st));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(ctx->d_sa_ls, 0, (size_t)n_sa * 4, st));

    Tmp tmp;
    // packed text
    const int64_t words = (n >> 5) + 3;
    uint64_t *pk; uint8_t *bwt; unsigned long long *hist; long long *d_sent; unsigned int *d_counter;
    BWAGPU_CUDA(ctx, tmp.get(&pk, words));
    BWAGPU_CUDA(ctx, tmp.get(&bwt, n_cp * 64 + 64));
    fmb_pack_kernel<<<(unsigned)((words + 255) / 256), 256, 0, st>>>(text, n, pk, words);
    FmbText t = { pk, n };
    // bucket prefix length: buckets of at most ~2^27 suffixes (BWAGPU_FMB_P overrides, for tests)
    int P = 0;
    while (P < 6 && (n_rows >> (2 * P)) > (1LL << 27)) ++P;
    if (getenv("BWAGPU_FMB_P")) { P = atoi(getenv("BWAGPU_FMB_P")); if (P < 0) P = 0; if (P > 6) P = 6; }
    const int nb = 1 << (2 * P);
    BWAGPU_CUDA(ctx, tmp.get(&hist, nb));
    BWAGPU_CUDA(ctx, tmp.get(&d_sent, 1));
    BWAGPU_CUDA(ctx, tmp.get(&d_counter, 1));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(hist, 0, (size_t)nb * 8, st));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_sent, 0xFF, 8, st));
    fmb_hist_kernel<<<ctx->sm_count * 8, 256, (size_t)nb * 4, st>>>(t, P, hist);
    std::vector<unsigned long long> h_hist(nb);
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(h_hist.data(), hist, (size_t)nb * 8, cudaMemcpyDeviceToHost, st));
    BWAGPU_CUDA(ctx, cudaStreamSynchronize(st));
    int64_t maxb = 0, total = 0;
    for (int b = 0; b < nb; ++b) { if ((int64_t)h_hist[b] > maxb) maxb = (int64_t)h_hist[b]; total += (int64_t)h_hist[b]; }
    if (total != n_rows) BWAGPU_FAIL(ctx, BWAGPU_ERR_CUDA, "index_build_fm: bucket histogram covers %lld of %lld suffixes", (long long)total, (long long)n_rows);
    if (maxb >= (1LL << 31) - 8) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "index_build_fm: a suffix bucket holds %lld entries (text too repetitive for P = %d)", (long long)maxb, P);
    // per-bucket working set
    uint64_t *keys, *keys2, *Ka, *Kb; int64_t *vals, *vals2, *Va, *Vb; int32_t *Sa, *Sb, *tied, *head, *cpos, *spos, *offA, *offB;
    uint8_t *ssA, *ssB;
    BWAGPU_CUDA(ctx, tmp.get(&keys, maxb)); BWAGPU_CUDA(ctx, tmp.get(&keys2, maxb));
    BWAGPU_CUDA(ctx, tmp.get(&vals, maxb)); BWAGPU_CUDA(ctx, tmp.get(&vals2, maxb));
    BWAGPU_CUDA(ctx, tmp.get(&tied, maxb + 1)); BWAGPU_CUDA(ctx, tmp.get(&head, maxb + 1));
    BWAGPU_CUDA(ctx, tmp.get(&cpos, maxb + 1)); BWAGPU_CUDA(ctx, tmp.get(&spos, maxb + 1));
    size_t sort_bytes = 0, scan_bytes = 0;
    cub::DeviceRadixSort::SortPairs((void *)nullptr, sort_bytes, keys, keys2, vals, vals2, (int)maxb, 0, 63, st);
    cub::DeviceScan::ExclusiveSum((void *)nullptr, scan_bytes, tied, cpos, (int)(maxb + 1), st);
    void *sort_tmp, *scan_tmp;
    { uint8_t *q; BWAGPU_CUDA(ctx, tmp.get(&q, (int64_t)sort_bytes + 256)); sort_tmp = q; }
    { uint8_t *q; BWAGPU_CUDA(ctx, tmp.get(&q, (int64_t)scan_bytes + 256)); scan_tmp = q; }
    // refinement buffers are sized lazily from the first round's tie count
    int64_t ref_cap = 0; void *seg_tmp = nullptr; size_t seg_cap = 0;
    Ka = Kb = nullptr; Va = Vb = nullptr; Sa = Sb = nullptr; offA = offB = nullptr; ssA = ssB = nullptr;
    // Inside a bucket the first P symbols of every key are equal, so only the bits below them are sorted — except in the
    // (at most P + 1) buckets that also hold a suffix shorter than P bases: it was filed under its A-padded prefix but its
    // key has end codes there, so those buckets sort the whole key.
    const int end_bit = 63 - 3 * P > 0 ? 63 - 3 * P : 1;
    std::vector<uint8_t> full_sort(nb, 0);
    if (P > 0) {
        uint8_t tail[8] = {0};
        const int nt = (int)(n < P ? n : P);
        if (nt > 0) BWAGPU_CUDA(ctx, cudaMemcpy(tail, text + n - nt, (size_t)nt, cudaMemcpyDeviceToHost));
        for (int len = 0; len < P && len <= nt; ++len) {          // the suffix of `len` bases
            unsigned pre = 0;
            for (int k = 0; k < P; ++k) pre = (pre << 2) | (k < len ? (unsigned)(tail[nt - len + k] & 3) : 0u);
            full_sort[pre] = 1;
        }
    }

    int64_t row0 = 0, max_rounds = 0;
    for (int b = 0; b < nb; ++b) {
        const int64_t cnt = (int64_t)h_hist[b];
        if (cnt == 0) continue;
        BWAGPU_CUDA(ctx, cudaMemsetAsync(d_counter, 0, 4, st));
        fmb_gather_kernel<<<ctx->sm_count * 16, 256, 0, st>>>(t, P, (unsigned)b, keys, vals, d_counter);
        size_t sb = sort_bytes;
        BWAGPU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(sort_tmp, sb, keys, keys2, vals, vals2, (int)cnt, 0, full_sort[b] ? 63 : end_bit, st));
        // runs of equal keys in the sorted bucket
        const unsigned g1 = (unsigned)((cnt + 256) / 256);
        fmb_tie_kernel<<<g1, 256, 0, st>>>(keys2, nullptr, (int)cnt, tied, head);
        size_t cb = scan_bytes;
        BWAGPU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(scan_tmp, cb, tied, cpos, (int)(cnt + 1), st));
        cb = scan_bytes;
        BWAGPU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(scan_tmp, cb, head, spos, (int)(cnt + 1), st));
        int32_t h_m = 0, h_seg = 0;
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(&h_m, cpos + cnt, 4, cudaMemcpyDeviceToHost, st));
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(&h_seg, spos + cnt, 4, cudaMemcpyDeviceToHost, st));
        BWAGPU_CUDA(ctx, cudaStreamSynchronize(st));
        if (h_m > 0) {
            if (h_m > ref_cap) {                     // (re)allocate the refinement buffers
                ref_cap = (int64_t)h_m + h_m / 4 + 1024;
                BWAGPU_CUDA(ctx, tmp.get(&Ka, ref_cap)); BWAGPU_CUDA(ctx, tmp.get(&Kb, ref_cap));
                BWAGPU_CUDA(ctx, tmp.get(&Va, ref_cap)); BWAGPU_CUDA(ctx, tmp.get(&Vb, ref_cap));
                BWAGPU_CUDA(ctx, tmp.get(&Sa, ref_cap)); BWAGPU_CUDA(ctx, tmp.get(&Sb, ref_cap));
                BWAGPU_CUDA(ctx, tmp.get(&offA, ref_cap / 2 + 8)); BWAGPU_CUDA(ctx, tmp.get(&offB, ref_cap / 2 + 8));
                BWAGPU_CUDA(ctx, tmp.get(&ssA, ref_cap)); BWAGPU_CUDA(ctx, tmp.get(&ssB, ref_cap));
            }
            fmb_compact_kernel<<<g1, 256, 0, st>>>(vals2, nullptr, tied, head, cpos, spos, (int)cnt, Va, Sa, ssA, offA);
            int m = h_m, nseg = h_seg;
            int64_t h = FMB_SYMS, rounds = 0;
            while (m > 0) {
                if (h > n + FMB_SYMS) BWAGPU_FAIL(ctx, BWAGPU_ERR_CUDA, "index_build_fm: refinement did not terminate (bucket %d)", b);
                const unsigned gm = (unsigned)((m + 256) / 256);
                fmb_keygen_kernel<<<gm, 256, 0, st>>>(t, Va, m, h, Ka);
                size_t need = 0;
                cub::DeviceSegmentedSort::SortPairs((void *)nullptr, need, Ka, Kb, Va, Vb, m, nseg, offA, offA + 1, st);
                if (need > seg_cap) { uint8_t *q; seg_cap = need + need / 4 + 4096; BWAGPU_CUDA(ctx, tmp.get(&q, (int64_t)seg_cap)); seg_tmp = q; }
                BWAGPU_CUDA(ctx, cub::DeviceSegmentedSort::SortPairs(seg_tmp, need, Ka, Kb, Va, Vb, m, nseg, offA, offA + 1, st));
                fmb_scatter_kernel<<<gm, 256, 0, st>>>(Vb, Sa, m, vals2);
                fmb_tie_kernel<<<gm, 256, 0, st>>>(Kb, ssA, m, tied, head);
                cb = scan_bytes;
                BWAGPU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(scan_tmp, cb, tied, cpos, m + 1, st));
                cb = scan_bytes;
                BWAGPU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(scan_tmp, cb, head, spos, m + 1, st));
                int32_t m2 = 0, s2 = 0;
                BWAGPU_CUDA(ctx, cudaMemcpyAsync(&m2, cpos + m, 4, cudaMemcpyDeviceToHost, st));
                BWAGPU_CUDA(ctx, cudaMemcpyAsync(&s2, spos + m, 4, cudaMemcpyDeviceToHost, st));
                BWAGPU_CUDA(ctx, cudaStreamSynchronize(st));
                if (m2 > 0) fmb_compact_kernel<<<gm, 256, 0, st>>>(Vb, Sa, tied, head, cpos, spos, m, Va, Sb, ssB, offB);
                { int32_t *x = Sa; Sa = Sb; Sb = x; } { uint8_t *x = ssA; ssA = ssB; ssB = x; } { int32_t *x = offA; offA = offB; offB = x; }
                m = m2; nseg = s2; h += FMB_SYMS; ++rounds;
            }
            if (rounds > max_rounds) max_rounds = rounds;
        }
        fmb_emit_kernel<<<(unsigned)((cnt + 255) / 256), 256, 0, st>>>(vals2, cnt, row0, text, bwt, (int8_t *)ctx->d_sa_ms,
                                                                      (uint32_t *)ctx->d_sa_ls, d_sent);
        BWAGPU_CUDA(ctx, cudaGetLastError());
        row0 += cnt;
    }
    // CP_OCC
    int64_t *cnt4, *pre4;
    BWAGPU_CUDA(ctx, tmp.get(&cnt4, 4 * (n_cp + 1))); BWAGPU_CUDA(ctx, tmp.get(&pre4, 4 * (n_cp + 1)));
    fmb_cp1_kernel<<<(unsigned)((n_cp + 255) / 256), 256, 0, st>>>(bwt, n_rows, n_cp, (int64_t *)ctx->d_cp_occ, cnt4);
    size_t cpb = 0;
    cub::DeviceScan::ExclusiveSum((void *)nullptr, cpb, cnt4, pre4, (int)(n_cp + 1), st);
    uint8_t *cp_tmp; BWAGPU_CUDA(ctx, tmp.get(&cp_tmp, (int64_t)cpb + 256));
    for (int b4 = 0; b4 < 4; ++b4) {
        size_t x = cpb;
        BWAGPU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(cp_tmp, x, cnt4 + b4 * (n_cp + 1), pre4 + b4 * (n_cp + 1), (int)(n_cp + 1), st));
    }
    fmb_cp2_kernel<<<(unsigned)((n_cp + 255) / 256), 256, 0, st>>>(pre4, n_cp, (int64_t *)ctx->d_cp_occ);
    { const int rc = bwagpu_fm_convert_layout(ctx, ctx->d_cp_occ, ctx->d_cp_occ, n_cp, 1); if (rc) return rc; }   // st == ctx->stream
    int64_t tot[4]; long long h_sent = -1;
    for (int b4 = 0; b4 < 4; ++b4)
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(&tot[b4], pre4 + b4 * (n_cp + 1) + n_cp, 8, cudaMemcpyDeviceToHost, st));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&h_sent, d_sent, 8, cudaMemcpyDeviceToHost, st));
    BWAGPU_CUDA(ctx, cudaStreamSynchronize(st));
    BWAGPU_CUDA(ctx, cudaGetLastError());
    if (h_sent < 0 || tot[0] + tot[1] + tot[2] + tot[3] != n)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_CUDA, "index_build_fm: inconsistent result (sentinel %lld, %lld of %lld symbols)", h_sent,
                    (long long)(tot[0] + tot[1] + tot[2] + tot[3]), (long long)n);
    // count[] in the LOADED form (file value + 1, src/FMI_search.cpp:1304-1307)
    ctx->fm_count[0] = 1;
    for (int b4 = 0; b4 < 4; ++b4) ctx->fm_count[b4 + 1] = ctx->fm_count[b4] + tot[b4];
    ctx->fm_sentinel = h_sent;
    ctx->fm_ref_seq_len = n_rows;
    ctx->has_fm = true;
    ctx->launches += 8;
    if (getenv("BWAGPU_VERBOSE"))
        fprintf(stderr, "[bwagpu] index_build_fm: %lld rows, %d buckets (P = %d, largest %lld), at most %lld refinement rounds\n",
                (long long)n_rows, nb, P, (long long)maxb, (long long)max_rounds);
    return BWAGPU_OK;
}

extern "C" int bwagpu_index_get_fm(bwagpu_ctx *ctx, int64_t *reference_seq_len, int64_t *count, void *cp_occ,
                                   int8_t *sa_ms_byte, uint32_t *sa_ls_word, int64_t *sentinel_index)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!ctx->has_fm) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_get_fm: no FM-index resident");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    const int64_t n_rows = ctx->fm_ref_seq_len, n_cp = (n_rows >> 6) + 1, n_sa = (n_rows >> 3) + 1;
    if (reference_seq_len) *reference_seq_len = n_rows;
    if (count) for (int i = 0; i < 5; ++i) count[i] = ctx->fm_count[i];
    if (sentinel_index) *sentinel_index = ctx->fm_sentinel;
    void *d_ref_order = nullptr;
    if (cp_occ) {                                  // the replica is held in the paired word order: hand back the reference's
        BWAGPU_CUDA(ctx, cudaMalloc(&d_ref_order, (size_t)n_cp * 64));
        const int rc = bwagpu_fm_convert_layout(ctx, ctx->d_cp_occ, d_ref_order, n_cp, 0);
        if (rc) { cudaFree(d_ref_order); return rc; }
        const cudaError_t e = cudaMemcpyAsync(cp_occ, d_ref_order, (size_t)n_cp * 64, cudaMemcpyDeviceToHost, ctx->stream);
        if (e != cudaSuccess) { cudaFree(d_ref_order); BWAGPU_CUDA(ctx, e); }
    }
    if (sa_ms_byte) BWAGPU_CUDA(ctx, cudaMemcpyAsync(sa_ms_byte, ctx->d_sa_ms, (size_t)n_sa, cudaMemcpyDeviceToHost, ctx->stream));
    if (sa_ls_word) BWAGPU_CUDA(ctx, cudaMemcpyAsync(sa_ls_word, ctx->d_sa_ls, (size_t)n_sa * 4, cudaMemcpyDeviceToHost, ctx->stream));
    const cudaError_t e_sync = bwagpu_stream_sync(ctx);
    if (d_ref_order) cudaFree(d_ref_order);
    BWAGPU_CUDA(ctx, e_sync);
    return BWAGPU_OK;
}
