// bwagpu_emf_build_device / bwagpu_emf_get  (include/bwagpu.h): the exact-match-filter table built in HBM.
//
// The reference's `perfect-index` (src/perfect_index.cpp:614-757, 926-990) inserts the l_pac seeds one after the other,
// so the slot a collision entry lands in — and with it the byte layout of the table — depends on the insertion order;
// bwagpu_emf_build reproduces that on the host, byte for byte, and like the reference needs minutes to hours at human scale.
// This builder produces a table with the SAME CONTENT in the same format, laid out by a parallel rule:
//   * every position that starts a seed is filed under its hash slot (the reference's hash of the canonical
//     orientation, __get_hash_idx_fw/rc, perfect.h:466-706) and the (slot, position) pairs are radix-sorted;
//   * one thread per occupied slot groups its positions into classes of identical canonical seeds; the smallest position
//     of a class is the entry's location (what sequential insertion in ascending order makes the primary), later ones go
//     to its multi-location list, same-text copies first, reverse-complement copies second, ascending (perfect.h:113-129);
//   * the classes of a slot form a balanced search tree rooted IN the slot (rebuild_perfect_table_for_mapping's shape);
//     the other nodes take the next free slots from the list of empty slots, in slot order (the reference takes the next
//     free slot after the chain's tail);
//   * multi-location lists are laid out in slot order by an exclusive scan; lists with 256 or more copies on one strand
//     use the table's second encoding in a region behind the short ones.
// find_perfect_match_entry (the reference's or emf_probe_kernel) returns for every read the same code and the same
// location on this table as on the reference-built one; only the list ids inside seq.perfect.flags (bits 2 and up, an
// index into THIS table's loc_table) differ, and they dereference to the same lists (tests/test_emf_gpu.py).
// Holes (runs of N in the reference) are not supported here: the host builder handles them.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>
#include <algorithm>
#include <stdlib.h>
#include <vector>
#include "ctx.cuh"

#define EB_NO 0xFFFFFFFFu
#define EB_MAXC 128                     // distinct seeds per hash slot handled (the XOR-of-words hash has weak spots on repeat families)
#define EB_MAXW 8                       // seed words: seeds up to 256 bases

struct EbText { const uint64_t *pk; int64_t l_pac; int L, nw, r; };   // packed forward + reverse-complement text

__device__ __forceinline__ uint64_t eb_win32(const uint64_t *pk, int64_t i)
{
    const int64_t w = i >> 5;
    const int s = (int)(i & 31) * 2;
    const uint64_t a = pk[w], b = pk[w + 1];
    return s ? (a << s) | (b >> (64 - s)) : a;
}

// word j of the seed at position p read in orientation fw (1: as it stands; 0: reverse complement): 32 bases, first base
// in the top bits; the last word holds the remaining r bases in its LOW bits (how the reference's hash packs them)
__device__ __forceinline__ uint64_t eb_word(const EbText &t, uint32_t p, int fw, int j)
{
    const int64_t start = fw ? (int64_t)p : 2 * t.l_pac - (int64_t)p - t.L;
    const uint64_t w = eb_win32(t.pk, start + 32 * j);
    return j == t.nw - 1 && t.r < 32 ? w >> (64 - 2 * t.r) : w;
}

// __compare_fw_rc (perfect.h:345-351): is the forward reading <= the reverse complement over the first half?
__device__ __forceinline__ int eb_fw_less(const EbText &t, uint32_t p)
{
    int left = (t.L + 1) / 2;
    for (int j = 0; left > 0; ++j, left -= 32) {
        uint64_t a = eb_win32(t.pk, (int64_t)p + 32 * j), b = eb_win32(t.pk, 2 * t.l_pac - (int64_t)p - t.L + 32 * j);
        if (left < 32) { a >>= 64 - 2 * left; b >>= 64 - 2 * left; }
        if (a != b) return a < b;
    }
    return 1;
}

__device__ __forceinline__ uint32_t eb_slot(const EbText &t, uint32_t p, int fw, uint32_t n_entry)
{
    uint64_t h = 0;
    for (int j = 0; j < t.nw; ++j) h ^= eb_word(t, p, fw, j);
    h ^= h >> 33; h *= 0xff51afd7ed558ccdULL; h ^= h >> 33; h *= 0xc4ceb9fe1a85ec53ULL; h ^= h >> 33;
    return (uint32_t)(h % n_entry);
}

// canonical seeds of a and b compared as strings
__device__ __forceinline__ int eb_cmp(const EbText &t, uint32_t a, int afw, uint32_t b, int bfw)
{
    for (int j = 0; j < t.nw; ++j) {
        const uint64_t x = eb_word(t, a, afw, j), y = eb_word(t, b, bfw, j);
        if (x != y) return x > y ? 1 : -1;
    }
    return 0;
}

__global__ void eb_pack_kernel(const uint8_t *__restrict__ text, int64_t n, uint64_t *__restrict__ pk, int64_t words)
{
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= words) return;
    uint64_t v = 0;
    for (int k = 0; k < 32; ++k) {
        const int64_t i = w * 32 + k;
        v |= (i < n ? (uint64_t)(text[i] & 3) : 0ull) << (62 - 2 * k);
    }
    pk[w] = v;
}

// (slot, position) of every position that starts a seed: inside one contig (calc_loc_key, perfect_index.cpp:1045-1164)
__global__ void eb_key_kernel(EbText t, uint32_t n_entry, int n_seqs, const int64_t *__restrict__ coff, const int32_t *__restrict__ clen,
                              uint32_t *__restrict__ keys, uint32_t *__restrict__ vals, unsigned long long *__restrict__ n_valid)
{
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= t.l_pac) return;
    int lo = 0, hi = n_seqs - 1;                       // contig holding p
    while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (coff[mid] <= p) lo = mid; else hi = mid - 1; }
    const bool ok = p + t.L <= coff[lo] + clen[lo];
    uint32_t k = EB_NO;
    if (ok) k = eb_slot(t, (uint32_t)p, eb_fw_less(t, (uint32_t)p), n_entry);
    keys[p] = k; vals[p] = (uint32_t)p;
    const unsigned m = __ballot_sync(__activemask(), ok);
    if (ok && (threadIdx.x & 31) == __ffs(m) - 1) atomicAdd(n_valid, (unsigned long long)__popc(m));
}

struct EbClass { uint32_t rep; int32_t fw; uint32_t nfw, nrc; };

// classes of identical canonical seeds among the positions vals[i0 .. i1) of one slot (ascending positions).
// Returns the number of classes, or -1 when there are more than EB_MAXC.
__device__ int eb_classify(const EbText &t, const uint32_t *__restrict__ vals, int64_t i0, int64_t i1, EbClass *cls)
{
    int nc = 0;
    for (int64_t i = i0; i < i1; ++i) {
        const uint32_t q = vals[i];
        const int qfw = eb_fw_less(t, q);
        int c = 0;
        for (; c < nc; ++c) if (eb_cmp(t, q, qfw, cls[c].rep, cls[c].fw) == 0) break;
        if (c == nc) {
            if (nc == EB_MAXC) return -1;
            cls[nc].rep = q; cls[nc].fw = qfw; cls[nc].nfw = cls[nc].nrc = 0;
            ++nc;
        } else if (qfw == cls[c].fw) ++cls[c].nfw;       // same text as the primary (__seedmatch == 1)
        else ++cls[c].nrc;                               // its reverse complement   (__seedmatch == 2)
    }
    return nc;
}

__device__ __forceinline__ bool eb_small(const EbClass &c) { return c.nfw < 256 && c.nrc < 256; }

// pass A: per occupied slot (thread at the first element of its run), extra entries and short-list words it needs
__global__ void eb_count_kernel(EbText t, const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals, int64_t n_valid,
                                uint32_t *__restrict__ extra, uint32_t *__restrict__ mainw, uint8_t *__restrict__ occupied,
                                unsigned long long *__restrict__ stats /* runs, many words, overflow */)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t k = i < n_valid ? keys[i] : EB_NO;
    const bool head = i < n_valid && (i == 0 || keys[i - 1] != k);
    const unsigned hm = __ballot_sync(0xffffffffu, head);            // occupied slots, counted per warp
    if (hm && (threadIdx.x & 31) == 0) atomicAdd(&stats[0], (unsigned long long)__popc(hm));
    if (i > n_valid) return;
    if (!head) { extra[i] = 0; mainw[i] = 0; return; }
    int64_t i1 = i + 1;
    while (i1 < n_valid && keys[i1] == k) ++i1;
    EbClass cls[EB_MAXC];
    const int nc = eb_classify(t, vals, i, i1, cls);
    occupied[k] = 0;                                  // the array starts as all ones: what is left set is a free slot
    if (nc < 0) { atomicAdd(&stats[2], 1ull); extra[i] = 0; mainw[i] = 0; return; }
    uint32_t mw = 0; unsigned long long many = 0;
    for (int c = 0; c < nc; ++c) {
        const uint32_t n = cls[c].nfw + cls[c].nrc;
        if (!n) continue;
        if (eb_small(cls[c])) mw += 1 + n; else { mw += 1; many += 2 + n; }
    }
    extra[i] = (uint32_t)(nc - 1); mainw[i] = mw;
    if (many) atomicAdd(&stats[1], many);
}

__global__ void eb_init_table_kernel(uint4 *__restrict__ T, int64_t n)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) T[i] = make_uint4(0u, EB_NO, EB_NO, EB_NO);
}

// pass B: entries, search trees and multi-location lists of one occupied slot
__global__ void eb_fill_kernel(EbText t, const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals, int64_t n_valid,
                               const uint32_t *__restrict__ extra_off, const uint32_t *__restrict__ main_off,
                               const uint32_t *__restrict__ empties, uint4 *__restrict__ T, uint32_t *__restrict__ loc,
                               uint32_t many_base, unsigned long long *__restrict__ many_cursor)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_valid) return;
    const uint32_t k = keys[i];
    if (i > 0 && keys[i - 1] == k) return;
    int64_t i1 = i + 1;
    while (i1 < n_valid && keys[i1] == k) ++i1;
    EbClass cls[EB_MAXC];
    const int nc = eb_classify(t, vals, i, i1, cls);
    if (nc < 0) return;
    // multi-location lists: header(s) now, members in a second walk over the run
    uint32_t ml[EB_MAXC], fwc[EB_MAXC], rcc[EB_MAXC];
    uint32_t cur = 1 + main_off[i];                                  // entry 0 of loc_table is never used
    for (int c = 0; c < nc; ++c) {
        const uint32_t n = cls[c].nfw + cls[c].nrc;
        ml[c] = 0; fwc[c] = rcc[c] = 0;
        if (!n) continue;
        ml[c] = cur;
        if (eb_small(cls[c])) {
            loc[cur] = (cls[c].nfw << 16) + cls[c].nrc;
            fwc[c] = cur + 1; rcc[c] = cur + 1 + cls[c].nfw;
            cur += 1 + n;
        } else {
            const uint32_t at = many_base + (uint32_t)atomicAdd(many_cursor, (unsigned long long)(2 + n));
            loc[cur] = 0x80000000u | at;
            loc[at] = cls[c].nfw; loc[at + 1] = cls[c].nrc;
            fwc[c] = at + 2; rcc[c] = at + 2 + cls[c].nfw;
            cur += 1;
        }
    }
    for (int64_t j = i; j < i1; ++j) {
        const uint32_t q = vals[j];
        const int qfw = eb_fw_less(t, q);
        for (int c = 0; c < nc; ++c) {
            if (q == cls[c].rep) break;
            if (!ml[c] || eb_cmp(t, q, qfw, cls[c].rep, cls[c].fw) != 0) continue;
            if (qfw == cls[c].fw) loc[fwc[c]++] = q; else loc[rcc[c]++] = q;
            break;
        }
    }
    // classes in string order (insertion sort; nc is tiny)
    int ord[EB_MAXC];
    for (int c = 0; c < nc; ++c) {
        int j = c;
        while (j > 0 && eb_cmp(t, cls[c].rep, cls[c].fw, cls[ord[j - 1]].rep, cls[ord[j - 1]].fw) < 0) { ord[j] = ord[j - 1]; --j; }
        ord[j] = c;
    }
    // balanced tree rooted in the hash slot; the other nodes take this slot's share of the empty slots
    const uint32_t *mine = empties + extra_off[i];
    int next = 0;
    uint32_t st_slot[12]; int st_lo[12], st_hi[12], sp = 0;          // depth <= log2(EB_MAXC) + 2
    st_slot[0] = k; st_lo[0] = 0; st_hi[0] = nc - 1; sp = 1;
    while (sp > 0) {
        --sp;
        const uint32_t slot = st_slot[sp]; const int lo = st_lo[sp], hi = st_hi[sp];
        const int mid = (lo + hi) / 2;
        const EbClass &e = cls[ord[mid]];
        const uint32_t left = mid > lo ? mine[next++] : EB_NO, right = mid < hi ? mine[next++] : EB_NO;
        T[slot] = make_uint4((uint32_t)e.fw | (slot != k ? 2u : 0u) | (ml[ord[mid]] << 2), e.rep, left, right);
        if (right != EB_NO) { st_slot[sp] = right; st_lo[sp] = mid + 1; st_hi[sp] = hi; ++sp; }
        if (left != EB_NO) { st_slot[sp] = left; st_lo[sp] = lo; st_hi[sp] = mid - 1; ++sp; }
    }
}

namespace {
struct EbTmp {                    // cudaMalloc'ed working set, freed on every exit path (or early via drop())
    std::vector<void *> p;
    ~EbTmp() { for (void *q : p) if (q) cudaFree(q); }
    template <typename T> cudaError_t get(T **out, int64_t n)
    {
        void *q = nullptr;
        cudaError_t e = cudaMalloc(&q, (size_t)(n > 0 ? n : 1) * sizeof(T));
        if (e == cudaSuccess) p.push_back(q);
        *out = (T *)q;
        return e;
    }
    void drop(void *q) { for (auto &x : p) if (x == q) { cudaFree(q); x = nullptr; } }
};
}  // namespace

extern "C" int bwagpu_emf_build_device(bwagpu_ctx *ctx, int32_t seed_len, double slack)
{
    BWAGPU_OWNS_INDEX(ctx);
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!ctx->d_ref_string) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "emf_build_device: no reference text loaded (bwagpu_index_load_ref)");
    if (seed_len < 2 || seed_len > 32 * EB_MAXW) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "emf_build_device: seed_len %d outside [2, %d]", seed_len, 32 * EB_MAXW);
    const int64_t l_pac = ctx->l_pac;
    if (l_pac >= (int64_t)0xFFFFFFFFu) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "emf_build_device: reference longer than 2^32 - 1 bases");
    const uint64_t n_entry64 = (uint64_t)((double)l_pac * slack);          // perfect_index.cpp:1356
    if (n_entry64 == 0 || n_entry64 > 0xFFFFFFFFull || n_entry64 < (uint64_t)l_pac)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "emf_build_device: slack %.3f gives an unusable table size", slack);
    const uint32_t n_entry = (uint32_t)n_entry64;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = bwagpu_emf_unload(ctx);
    if (rc) return rc;
    cudaStream_t st = ctx->stream;
    EbTmp tmp;
    // packed text, keys
    const int64_t n_text = 2 * l_pac, words = (n_text >> 5) + 3;
    uint64_t *pk; uint32_t *keys, *vals, *keys2, *vals2; unsigned long long *d_stats;
    BWAGPU_CUDA(ctx, tmp.get(&pk, words));
    BWAGPU_CUDA(ctx, tmp.get(&keys, l_pac)); BWAGPU_CUDA(ctx, tmp.get(&vals, l_pac));
    BWAGPU_CUDA(ctx, tmp.get(&keys2, l_pac)); BWAGPU_CUDA(ctx, tmp.get(&vals2, l_pac));
    BWAGPU_CUDA(ctx, tmp.get(&d_stats, 8));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_stats, 0, 64, st));
    eb_pack_kernel<<<(unsigned)((words + 255) / 256), 256, 0, st>>>((const uint8_t *)ctx->d_ref_string, n_text, pk, words);
    EbText t; t.pk = pk; t.l_pac = l_pac; t.L = seed_len; t.nw = (seed_len + 31) / 32; t.r = seed_len - 32 * (t.nw - 1);
    eb_key_kernel<<<(unsigned)((l_pac + 255) / 256), 256, 0, st>>>(t, n_entry, ctx->n_contigs, (const int64_t *)ctx->d_contig_off,
                                                                   (const int32_t *)ctx->d_contig_len, keys, vals, d_stats + 3);
    size_t sort_bytes = 0;
    cub::DeviceRadixSort::SortPairs((void *)nullptr, sort_bytes, keys, keys2, vals, vals2, l_pac, 0, 32, st);
    { uint8_t *q; BWAGPU_CUDA(ctx, tmp.get(&q, (int64_t)sort_bytes + 256));
      BWAGPU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(q, sort_bytes, keys, keys2, vals, vals2, l_pac, 0, 32, st));
      BWAGPU_CUDA(ctx, cudaStreamSynchronize(st)); tmp.drop(q); }
    unsigned long long h_stats[8];
    BWAGPU_CUDA(ctx, cudaMemcpy(h_stats, d_stats, 64, cudaMemcpyDeviceToHost));
    const int64_t n_valid = (int64_t)h_stats[3];
    tmp.drop(keys); tmp.drop(vals);                      // the sorted copy lives in keys2 / vals2
    // pass A
    uint32_t *extra, *mainw; uint8_t *occupied;
    BWAGPU_CUDA(ctx, tmp.get(&extra, n_valid + 1)); BWAGPU_CUDA(ctx, tmp.get(&mainw, n_valid + 1));
    BWAGPU_CUDA(ctx, tmp.get(&occupied, (int64_t)n_entry));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(occupied, 1, n_entry, st));
    eb_count_kernel<<<(unsigned)((n_valid + 256) / 256), 256, 0, st>>>(t, keys2, vals2, n_valid, extra, mainw, occupied, d_stats);
    size_t scan_bytes = 0;
    cub::DeviceScan::ExclusiveSum((void *)nullptr, scan_bytes, extra, extra, n_valid + 1, st);
    { uint8_t *q; BWAGPU_CUDA(ctx, tmp.get(&q, (int64_t)scan_bytes + 256));
      size_t b = scan_bytes;
      BWAGPU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(q, b, extra, extra, n_valid + 1, st));
      b = scan_bytes;
      BWAGPU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(q, b, mainw, mainw, n_valid + 1, st));
      BWAGPU_CUDA(ctx, cudaStreamSynchronize(st)); tmp.drop(q); }
    uint32_t tot_extra = 0, tot_main = 0;
    BWAGPU_CUDA(ctx, cudaMemcpy(&tot_extra, extra + n_valid, 4, cudaMemcpyDeviceToHost));
    BWAGPU_CUDA(ctx, cudaMemcpy(&tot_main, mainw + n_valid, 4, cudaMemcpyDeviceToHost));
    BWAGPU_CUDA(ctx, cudaMemcpy(h_stats, d_stats, 64, cudaMemcpyDeviceToHost));
    if (h_stats[2]) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "emf_build_device: %llu hash slots hold more than %d distinct seeds (use bwagpu_emf_build)", h_stats[2], EB_MAXC);
    const uint64_t n_loc0 = 1ull + tot_main + h_stats[1];
    if (n_loc0 >= (1ull << 30)) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "emf_build_device: multi-location lists need %llu words (limit 2^30)", (unsigned long long)n_loc0);
    const uint32_t n_loc = (uint32_t)(n_loc0 + (16 - (n_loc0 % 16)));      // build_loc_to_loc_table's padding
    // empty slots, in slot order
    uint32_t *empties; long long *d_nsel;
    BWAGPU_CUDA(ctx, tmp.get(&empties, (int64_t)tot_extra + 1));
    BWAGPU_CUDA(ctx, tmp.get(&d_nsel, 1));
    {   // select slots with occupied == 0 until tot_extra of them are found: chunks keep the output bounded
        uint8_t *free_flag = occupied;                   // 1 = free
        const int64_t chunk = 1 << 28;
        int64_t got = 0;
        size_t sel_bytes = 0;
        cub::CountingInputIterator<uint32_t> it0(0);
        uint32_t *sel_out; BWAGPU_CUDA(ctx, tmp.get(&sel_out, chunk));
        cub::DeviceSelect::Flagged((void *)nullptr, sel_bytes, it0, free_flag, sel_out, d_nsel, chunk, st);
        uint8_t *q; BWAGPU_CUDA(ctx, tmp.get(&q, (int64_t)sel_bytes + 256));
        for (int64_t c0 = 0; c0 < (int64_t)n_entry && got < (int64_t)tot_extra; c0 += chunk) {
            const int64_t nc = std::min<int64_t>(chunk, (int64_t)n_entry - c0);
            cub::CountingInputIterator<uint32_t> it((uint32_t)c0);
            size_t b = sel_bytes;
            BWAGPU_CUDA(ctx, cub::DeviceSelect::Flagged(q, b, it, free_flag + c0, sel_out, d_nsel, nc, st));
            long long h_n = 0;
            BWAGPU_CUDA(ctx, cudaMemcpyAsync(&h_n, d_nsel, 8, cudaMemcpyDeviceToHost, st));
            BWAGPU_CUDA(ctx, cudaStreamSynchronize(st));
            const int64_t take = std::min<int64_t>(h_n, (int64_t)tot_extra - got);
            BWAGPU_CUDA(ctx, cudaMemcpyAsync(empties + got, sel_out, (size_t)take * 4, cudaMemcpyDeviceToDevice, st));
            got += take;
        }
        BWAGPU_CUDA(ctx, cudaStreamSynchronize(st));
        tmp.drop(q); tmp.drop(sel_out);
        if (got < (int64_t)tot_extra) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "emf_build_device: table full (slack %.3f too small)", slack);
    }
    tmp.drop(occupied);
    // the tables, pass B
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_emf_seed, (size_t)n_entry * 16));
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_emf_loc, (size_t)n_loc * 4));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(ctx->d_emf_loc, 0, (size_t)n_loc * 4, st));
    eb_init_table_kernel<<<(unsigned)(((int64_t)n_entry + 255) / 256), 256, 0, st>>>((uint4 *)ctx->d_emf_seed, n_entry);
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_stats + 4, 0, 8, st));
    eb_fill_kernel<<<(unsigned)((n_valid + 255) / 256), 256, 0, st>>>(t, keys2, vals2, n_valid, extra, mainw, empties, (uint4 *)ctx->d_emf_seed,
                                                                      (uint32_t *)ctx->d_emf_loc, 1u + tot_main, d_stats + 4);
    BWAGPU_CUDA(ctx, cudaStreamSynchronize(st));
    BWAGPU_CUDA(ctx, cudaGetLastError());
    // occupied slots = runs: every entry but the collision ones has its own slot
    BWAGPU_CUDA(ctx, cudaMemcpy(h_stats, d_stats, 64, cudaMemcpyDeviceToHost));
    memset(&ctx->emf, 0, sizeof ctx->emf);
    ctx->emf.seed_len = seed_len; ctx->emf.num_loc_entry = n_loc; ctx->emf.num_seed_entry = n_entry; ctx->emf.num_seed_load = n_entry;
    ctx->emf.seq_len = (uint32_t)l_pac; ctx->emf.num_seed_key = (uint32_t)h_stats[0]; ctx->emf.num_seed_used = (uint32_t)(h_stats[0] + tot_extra);
    ctx->launches += 8;
    return BWAGPU_OK;
}

extern "C" int bwagpu_emf_get(bwagpu_ctx *ctx, bwagpu_emf_desc *desc, uint32_t *loc_table, bwagpu_seed_entry *seed_table)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!ctx->d_emf_seed) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "emf_get: no table resident");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    if (desc) { *desc = ctx->emf; if (desc->num_seed_load == desc->num_seed_entry) desc->num_seed_load = 0; }
    if (loc_table) BWAGPU_CUDA(ctx, cudaMemcpyAsync(loc_table, ctx->d_emf_loc, (size_t)ctx->emf.num_loc_entry * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (seed_table) BWAGPU_CUDA(ctx, cudaMemcpyAsync(seed_table, ctx->d_emf_seed, (size_t)ctx->emf.num_seed_load * 16, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}
