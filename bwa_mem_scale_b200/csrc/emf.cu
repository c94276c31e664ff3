// EMF (exact-match filter) stage of the C ABI: bwagpu_emf_load / _unload / _probe{,_device} and the
// offline host builder bwagpu_emf_build (include/bwagpu.h).
//
// Probe = find_perfect_match_entry (reference src/perfect_map.cpp:638-659): a read that occurs
// verbatim (either strand) in the reference is answered from a hash table keyed by the
// "canonical" orientation of its first seed_len bases and never reaches seeding or extension.
//
// emf_probe_kernel: ONE WARP PER READ.  Every string operation of the probe is a lane-parallel
// pass over bases (lane l takes bases l, l+32, ...):
//   N test               OR of (code & 0xC) over the read               -> __any_sync
//   canonical strand     lexicographic compare of first half vs revcomp -> ballot + first set lane
//   hash                 2-bit pack of 32 bases per 64-bit word         -> __reduce_or_sync halves
//   table walk           16-byte entry broadcast load, compare of the reference text at
//                        ent.location with the read (coalesced 32-byte sectors), left/right hop
//   longer reads         remainder compare at the primary and every multi-location
// HBM traffic per read is its own bytes + one 32-byte sector per table hop + ~6 sectors of
// reference text per compare: the kernel is bound by dependent-load latency (3-4 round trips per
// read), hidden by 64 resident warps per SM.
#include <stdlib.h>
#include <algorithm>
#include <thread>
#include <vector>
#include "ctx.cuh"

#define EMF_NO_ENTRY 0xFFFFFFFFu
#define EMF_FULL 0xFFFFFFFFu

struct EmfView {
    const uint8_t *ref;          // .0123 text
    const uint32_t *loc;         // loc_table
    const uint4 *seed;           // seed_table (flags, location, left, right)
    int L;                       // seed_len
    uint32_t n_entry, n_load, seq_len;
};

// lexicographic compare of two length-n strings given as per-position functors; warp-uniform result
template <class FA, class FB>
__device__ __forceinline__ int emf_warp_cmp(int n, int lane, FA a, FB b)
{
    for (int base = 0; base < n; base += 32) {
        const int j = base + lane;
        int x = 0, y = 0;
        if (j < n) { x = a(j); y = b(j); }
        const unsigned m = __ballot_sync(EMF_FULL, x != y);
        if (m) return __shfl_sync(EMF_FULL, x > y ? 1 : -1, __ffs(m) - 1);
    }
    return 0;
}

template <class FA, class FB>
__device__ __forceinline__ bool emf_warp_equal(int n, int lane, FA a, FB b)
{
    for (int base = 0; base < n; base += 32) {
        const int j = base + lane;
        const bool ne = j < n && a(j) != b(j);
        if (__any_sync(EMF_FULL, ne)) return false;
    }
    return true;
}

__device__ __forceinline__ uint64_t emf_fmix64(uint64_t k)      // perfect.h:425-433
{
    k ^= k >> 33; k *= 0xff51afd7ed558ccdULL;
    k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL;
    k ^= k >> 33;
    return k;
}

// __seedmatch_further (perfect.h:475-491): the read's bases beyond seed_len against the text next to loc
__device__ __forceinline__ bool emf_further(const EmfView &v, uint32_t loc, const uint8_t *q, int n, int is_rev, int lane)
{
    const int extra = n - v.L;
    if (!is_rev) {
        if ((uint64_t)loc + (uint64_t)extra >= v.seq_len) return false;
        const uint8_t *r = v.ref + loc + v.L, *s = q + v.L;
        return emf_warp_equal(extra, lane, [&](int j) { return (int)r[j]; }, [&](int j) { return (int)s[j]; });
    }
    if (loc < (uint32_t)extra) return false;
    const uint8_t *r = v.ref + loc - extra;
    return emf_warp_equal(extra, lane, [&](int j) { return (int)r[j]; }, [&](int j) { return 3 - (q[n - 1 - j] & 3); });
}

__global__ void __launch_bounds__(256)
emf_probe_kernel(EmfView v, const uint8_t *__restrict__ reads, const int64_t *__restrict__ off,
                 const int32_t *__restrict__ len, int nseq, int32_t *__restrict__ ret,
                 uint2 *__restrict__ perfect, uint8_t *__restrict__ skip)
{
    const int lane = threadIdx.x & 31;
    const int nwarp = (gridDim.x * blockDim.x) >> 5;
    for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < nseq; r += nwarp) {
        const uint8_t *q = reads + off[r];
        const int n = len[r];
        const int L = v.L;
        int code = BWAGPU_EMF_NOT_MATCHED;
        uint32_t o_flags = 0, o_loc = 0;
        if (v.seed == nullptr || n < L) code = BWAGPU_EMF_NO_TABLE;
        else {
            int bad = 0;
            for (int j = lane; j < n; j += 32) bad |= q[j] & 0xC;
            if (__any_sync(EMF_FULL, bad)) code = BWAGPU_EMF_WITH_N;
        }
        if (code == BWAGPU_EMF_NOT_MATCHED) {
            // __compare_fw_rc (perfect.h:345-351): first halves of the two orientations only
            const int half = (L + 1) / 2;
            const int fw_less = emf_warp_cmp(half, lane, [&](int j) { return (int)q[j]; },
                                             [&](int j) { return 3 - (int)q[L - 1 - j]; }) <= 0;
            auto canon = [&](int j) { return fw_less ? (int)q[j] : 3 - (int)q[L - 1 - j]; };
            // __get_hash_idx_fw / _rc (perfect.h:466-706)
            uint64_t h = 0;
            for (int base = 0; base < L; base += 32) {
                const int nb = min(32, L - base);
                uint64_t c = 0;
                if (lane < nb) c = (uint64_t)(canon(base + lane) & 3) << (2 * (nb - 1 - lane));
                const uint32_t hi = __reduce_or_sync(EMF_FULL, (uint32_t)(c >> 32));
                const uint32_t lo = __reduce_or_sync(EMF_FULL, (uint32_t)c);
                h ^= ((uint64_t)hi << 32) | lo;
            }
            uint32_t idx = (uint32_t)(emf_fmix64(h) % v.n_entry);
            bool first = true;
            while (idx < v.n_load) {                                       // get_seed_entry (perfect.h:749-755)
                const uint4 e = __ldg(v.seed + idx);                       // x flags, y location, z left, w right
                if (first && (e.y == EMF_NO_ENTRY || (e.x & 2))) break;    // is_hash_matched_entry
                first = false;
                const uint8_t *t = v.ref + e.y;
                const int efw = e.x & 1;
                const int cmp = emf_warp_cmp(L, lane, [&](int j) { return efw ? (int)t[j] : 3 - (int)t[L - 1 - j]; }, canon);
                if (cmp == 0) {
                    int is_rev = efw == fw_less ? 0 : 1;
                    uint32_t loc = EMF_NO_ENTRY;
                    if (n == L || emf_further(v, e.y, q, n, is_rev, lane)) loc = e.y;
                    else if (e.x >> 2) {                                   // seedmatch_further (perfect_map.cpp:528-581)
                        const uint32_t ml = e.x >> 2, head = __ldg(v.loc + ml);
                        uint32_t st, nfw, nrc;
                        if (head & 0x80000000u) { st = head & 0x7FFFFFFFu; nfw = __ldg(v.loc + st); nrc = __ldg(v.loc + st + 1); st += 2; }
                        else { nfw = (head >> 16) & 0xFFFF; nrc = head & 0xFFFF; st = ml + 1; }
                        for (uint32_t i = 0; i < nfw && loc == EMF_NO_ENTRY; ++i) {
                            const uint32_t c = __ldg(v.loc + st + i);
                            if (emf_further(v, c, q, n, is_rev, lane)) loc = c;
                        }
                        if (loc == EMF_NO_ENTRY) {
                            is_rev = !is_rev;
                            for (uint32_t i = 0; i < nrc && loc == EMF_NO_ENTRY; ++i) {
                                const uint32_t c = __ldg(v.loc + st + nfw + i);
                                if (emf_further(v, c, q, n, is_rev, lane)) loc = c;
                            }
                        }
                    }
                    if (loc == EMF_NO_ENTRY) code = BWAGPU_EMF_SEED_ONLY_MATCHED;
                    else {
                        o_loc = loc;
                        o_flags = is_rev ? (e.x | 3u) : ((e.x & ~2u) | 1u);
                        code = is_rev ? BWAGPU_EMF_RC_MATCHED : BWAGPU_EMF_FW_MATCHED;
                    }
                    break;
                }
                idx = cmp > 0 ? e.z : e.w;
            }
        }
        if (lane == 0) {
            ret[r] = code;
            perfect[r] = make_uint2(o_flags, o_loc);
            if (skip) skip[r] = code == BWAGPU_EMF_FW_MATCHED || code == BWAGPU_EMF_RC_MATCHED;
        }
    }
}

// ---------------------------------------------------------------------------------------------------
extern "C" int bwagpu_emf_unload(bwagpu_ctx *ctx)
{
    BWAGPU_OWNS_INDEX(ctx);
    if (!ctx) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    if (ctx->d_emf_loc) cudaFree(ctx->d_emf_loc);
    if (ctx->d_emf_seed) cudaFree(ctx->d_emf_seed);
    ctx->d_emf_loc = ctx->d_emf_seed = nullptr;
    memset(&ctx->emf, 0, sizeof(ctx->emf));
    return BWAGPU_OK;
}

extern "C" int bwagpu_emf_load(bwagpu_ctx *ctx, const bwagpu_emf_desc *d, const uint32_t *loc_table,
                               const bwagpu_seed_entry *seed_table)
{
    BWAGPU_OWNS_INDEX(ctx);
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!d || !loc_table || !seed_table || d->seed_len <= 0 || d->num_seed_entry == 0 || d->num_loc_entry == 0)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "emf_load: incomplete table descriptor");
    if (!ctx->d_ref_string) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "emf_load: load the reference text first (bwagpu_index_load_ref)");
    if ((int64_t)d->seq_len != ctx->l_pac)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "emf_load: table built for seq_len %u, resident reference has l_pac %lld", d->seq_len, (long long)ctx->l_pac);
    int rc = bwagpu_emf_unload(ctx);
    if (rc) return rc;
    ctx->emf = *d;
    if (ctx->emf.num_seed_load == 0 || ctx->emf.num_seed_load > d->num_seed_entry) ctx->emf.num_seed_load = d->num_seed_entry;
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_emf_loc, (size_t)d->num_loc_entry * 4));
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_emf_seed, (size_t)ctx->emf.num_seed_load * 16));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_emf_loc, loc_table, (size_t)d->num_loc_entry * 4, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_emf_seed, seed_table, (size_t)ctx->emf.num_seed_load * 16, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}

extern "C" int bwagpu_emf_probe_device(bwagpu_ctx *ctx, const uint8_t *d_reads, const int64_t *d_off, const int32_t *d_len,
                                       int32_t nseq, int32_t *d_ret, bwagpu_perfect *d_perfect, uint8_t *d_skip)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (nseq < 0 || (nseq > 0 && (!d_reads || !d_off || !d_len || !d_ret || !d_perfect)))
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "emf_probe: null buffer");
    if (nseq == 0) return BWAGPU_OK;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    EmfView v;
    v.ref = (const uint8_t *)ctx->d_ref_string;
    v.loc = (const uint32_t *)ctx->d_emf_loc;
    v.seed = (const uint4 *)ctx->d_emf_seed;       // nullptr = no table: every read answers NO_TABLE
    v.L = ctx->emf.seed_len; v.n_entry = ctx->emf.num_seed_entry; v.n_load = ctx->emf.num_seed_load; v.seq_len = ctx->emf.seq_len;
    const int wpb = 8;
    const int64_t want = ((int64_t)nseq + wpb - 1) / wpb;
    const int64_t cap = (int64_t)ctx->sm_count * 8 * 4;
    const unsigned grid = (unsigned)std::max<int64_t>(1, std::min(want, cap));
    BWAGPU_KERNEL_BEGIN(ctx);
    emf_probe_kernel<<<grid, wpb * 32, 0, ctx->stream>>>(v, d_reads, d_off, d_len, nseq, d_ret, (uint2 *)d_perfect, d_skip);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_EMF);
    return BWAGPU_OK;
}

extern "C" int bwagpu_emf_probe(bwagpu_ctx *ctx, const uint8_t *reads, const int64_t *off, const int32_t *len,
                                int32_t nseq, int32_t *ret, bwagpu_perfect *perfect, uint8_t *skip)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (nseq < 0 || (nseq > 0 && (!reads || !off || !len || !ret || !perfect))) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "emf_probe: null buffer");
    if (nseq == 0) return BWAGPU_OK;
    int64_t total = 0;
    for (int32_t i = 0; i < nseq; ++i) {
        if (len[i] < 0 || off[i] < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "emf_probe: negative offset/length at read %d", i);
        total = std::max(total, off[i] + len[i]);
    }
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    void *d_reads, *d_off, *d_len, *d_ret, *d_perf, *d_skip;
    int rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_READS, total + 64, &d_reads))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_OFF, (int64_t)nseq * 8, &d_off))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_LEN, (int64_t)nseq * 4, &d_len))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_MISC0, (int64_t)nseq * 4, &d_ret))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_MISC1, (int64_t)nseq * 8, &d_perf))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_SKIP, (int64_t)nseq, &d_skip))) return rc;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_reads, reads, (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_off, off, (size_t)nseq * 8, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_len, len, (size_t)nseq * 4, cudaMemcpyHostToDevice, ctx->stream));
    rc = bwagpu_emf_probe_device(ctx, (const uint8_t *)d_reads, (const int64_t *)d_off, (const int32_t *)d_len, nseq,
                                 (int32_t *)d_ret, (bwagpu_perfect *)d_perf, (uint8_t *)d_skip);
    if (rc) return rc;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(ret, d_ret, (size_t)nseq * 4, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(perfect, d_perf, (size_t)nseq * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (skip) BWAGPU_CUDA(ctx, cudaMemcpyAsync(skip, d_skip, (size_t)nseq, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}

// ---------------------------------------------------------------------------------------------------
// Offline builder (host).  The table layout is decided by the ORDER of insertions (linear probing
// for collision entries, eviction of a collision entry when the owner of its slot arrives), so the
// insertion pass is sequential, as in the reference (src/perfect_index.cpp:1166-1232); only the
// per-location hash keys are computed in parallel beforehand.
namespace {

struct EmfBuilder {
    const uint8_t *ref;
    int L;
    uint32_t n_entry;
    bwagpu_seed_entry *T;
    std::vector<uint32_t> key;                 // hash slot of every location (EMF_NO_ENTRY = not a seed)
    std::vector<uint8_t> fwl;                  // canonical-strand bit of every location
    struct Bloc { std::vector<uint32_t> fw, rc; };
    std::vector<Bloc> blocs;                   // multi-location lists, id 0 unused
    uint32_t used = 0, keys = 0;

    int base(uint32_t loc, int fw, int j) const { return fw ? ref[loc + j] : 3 - ref[loc + L - 1 - j]; }
    int cmp(uint32_t a, int afw, uint32_t b, int bfw, int n) const       // both strings of length L, first n bases
    {
        for (int j = 0; j < n; ++j) { const int x = base(a, afw, j), y = base(b, bfw, j); if (x != y) return x > y ? 1 : -1; }
        return 0;
    }
    uint32_t hash_slot(uint32_t loc, int fw) const
    {
        uint64_t h = 0, s = 0;
        int j = 0;
        for (; j + 32 <= L; j += 32) { s = 0; for (int t = 0; t < 32; ++t) s = (s << 2) | (uint64_t)base(loc, fw, j + t); h ^= s; }
        if (j < L) { s = 0; for (; j < L; ++j) s = (s << 2) | (uint64_t)base(loc, fw, j); h ^= s; }
        h ^= h >> 33; h *= 0xff51afd7ed558ccdULL; h ^= h >> 33; h *= 0xc4ceb9fe1a85ec53ULL; h ^= h >> 33;
        return (uint32_t)(h % n_entry);
    }
    // __seedmatch (perfect.h:412-424): 1 = same text, 2 = reverse complement, 0 = different
    int match(uint32_t a, uint32_t b) const
    {
        if (memcmp(ref + a, ref + b, (size_t)L) == 0) return 1;
        for (int j = 0; j < L; ++j) if (ref[a + j] != 3 - ref[b + L - 1 - j]) return 0;
        return 2;
    }
    uint32_t empty_from(uint32_t from) const                              // get_empty_idx (:215-229)
    {
        if (T[from].location == EMF_NO_ENTRY) return from;
        for (uint32_t i = (from + 1) % n_entry; i != from; i = (i + 1) % n_entry)
            if (T[i].location == EMF_NO_ENTRY) return i;
        return EMF_NO_ENTRY;
    }
    static void init(bwagpu_seed_entry &e, uint32_t loc, int fw_less, int collision)
    {
        e.flags = (fw_less ? 1u : 0u) | (collision ? 2u : 0u); e.location = loc; e.left = e.right = EMF_NO_ENTRY;
    }
    // __add_to_hash (:614-757)
    bool add(uint32_t loc)
    {
        const uint32_t k = key[loc];
        bwagpu_seed_entry &home = T[k];
        if (home.flags & 2) {                     // a foreign collision entry sits in our slot: evict it
            const uint32_t to = empty_from(k);
            if (to == EMF_NO_ENTRY) return false;
            T[to] = home;
            uint32_t p = key[home.location];      // head of the chain the evicted entry belongs to
            while (T[p].right != k && T[p].right != EMF_NO_ENTRY) p = T[p].right;
            T[p].right = to;
            init(home, EMF_NO_ENTRY, 0, 0);
        }
        if (home.location == EMF_NO_ENTRY) { init(home, loc, fwl[loc], 0); ++used; ++keys; return true; }
        uint32_t prev = EMF_NO_ENTRY, cur = k;
        int m = 0;
        while (cur != EMF_NO_ENTRY) {
            m = match(loc, T[cur].location);
            if (m) break;
            prev = cur; cur = T[cur].right;
        }
        if (!m) {
            const uint32_t to = empty_from(prev);
            if (to == EMF_NO_ENTRY) return false;
            init(T[to], loc, fwl[loc], 1);
            T[prev].right = to;
            ++used;
            return true;
        }
        uint32_t ml = T[cur].flags >> 2;
        if (ml == 0) {
            if (blocs.empty()) blocs.resize(1);
            ml = (uint32_t)blocs.size();
            blocs.emplace_back();
            T[cur].flags = (T[cur].flags & 3u) | (ml << 2);
        }
        (m == 1 ? blocs[ml].fw : blocs[ml].rc).push_back(loc);
        return true;
    }
};

}  // namespace

extern "C" void bwagpu_emf_free_tables(uint32_t *loc_table, bwagpu_seed_entry *seed_table) { free(loc_table); free(seed_table); }

extern "C" int bwagpu_emf_build(const uint8_t *ref_string, int64_t l_pac, int32_t n_seqs, const int64_t *contig_off,
                                const int32_t *contig_len, int32_t n_holes, const int64_t *hole_off, const int32_t *hole_len,
                                int32_t seed_len, double slack, int32_t n_threads, bwagpu_emf_desc *desc,
                                uint32_t **loc_table, bwagpu_seed_entry **seed_table)
{
    if (!ref_string || !desc || !loc_table || !seed_table || n_seqs <= 0 || !contig_off || !contig_len || seed_len <= 0 ||
        (n_holes > 0 && (!hole_off || !hole_len)) || l_pac <= 0 || l_pac >= (int64_t)0xFFFFFFFFu)
        return BWAGPU_ERR_ARG;
    const uint64_t n_entry64 = (uint64_t)((double)l_pac * slack);          // perfect_index.cpp:1356
    if (n_entry64 == 0 || n_entry64 > 0xFFFFFFFFull) return BWAGPU_ERR_RANGE;
    EmfBuilder b;
    b.ref = ref_string; b.L = seed_len; b.n_entry = (uint32_t)n_entry64;
    b.T = (bwagpu_seed_entry *)malloc((size_t)b.n_entry * sizeof(bwagpu_seed_entry));
    if (!b.T) return BWAGPU_ERR_NOMEM;
    for (uint32_t i = 0; i < b.n_entry; ++i) EmfBuilder::init(b.T[i], EMF_NO_ENTRY, 0, 0);
    // which locations start a seed: inside one contig and clear of every hole (calc_loc_key :1045-1164)
    b.key.assign((size_t)l_pac, EMF_NO_ENTRY);
    b.fwl.assign((size_t)l_pac, 0);
    std::vector<uint8_t> ok((size_t)l_pac, 0);
    for (int32_t s = 0; s < n_seqs; ++s)
        for (int64_t p = contig_off[s]; p + seed_len <= contig_off[s] + contig_len[s] && p < l_pac; ++p) ok[(size_t)p] = 1;
    for (int32_t h = 0; h < n_holes; ++h)
        for (int64_t p = std::max<int64_t>(0, hole_off[h] - seed_len + 1); p < hole_off[h] + hole_len[h] && p < l_pac; ++p) ok[(size_t)p] = 0;
    const int nt = std::max(1, std::min(n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency(), 64));
    {
        std::vector<std::thread> th;
        for (int t = 0; t < nt; ++t)
            th.emplace_back([&, t]() {
                const int64_t lo = l_pac * t / nt, hi = l_pac * (t + 1) / nt;
                for (int64_t p = lo; p < hi; ++p) {
                    if (!ok[(size_t)p]) continue;
                    const int half = (seed_len + 1) / 2;
                    const int fw = b.cmp((uint32_t)p, 1, (uint32_t)p, 0, half) <= 0;      // __compare_fw_rc
                    b.fwl[(size_t)p] = (uint8_t)fw;
                    b.key[(size_t)p] = b.hash_slot((uint32_t)p, fw);
                }
            });
        for (auto &t : th) t.join();
    }
    for (int64_t p = 0; p < l_pac; ++p)
        if (ok[(size_t)p] && !b.add((uint32_t)p)) { free(b.T); return BWAGPU_ERR_RANGE; }     // table full: slack < 1?

    // ---- rebuild_perfect_table_for_mapping (:926-990) ----
    // 1. multi-location lists -> loc_table (build_loc_to_loc_table :308-397)
    const uint32_t nb = (uint32_t)std::max<size_t>(b.blocs.size(), 1);
    uint32_t n = 1, i_many = 1;
    for (uint32_t i = 1; i < nb; ++i) {
        const uint32_t c = (uint32_t)(b.blocs[i].fw.size() + b.blocs[i].rc.size());
        if (b.blocs[i].fw.size() < 256 && b.blocs[i].rc.size() < 256) { n += 1 + c; i_many += 1 + c; }
        else { n += 3 + c; i_many += 1; }
    }
    n = n + (16 - (n % 16));
    uint32_t *lt = (uint32_t *)calloc(n, sizeof(uint32_t));
    if (!lt) { free(b.T); return BWAGPU_ERR_NOMEM; }
    std::vector<uint32_t> remap(nb, 0);
    uint32_t w = 1;
    for (uint32_t i = 1; i < nb; ++i) {
        const auto &bl = b.blocs[i];
        uint32_t at;
        remap[i] = w;
        if (bl.fw.size() < 256 && bl.rc.size() < 256) {
            lt[w++] = ((uint32_t)bl.fw.size() << 16) + (uint32_t)bl.rc.size();
            at = w; w += (uint32_t)(bl.fw.size() + bl.rc.size());
        } else {
            lt[w++] = 0x80000000u | i_many;
            lt[i_many++] = (uint32_t)bl.fw.size();
            lt[i_many++] = (uint32_t)bl.rc.size();
            at = i_many; i_many += (uint32_t)(bl.fw.size() + bl.rc.size());
        }
        for (uint32_t x : bl.fw) lt[at++] = x;
        for (uint32_t x : bl.rc) lt[at++] = x;
    }
    // 2. every collision chain -> balanced search tree rooted in the hash slot (:861-924)
    std::vector<uint32_t> slots;
    std::vector<bwagpu_seed_entry> nodes;
    auto fix_ml = [&](bwagpu_seed_entry &e) { const uint32_t ml = e.flags >> 2; if (ml) e.flags = (e.flags & 3u) | (remap[ml] << 2); };
    for (uint32_t i = 0; i < b.n_entry; ++i) {
        bwagpu_seed_entry &e = b.T[i];
        if (e.location == EMF_NO_ENTRY || (e.flags & 2)) continue;
        if (e.right == EMF_NO_ENTRY) { fix_ml(e); continue; }
        slots.clear(); nodes.clear();
        for (uint32_t c = i; c != EMF_NO_ENTRY; c = b.T[c].right) { slots.push_back(c); nodes.push_back(b.T[c]); }
        for (auto &x : nodes) fix_ml(x);
        // the seeds of one chain are pairwise different, so any sort gives the reference's order
        std::sort(nodes.begin(), nodes.end(), [&](const bwagpu_seed_entry &x, const bwagpu_seed_entry &y) {
            return b.cmp(x.location, x.flags & 1, y.location, y.flags & 1, seed_len) < 0; });
        size_t next = 1;
        // pre-order placement; both children take their slots before either subtree is laid out
        struct Job { uint32_t slot; int lo, hi; };
        std::vector<Job> stack{{slots[0], 0, (int)nodes.size() - 1}};
        while (!stack.empty()) {
            const Job j = stack.back(); stack.pop_back();
            const int mid = (j.lo + j.hi) / 2;
            bwagpu_seed_entry &d = b.T[j.slot];
            d = nodes[mid];
            d.left = mid > j.lo ? slots[next++] : EMF_NO_ENTRY;
            d.right = mid < j.hi ? slots[next++] : EMF_NO_ENTRY;
            if (d.right != EMF_NO_ENTRY) stack.push_back({d.right, mid + 1, j.hi});   // left subtree is laid out first
            if (d.left != EMF_NO_ENTRY) stack.push_back({d.left, j.lo, mid - 1});
        }
        for (size_t c = 0; c < slots.size(); ++c) b.T[slots[c]].flags = (b.T[slots[c]].flags & ~2u) | (c ? 2u : 0u);
    }
    desc->seed_len = seed_len; desc->num_loc_entry = n; desc->num_seed_entry = b.n_entry; desc->num_seed_load = 0;
    desc->seq_len = (uint32_t)l_pac; desc->num_seed_used = b.used; desc->num_seed_key = b.keys;
    *loc_table = lt; *seed_table = b.T;
    return BWAGPU_OK;
}
