// ERT seeding stage of the C ABI: bwagpu_index_load_ert, bwagpu_seed_ert{,_device}  (include/bwagpu.h).
// Replaces the per-read seeding body of mem_kernel1_core_ert (reference src/bwamem.cpp:1124-1193); the
// per-read algorithm lives in ert_core.cuh.
//
// One read per thread.  A read's walk is a chain of dependent small loads (8-byte k-mer entry, then
// 1-5 byte tree records, then a few text bytes), about 0.5-1 KB per read in total, so the stage is bound
// by load latency, not bandwidth; parallelism comes from the number of reads in flight.  Each read writes
// its MEMs and hits into a private slab; the few reads that outgrow the slab (hits inside long repeats) are
// collected in a list and re-run by a second launch with slabs sized for the largest of them, so the common
// slab stays small.  A second kernel turns the slabs into the batch-wide, per-read sorted MEM list and the
// coordinate list in the layout of the FM-index path (bwagpu_seed_fm_device), which is what chaining reads.
#include <cub/device/device_scan.cuh>
#include <limits.h>
#include <stdlib.h>
#include "ctx.cuh"
#include "ert_core.cuh"
#include "ert_build.cuh"

#ifndef ERT_TPB
#define ERT_TPB 128
#endif
#ifndef ERT_MINB
#define ERT_MINB 1
#endif
#define ERT_KMER_ENTRIES (1LL << 30)

struct ErtStats { unsigned long long n_bytes, n_hits; unsigned int max_mem, max_hit, n_ovf, pad_; int bad_site, bad_read; long long bad_val; unsigned long long bad_u, bad_v; };

struct ErtSlab { ErtMemRec *mems; int64_t *hits; uint64_t *keys; int mem_cap, hit_cap; };

__global__ void __launch_bounds__(ERT_TPB, ERT_MINB)
ert_seed_kernel(ErtView ix, ErtOpt o, int max_occ, const uint8_t *__restrict__ reads, const int64_t *__restrict__ off,
                const int32_t *__restrict__ len, const uint8_t *__restrict__ skip, const int32_t *__restrict__ list,
                int n_items, ErtSlab sl, int32_t *cnt_mem, int64_t *cnt_coord, int32_t *slab_of, int32_t *ovf_list,
                ErtStats *st)
{
    const int item = blockIdx.x * blockDim.x + threadIdx.x;
    if (item >= n_items) return;
    const int r = list ? list[item] : item;
    if (!list) { cnt_mem[r] = 0; cnt_coord[r] = 0; slab_of[r] = -1; }
    if ((skip && skip[r]) || len[r] > ERT_MAX_READ) return;
    ErtWalker w;
    w.ix = ix;
    w.fwd = reads + off[r]; w.l_seq = len[r];
    w.mems = sl.mems + (int64_t)item * sl.mem_cap; w.mem_cap = sl.mem_cap;
    w.hits = sl.hits + (int64_t)item * sl.hit_cap; w.hit_cap = sl.hit_cap;
    w.seed_read(o);
#ifdef ERT_GUARD
    if (w.bad_site && atomicCAS(&st->bad_site, 0, w.bad_site) == 0) { st->bad_read = r; st->bad_val = w.bad_val; st->bad_u = w.bad_u; st->bad_v = w.bad_v; }
#endif
    if (w.n_mem > sl.mem_cap || w.hit_peak > sl.hit_cap) {
        atomicMax(&st->max_mem, (unsigned)w.n_mem);
        atomicMax(&st->max_hit, (unsigned)w.hit_peak);
        if (!list) ovf_list[atomicAdd(&st->n_ovf, 1u)] = r;
        return;
    }
    uint64_t *keys = sl.keys + (int64_t)item * sl.mem_cap;
    int64_t n_coord = 0, n_hits = 0;
    for (int k = 0; k < w.n_mem; ++k) {
        const ErtMemRec &m = w.mems[k];
        keys[k] = ((uint64_t)m.start << 40) | ((uint64_t)m.end << 20) | (uint64_t)k;
        const int64_t hc = m.hitcount, step = hc > max_occ ? hc / max_occ : 1;
        const int64_t c = (hc + step - 1) / step;
        n_coord += c > max_occ ? max_occ : c;
        n_hits += hc;
    }
    ErtSort::introsort(keys, w.n_mem);
    cnt_mem[r] = w.n_mem; cnt_coord[r] = n_coord;
    if (list) slab_of[r] = item;
    atomicAdd(&st->n_bytes, (unsigned long long)w.n_bytes);
    atomicAdd(&st->n_hits, (unsigned long long)n_hits);
}

// slabs -> batch-wide lists: MEMs of read r at out[mem_off[r] ..] in sorted order, their coordinates behind
// coord_base[r] in MEM order then hit order (every `step`-th hit, at most max_occ: mem_chain_new's loop,
// src/bwamem.cpp:985-987)
__global__ void ert_emit_kernel(int nseq, int max_occ, int64_t l_pac, ErtSlab s1, ErtSlab s2, const int32_t *__restrict__ cnt_mem,
                                const int32_t *__restrict__ slab_of, const int64_t *__restrict__ mem_off,
                                const int64_t *__restrict__ coord_base, bwagpu_smem *__restrict__ out,
                                int64_t *__restrict__ coord_off, int64_t *__restrict__ coords)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nseq) return;
    const int n = cnt_mem[r];
    int64_t c = coord_base[r];
    if (r == nseq - 1) coord_off[mem_off[nseq]] = coord_base[nseq];
    if (n == 0) return;
    const int so = slab_of[r];
    const ErtSlab &sl = so < 0 ? s1 : s2;
    const int64_t item = so < 0 ? r : so;
    const ErtMemRec *mems = sl.mems + item * sl.mem_cap;
    const int64_t *hits = sl.hits + item * sl.hit_cap;
    const uint64_t *keys = sl.keys + item * sl.mem_cap;
    const int64_t m0 = mem_off[r];
    for (int k = 0; k < n; ++k) {
        const ErtMemRec m = mems[keys[k] & 0xFFFFF];
        bwagpu_smem o;
        o.rid = (uint32_t)r; o.m = (uint32_t)m.start; o.n = (uint32_t)(m.end - 1);
        o.k = c; o.l = 0; o.s = m.hitcount;
        out[m0 + k] = o;
        coord_off[m0 + k] = c;
        const int64_t hc = m.hitcount, step = hc > max_occ ? hc / max_occ : 1;
        int cnt = 0;
        for (int64_t q = 0; q < hc && cnt < max_occ; q += step, ++cnt)
            coords[c++] = ert_hit_rbeg(m, hits[m.hitbeg + q], l_pac);
    }
}

__global__ void ert_scatter_kernel(const uint32_t *__restrict__ idx, const uint64_t *__restrict__ val, int64_t n,
                                   uint64_t *__restrict__ table)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) table[idx[i]] = val[i];
}

extern "C" int bwagpu_index_unload_ert(bwagpu_ctx *ctx)
{
    BWAGPU_OWNS_INDEX(ctx);
    if (!ctx) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    if (ctx->d_ert_kmer) cudaFree(ctx->d_ert_kmer);
    if (ctx->d_ert_mlt) cudaFree(ctx->d_ert_mlt);
    ctx->d_ert_kmer = ctx->d_ert_mlt = nullptr;
    ctx->ert_mlt_bytes = 0;
    return BWAGPU_OK;
}

extern "C" int bwagpu_index_load_ert(bwagpu_ctx *ctx, const uint64_t *kmer_offsets, const uint32_t *sparse_idx,
                                     const uint64_t *sparse_val, int64_t n_sparse, const uint8_t *mlt_table,
                                     int64_t mlt_bytes)
{
    BWAGPU_OWNS_INDEX(ctx);
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!mlt_table || mlt_bytes <= 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_load_ert: mlt_table is empty");
    if (!kmer_offsets && (n_sparse < 0 || (n_sparse > 0 && (!sparse_idx || !sparse_val))))
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_load_ert: neither a dense k-mer table nor sparse entries");
    int rc = bwagpu_index_unload_ert(ctx);
    if (rc) return rc;
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_ert_kmer, (size_t)ERT_KMER_ENTRIES * 8));
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_ert_mlt, (size_t)mlt_bytes + 16));
    if (kmer_offsets) {
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_ert_kmer, kmer_offsets, (size_t)ERT_KMER_ENTRIES * 8, cudaMemcpyHostToDevice, ctx->stream));
    } else {
        BWAGPU_CUDA(ctx, cudaMemsetAsync(ctx->d_ert_kmer, 0, (size_t)ERT_KMER_ENTRIES * 8, ctx->stream));
        if (n_sparse > 0) {
            uint32_t *d_idx; uint64_t *d_val;
            if ((rc = bwagpu_scratch(ctx, SCR_MISC0, n_sparse * 4, (void **)&d_idx))) return rc;
            if ((rc = bwagpu_scratch(ctx, SCR_MISC1, n_sparse * 8, (void **)&d_val))) return rc;
            BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_idx, sparse_idx, (size_t)n_sparse * 4, cudaMemcpyHostToDevice, ctx->stream));
            BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_val, sparse_val, (size_t)n_sparse * 8, cudaMemcpyHostToDevice, ctx->stream));
            ert_scatter_kernel<<<(unsigned)((n_sparse + 255) / 256), 256, 0, ctx->stream>>>(d_idx, d_val, n_sparse, (uint64_t *)ctx->d_ert_kmer);
            ctx->launches++;
            BWAGPU_CUDA(ctx, cudaGetLastError());
        }
    }
    BWAGPU_CUDA(ctx, cudaMemsetAsync((uint8_t *)ctx->d_ert_mlt + mlt_bytes, 0, 16, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_ert_mlt, mlt_table, (size_t)mlt_bytes, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    ctx->ert_mlt_bytes = mlt_bytes;
    return BWAGPU_OK;
}

// ---- k-mer table entries of absent k-mers, from the FM-index ---------------------------------------------
// One thread per 11-base prefix (the low 22 bits of the k-mer index); the 4^4 completions are enumerated depth
// first so the prefix walk is shared and dead branches are cut: a prefix that no longer occurs gives all of
// its completions the same LEP mask.  A warp's threads hold consecutive prefixes, so for a fixed completion
// their 32 table entries are contiguous (256-byte stores).
#define ERT_FILL_PREFIX 11
__global__ void __launch_bounds__(128) ert_kmer_fill_kernel(FmView f, uint64_t *__restrict__ table)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;          // 4^11 threads exactly
    ErtKmerWalk w0, w1, w2, w3, w4;                                    // walk state after 11, 12, 13, 14, 15 bases
    w0.start(f, (int)(p & 3));
    for (int i = 1; i < ERT_FILL_PREFIX && w0.hits > 0; ++i) w0.step(f, i, (int)((p >> (2 * i)) & 3));
    if (w0.hits == 0) {
        const uint64_t e = (uint64_t)(w0.lep & 0x3FFF) << 2;
        for (uint32_t sfx = 0; sfx < 256; ++sfx) table[p | (sfx << 22)] = e;
        return;
    }
    w1 = w2 = w3 = w0;
    for (uint32_t t = 0; t < 256; ++t) {
        // base 11 is the slowest digit of t, base 14 the fastest: a level is re-walked only when its base changes
        const int b11 = (t >> 6) & 3, b12 = (t >> 4) & 3, b13 = (t >> 2) & 3, b14 = t & 3;
        if ((t & 63) == 0) { w1 = w0; w1.step(f, 11, b11); }
        if ((t & 15) == 0) { w2 = w1; if (w1.hits > 0) w2.step(f, 12, b12); }
        if ((t & 3) == 0) { w3 = w2; if (w2.hits > 0) w3.step(f, 13, b13); }
        w4 = w3; if (w3.hits > 0) w4.step(f, 14, b14);
        if (w4.hits == 0) {
            const uint32_t sfx = (uint32_t)(b11 | (b12 << 2) | (b13 << 4) | (b14 << 6));
            table[p | (sfx << 22)] = (uint64_t)(w4.lep & 0x3FFF) << 2;
        }
    }
}

extern "C" int bwagpu_index_complete_ert(bwagpu_ctx *ctx)
{
    BWAGPU_OWNS_INDEX(ctx);
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!ctx->has_fm) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_complete_ert: no FM-index loaded (bwagpu_index_load_fm)");
    if (!ctx->d_ert_kmer) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_complete_ert: no ERT index loaded (bwagpu_index_load_ert)");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    FmView f;
    f.count_tab = nullptr;
    f.cp_occ = (const int64_t *)ctx->d_cp_occ; f.sa_ms_byte = (const int8_t *)ctx->d_sa_ms; f.sa_ls_word = (const uint32_t *)ctx->d_sa_ls;
    for (int i = 0; i < 5; ++i) f.count[i] = ctx->fm_count[i];
    f.sentinel_index = ctx->fm_sentinel; f.ref_seq_len = ctx->fm_ref_seq_len;
    f.fma_all = f.fma_last = nullptr;
    BWAGPU_KERNEL_BEGIN(ctx);
    ert_kmer_fill_kernel<<<(1u << 22) / 128, 128, 0, ctx->stream>>>(f, (uint64_t *)ctx->d_ert_kmer);
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_ERT_EMIT);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}

extern "C" int bwagpu_index_get_ert_kmers(bwagpu_ctx *ctx, const uint32_t *idx, int64_t n, uint64_t *out)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!ctx->d_ert_kmer) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_get_ert_kmers: no ERT index loaded");
    if (n < 0 || (n > 0 && (!idx || !out))) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_get_ert_kmers: bad argument");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    for (int64_t i = 0; i < n; ++i)
        BWAGPU_CUDA(ctx, cudaMemcpy(out + i, (const uint64_t *)ctx->d_ert_kmer + idx[i], 8, cudaMemcpyDeviceToHost));
    return BWAGPU_OK;
}

extern "C" int bwagpu_index_get_ert_table(bwagpu_ctx *ctx, uint64_t *kmer_offsets)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!ctx->d_ert_kmer) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_get_ert_table: no ERT index loaded");
    if (!kmer_offsets) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_get_ert_table: output is NULL");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    BWAGPU_CUDA(ctx, cudaMemcpy(kmer_offsets, ctx->d_ert_kmer, (size_t)ERT_KMER_ENTRIES * 8, cudaMemcpyDeviceToHost));
    return BWAGPU_OK;
}

static int ert_check(bwagpu_ctx *ctx, const bwagpu_seed_opt *opt)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!opt) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "seed_ert: options are NULL");
    if (!ctx->d_ert_kmer) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "seed_ert: no ERT index loaded (bwagpu_index_load_ert)");
    if (!ctx->d_ref_string) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "seed_ert: no reference text loaded (bwagpu_index_load_ref)");
    if (opt->min_seed_len < ERT_KMER + ERT_XMER)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "seed_ert: min_seed_len %d below k-mer + x-mer = 19 (the reference asserts)", opt->min_seed_len);
    if (opt->max_occ < 1) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "seed_ert: max_occ must be positive");
    return BWAGPU_OK;
}

static int ert_scan_i32(bwagpu_ctx *ctx, const int32_t *d_in, int64_t *d_out, int64_t n)
{
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveScan((void *)nullptr, bytes, d_in, d_out, cub::Sum(), (int64_t)0, (int)n, ctx->stream);
    void *tmp;
    int rc = bwagpu_scratch(ctx, SCR_CUB, (int64_t)bytes + 256, &tmp);
    if (rc) return rc;
    BWAGPU_CUDA(ctx, cub::DeviceScan::ExclusiveScan(tmp, bytes, d_in, d_out, cub::Sum(), (int64_t)0, (int)n, ctx->stream));
    ctx->launches += 2;
    return BWAGPU_OK;
}

static int ert_scan_i64(bwagpu_ctx *ctx, const int64_t *d_in, int64_t *d_out, int64_t n)
{
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum((void *)nullptr, bytes, d_in, d_out, (int)n, ctx->stream);
    void *tmp;
    int rc = bwagpu_scratch(ctx, SCR_CUB, (int64_t)bytes + 256, &tmp);
    if (rc) return rc;
    BWAGPU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, bytes, d_in, d_out, (int)n, ctx->stream));
    ctx->launches += 2;
    return BWAGPU_OK;
}

static int ert_slab(bwagpu_ctx *ctx, int slot0, int64_t items, int mem_cap, int hit_cap, ErtSlab *sl)
{
    int rc;
    sl->mem_cap = mem_cap; sl->hit_cap = hit_cap;
    if ((rc = bwagpu_scratch(ctx, slot0, items * mem_cap * (int64_t)sizeof(ErtMemRec), (void **)&sl->mems))) return rc;
    if ((rc = bwagpu_scratch(ctx, slot0 + 1, items * hit_cap * 8, (void **)&sl->hits))) return rc;
    if ((rc = bwagpu_scratch(ctx, slot0 + 2, items * mem_cap * 8, (void **)&sl->keys))) return rc;
    return BWAGPU_OK;
}

// max_occ: hits kept per MEM (mem_chain_new's sampling); INT_MAX keeps all of them
static int ert_seed_dev(bwagpu_ctx *ctx, const bwagpu_seed_opt *opt, int max_occ, const uint8_t *d_reads,
                        const int64_t *d_off, const int32_t *d_len, const uint8_t *d_skip, int32_t nseq,
                        bwagpu_ert_result *res)
{
    memset(res, 0, sizeof *res);
    if (nseq == 0) return BWAGPU_OK;
    ErtView ix;
    ix.kmer_offsets = (const uint64_t *)ctx->d_ert_kmer; ix.mlt = (const uint8_t *)ctx->d_ert_mlt;
    ix.ref = (const uint8_t *)ctx->d_ref_string; ix.l_pac = ctx->l_pac; ix.mlt_bytes = ctx->ert_mlt_bytes;
    ErtOpt o = {opt->min_seed_len, (int)(opt->min_seed_len * opt->split_factor + .499), opt->split_width,
                (int)opt->max_mem_intv};
    if (ctx->ert_mem_cap < 16) ctx->ert_mem_cap = 16;
    if (ctx->ert_hit_cap < 128) ctx->ert_hit_cap = 128;
    int rc;
    ErtSlab s1, s2;
    memset(&s2, 0, sizeof s2);
    if ((rc = ert_slab(ctx, SCR_ERT_MEMS, nseq, ctx->ert_mem_cap, ctx->ert_hit_cap, &s1))) return rc;
    int32_t *cnt_mem, *slab_of, *list; int64_t *cnt_coord, *mem_off, *coord_base; ErtStats *d_st;
    if ((rc = bwagpu_scratch(ctx, SCR_ERT_CNT, (int64_t)(nseq + 1) * 12, (void **)&cnt_coord))) return rc;
    cnt_mem = (int32_t *)(cnt_coord + nseq + 1);
    if ((rc = bwagpu_scratch(ctx, SCR_ERT_SLAB, (int64_t)nseq * 4, (void **)&slab_of))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_ERT_LIST, (int64_t)nseq * 4, (void **)&list))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_SCAN, (int64_t)(nseq + 1) * 8, (void **)&mem_off))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_ERT_COFF, (int64_t)(nseq + 1) * 8, (void **)&coord_base))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_STATS, 256, (void **)&d_st))) return rc;
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_st, 0, sizeof(ErtStats), ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(cnt_mem + nseq, 0, 4, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(cnt_coord + nseq, 0, 8, ctx->stream));
    BWAGPU_KERNEL_BEGIN(ctx);
    ert_seed_kernel<<<(nseq + ERT_TPB - 1) / ERT_TPB, ERT_TPB, 0, ctx->stream>>>(ix, o, max_occ, d_reads, d_off, d_len, d_skip,
                                                                                nullptr, nseq, s1, cnt_mem, cnt_coord, slab_of,
                                                                                list, d_st);
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_ERT_SEED);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    ErtStats st;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&st, d_st, sizeof st, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
#ifdef ERT_GUARD
    fprintf(stderr, "ERT_GUARD pass 1: bad_site=%d read=%d val=0x%llx u=0x%llx v=0x%llx n_ovf=%u max_mem=%u max_hit=%u mlt_bytes=%lld\n", st.bad_site, st.bad_read, (unsigned long long)st.bad_val, st.bad_u, st.bad_v, st.n_ovf, st.max_mem, st.max_hit, (long long)ctx->ert_mlt_bytes);
#endif
    const int n_ovf = (int)st.n_ovf;
    for (int pass = 0; n_ovf > 0; ++pass) {
        // the reads that outgrew the common slab, re-run with room for the largest of them
        int mc = ctx->ert_mem_cap, hc = ctx->ert_hit_cap;
        while (mc < (int64_t)st.max_mem) mc *= 2;
        while (hc < (int64_t)st.max_hit) hc *= 2;
        if (pass > 8 || (int64_t)n_ovf * hc * 8 > (48LL << 30))
            BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "seed_ert: %d reads need %d hit slots each; scratch would exceed 48 GiB", n_ovf, hc);
        if ((rc = ert_slab(ctx, SCR_ERT_MEMS2, n_ovf, mc, hc, &s2))) return rc;
        BWAGPU_CUDA(ctx, cudaMemsetAsync(&d_st->max_mem, 0, 8, ctx->stream));
        BWAGPU_KERNEL_BEGIN(ctx);
        ert_seed_kernel<<<(n_ovf + ERT_TPB - 1) / ERT_TPB, ERT_TPB, 0, ctx->stream>>>(ix, o, max_occ, d_reads, d_off, d_len, d_skip,
                                                                                     list, n_ovf, s2, cnt_mem, cnt_coord, slab_of,
                                                                                     list, d_st);
        BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_ERT_SEED);
        BWAGPU_CUDA(ctx, cudaGetLastError());
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(&st, d_st, sizeof st, cudaMemcpyDeviceToHost, ctx->stream));
        BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
        if ((int64_t)st.max_mem <= mc && (int64_t)st.max_hit <= hc) break;      // hit_peak is exact, so one pass suffices
    }
    if ((rc = ert_scan_i32(ctx, cnt_mem, mem_off, nseq + 1))) return rc;
    if ((rc = ert_scan_i64(ctx, cnt_coord, coord_base, nseq + 1))) return rc;
    int64_t tot[2] = {0, 0};
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&tot[0], mem_off + nseq, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&tot[1], coord_base + nseq, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    bwagpu_smem *out; int64_t *coord_off, *coords;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_OUT, (tot[0] + 1) * (int64_t)sizeof(bwagpu_smem), (void **)&out))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SA_OFF, (tot[0] + 1) * 8, (void **)&coord_off))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SA_COORD, (tot[1] + 1) * 8, (void **)&coords))) return rc;
    BWAGPU_KERNEL_BEGIN(ctx);
    ert_emit_kernel<<<(nseq + 127) / 128, 128, 0, ctx->stream>>>(nseq, max_occ, ctx->l_pac, s1, s2, cnt_mem, slab_of, mem_off,
                                                                coord_base, out, coord_off, coords);
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_ERT_EMIT);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    res->d_smems = out; res->n_smem = tot[0];
    res->d_coords = coords; res->n_coords = tot[1];
    res->d_coord_off = coord_off; res->d_smem_off = mem_off;
    res->n_index_bytes = (int64_t)st.n_bytes; res->n_hits = (int64_t)st.n_hits; res->n_retry_reads = n_ovf;
    return BWAGPU_OK;
}

extern "C" int bwagpu_seed_ert_device(bwagpu_ctx *ctx, const bwagpu_seed_opt *opt, const uint8_t *d_reads,
                                      const int64_t *d_off, const int32_t *d_len, const uint8_t *d_skip,
                                      int32_t nseq, int32_t max_len, bwagpu_ert_result *res)
{
    int rc = ert_check(ctx, opt);
    if (rc) return rc;
    if (!res || nseq < 0 || max_len < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "seed_ert_device: bad argument");
    if (max_len > ERT_MAX_READ)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "seed_ert_device: reads longer than %d are not supported by ERT seeding", ERT_MAX_READ);
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    return ert_seed_dev(ctx, opt, opt->max_occ, d_reads, d_off, d_len, d_skip, nseq, res);
}

extern "C" int bwagpu_seed_ert(bwagpu_ctx *ctx, const bwagpu_seed_opt *opt, const uint8_t *enc_reads, const int64_t *off,
                               const int32_t *len, const uint8_t *skip, int32_t nseq, bwagpu_ert_mem *mems, int64_t mems_cap,
                               int64_t *hits, int64_t hits_cap, int64_t *mem_off, int64_t *n_mems, int64_t *n_hits)
{
    int rc = ert_check(ctx, opt);
    if (rc) return rc;
    if (!n_mems || !n_hits || nseq < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "seed_ert: bad argument");
    *n_mems = 0; *n_hits = 0;
    if (nseq == 0) { if (mem_off) mem_off[0] = 0; return BWAGPU_OK; }
    if (!enc_reads || !off || !len) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "seed_ert: NULL buffer");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    int64_t bytes = 0;
    for (int i = 0; i < nseq; ++i) {
        if (len[i] < 0 || off[i] < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "seed_ert: read %d has a negative length/offset", i);
        if (len[i] > ERT_MAX_READ) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "seed_ert: read %d longer than %d", i, ERT_MAX_READ);
        if (off[i] + len[i] > bytes) bytes = off[i] + len[i];
    }
    uint8_t *d_reads, *d_skip = nullptr; int64_t *d_off; int32_t *d_len;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_READS, bytes + 16, (void **)&d_reads))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_OFF, (int64_t)nseq * 8, (void **)&d_off))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_LEN, (int64_t)nseq * 4, (void **)&d_len))) return rc;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_reads, enc_reads, (size_t)bytes, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_off, off, (size_t)nseq * 8, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_len, len, (size_t)nseq * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (skip) {
        if ((rc = bwagpu_scratch(ctx, SCR_SEED_SKIP, nseq, (void **)&d_skip))) return rc;
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_skip, skip, (size_t)nseq, cudaMemcpyHostToDevice, ctx->stream));
    }
    bwagpu_ert_result r;
    if ((rc = ert_seed_dev(ctx, opt, INT_MAX, d_reads, d_off, d_len, d_skip, nseq, &r))) return rc;
    *n_mems = r.n_smem; *n_hits = r.n_coords;
    if (r.n_smem > mems_cap || r.n_coords > hits_cap)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "seed_ert: %lld MEMs / %lld hits do not fit the capacities %lld / %lld",
                    (long long)r.n_smem, (long long)r.n_coords, (long long)mems_cap, (long long)hits_cap);
    if (mem_off) BWAGPU_CUDA(ctx, cudaMemcpyAsync(mem_off, r.d_smem_off, (size_t)(nseq + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    bwagpu_smem *tmp = nullptr;
    if (r.n_smem > 0) {
        if (!mems) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "seed_ert: mems is NULL");
        tmp = (bwagpu_smem *)malloc((size_t)r.n_smem * sizeof(bwagpu_smem));
        if (!tmp) BWAGPU_FAIL(ctx, BWAGPU_ERR_NOMEM, "seed_ert: host staging allocation failed");
        cudaError_t e = cudaMemcpyAsync(tmp, r.d_smems, (size_t)r.n_smem * sizeof(bwagpu_smem), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess && r.n_coords > 0) {
            if (!hits) { free(tmp); BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "seed_ert: hits is NULL"); }
            e = cudaMemcpyAsync(hits, r.d_coords, (size_t)r.n_coords * 8, cudaMemcpyDeviceToHost, ctx->stream);
        }
        if (e == cudaSuccess) e = bwagpu_stream_sync(ctx);
        if (e != cudaSuccess) { free(tmp); BWAGPU_FAIL(ctx, BWAGPU_ERR_CUDA, "seed_ert: copy back -> %s", cudaGetErrorString(e)); }
        for (int64_t i = 0; i < r.n_smem; ++i) {
            mems[i].rid = (int32_t)tmp[i].rid; mems[i].start = (int32_t)tmp[i].m; mems[i].end = (int32_t)tmp[i].n + 1;
            mems[i].hitcount = (int32_t)tmp[i].s; mems[i].hit_off = tmp[i].k;
        }
        free(tmp);
    }
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}
