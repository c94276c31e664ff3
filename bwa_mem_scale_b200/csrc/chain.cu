// Seed chaining stage of the C ABI: bwagpu_chain_seeds{,_device}  (include/bwagpu.h).
// Replaces mem_chain_seeds + mem_chain_flt as worker_bwt runs them after seeding (reference
// src/bwamem.cpp:1341-1363; functions at :789-936 and :528-646).  One read per thread: the work per
// read is a short strictly sequential insertion into a B-tree whose shape is observable (see
// chain_core.cuh), so there is no intra-read parallelism to use; reads are independent.
// (mem_flt_chained_seeds, :494-526, only acts on reads longer than ~700 bp — min_l = 5.5*log(l) must
// not exceed 0.05*l — and is not part of this stage; such batches are rejected with BWAGPU_ERR_RANGE.)
#include <cub/device/device_scan.cuh>
#include <cub/device/device_radix_sort.cuh>
#include <stdlib.h>
#include "ctx.cuh"
#include "chain_core.cuh"

struct ChainIn {
    const bwagpu_smem *smems; const int64_t *smem_off;      // SMEMs of read r: [smem_off[r], smem_off[r+1])
    const int64_t *coords; const int64_t *coord_off;        // SA coordinates of SMEM i: [coord_off[i], coord_off[i+1])
    const int32_t *len; int32_t nseq; int64_t n_smem;
    int32_t per_read;                                        // mem_chain_new (ERT path): no batch-wide SMEM bound
    int32_t batch;                                           // reads per mem_chain_seeds call upstream (0 = this whole call)
};

__device__ __forceinline__ ChainScratch chain_scratch(bwagpu_seed *sd, WChain *ch, BtNode *nodes, int32_t *ord,
                                                      int32_t *lst, int64_t c0, int r)
{
    ChainScratch s;
    s.sd = sd + c0; s.ch = ch + c0; s.nodes = nodes + (c0 / 2 + 4 * (int64_t)r); s.ord = ord + c0; s.lst = lst + c0;
    return s;
}

// per read: build + filter; leaves the surviving chains in ord[] and their counts in cnt[0][r] (chains),
// cnt[1][r] (seeds)
#ifndef CHAIN_TPB
#define CHAIN_TPB 64
#endif
#ifndef CHAIN_MINB
#define CHAIN_MINB 16
#endif
// Reads whose chain list is longer than this leave mem_chain_flt's pairwise scan to chain_filter_warp_kernel: for
// reads from repeats nearly every chain stays in the kept list, the scan is quadratic, and it was 3/4 of this
// kernel's instructions at one active lane per warp (profiles/r02_ncu_chain_build_hot_lines.txt).
#ifndef CHAIN_WARP_N
#define CHAIN_WARP_N 48
#endif
// CHAIN_SORTED = 1: threads take the reads in order of decreasing work (number of seeds = SA coordinates of the read).
// A read is one thread's strictly sequential job and reads differ by orders of magnitude (a unique read has a dozen
// seeds, a repeat read hundreds): in read order a warp runs for as long as its heaviest read with one or two lanes
// active (1.65 of 32 at config-3 scale, profiles/r02_ncu_chain_build_hot_lines_after_warp_filter.txt); in work order
// the lanes of a warp finish together.  Which thread runs which read changes no result.
// MEASURED (scripts/gpu_r2_v.sh, 2 M reads at config-3 scale): 90.5 ms in work order against 48.4 ms in read order — with
// every resident warp on repeat reads at once their B-trees (tens of KB each) no longer fit L1/L2 and every node visit
// goes to DRAM, while in read order the one or two heavy lanes of a warp keep theirs on chip.  OFF; kept for the record.
#ifndef CHAIN_SORTED
#define CHAIN_SORTED 0
#endif
__global__ void chain_work_key_kernel(ChainIn in, uint32_t *__restrict__ keys, int32_t *__restrict__ idx)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= in.nseq) return;
    const int64_t n = in.coord_off[in.smem_off[r + 1]] - in.coord_off[in.smem_off[r]];
    keys[r] = (uint32_t)(n > 0xFFFF ? 0xFFFF : n);
    idx[r] = r;
}

// CHAIN_SORTED = 2: a middle way — only the heavy reads (>= CHAIN_HEAVY_MIN seeds) are placed, `hl` of them per warp in its
// first lanes, in order of decreasing work; every other lane takes a light read.  Heavy lanes then run `hl` wide instead of
// one or two, without every resident warp being all-heavy.
// MEASURED (scripts/gpu_r2_ab.sh, ms per 5 M reads): read order 117, hl = 2: 115, 4: 125, 8: 138, 16: 161 — the more heavy
// reads share a warp the slower: their B-tree walks take different branches, so the lanes do not run in lockstep anyway,
// and what is lost is the parallelism ACROSS warps.  OFF as well.
#ifndef CHAIN_HEAVY_MIN
#define CHAIN_HEAVY_MIN 64
#endif
__global__ void chain_heavy_count_kernel(const uint32_t *__restrict__ keys_sorted_desc, int n, int32_t *__restrict__ n_heavy)
{
    // first index whose key is below the threshold (keys are sorted in decreasing order)
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        int lo = 0, hi = n;
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (keys_sorted_desc[mid] >= CHAIN_HEAVY_MIN) lo = mid + 1; else hi = mid; }
        *n_heavy = lo;
    }
}
__global__ void chain_mix_perm_kernel(const int32_t *__restrict__ sorted, int n, const int32_t *__restrict__ n_heavy, int hl,
                                      int32_t *__restrict__ perm)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    int nh = *n_heavy;
    if ((int64_t)nh > (int64_t)(n >> 5) * hl) nh = (n >> 5) * hl;          // more heavy reads than heavy slots: the rest count as light
    const int warp = t >> 5, lane = t & 31;
    const int64_t before = (int64_t)warp * hl + (lane < hl ? lane : hl);    // heavy slots before slot t
    const int hb = before < nh ? (int)before : nh;
    const bool heavy_slot = lane < hl && before < nh;
    perm[t] = heavy_slot ? sorted[before] : sorted[nh + (t - hb)];
}

__global__ void __launch_bounds__(CHAIN_TPB, CHAIN_MINB) chain_build_kernel(ChainOpt o, RefView rv, ChainIn in, bwagpu_seed *sd, WChain *ch, BtNode *nodes,
                                   int32_t *ord, int32_t *lst, int64_t *cnt_chains, int64_t *cnt_seeds,
                                   float *frac_rep, int32_t *heavy_list, int32_t *heavy_cnt, const int32_t *__restrict__ perm)
{
    const int t_ = blockIdx.x * blockDim.x + threadIdx.x;
    if (t_ >= in.nseq) return;
    const int r = perm ? perm[t_] : t_;
    cnt_chains[r] = 0; cnt_seeds[r] = 0;
    const int64_t s0 = in.smem_off[r], s1 = in.smem_off[r + 1];
    if (s0 == s1 || in.len[r] < o.min_seed_len) return;                       // :821-828
    if (!in.per_read) {                                                       // `pos < num_smem - 1` (:819): a batch with <= 1 SMEM yields nothing
        int64_t nb = in.n_smem;
        if (in.batch > 0) {
            const int b0 = (r / in.batch) * in.batch, b1 = min(b0 + in.batch, in.nseq);
            nb = in.smem_off[b1] - in.smem_off[b0];
        }
        if (nb <= 1) return;
    }
    const int64_t c0 = in.coord_off[s0];
    ChainBuilder b;
    b.init(chain_scratch(sd, ch, nodes, ord, lst, c0, r));
    int bb = 0, e = 0, l_rep = 0;                                            // frac_rep (:838-848)
    for (int64_t i = s0; i < s1; ++i) {
        const int sb = (int)in.smems[i].m, se = (int)in.smems[i].n + 1;
        if (in.smems[i].s <= o.max_occ) continue;
        if (sb > e) { l_rep += e - bb; bb = sb; e = se; }
        else e = e > se ? e : se;
    }
    l_rep += e - bb;
    frac_rep[r] = (float)l_rep / in.len[r];
    for (int64_t i = s0; i < s1; ++i) {
        const bwagpu_smem p = in.smems[i];
        bwagpu_seed sdv;
        sdv.qbeg = (int32_t)p.m; sdv.len = sdv.score = (int32_t)(p.n + 1 - p.m); sdv.pad_ = -1;
        for (int64_t c = in.coord_off[i]; c < in.coord_off[i + 1]; ++c) {
            sdv.rbeg = in.coords[c];
            b.add_seed(o, rv, sdv);
        }
    }
    b.traverse();
    const int n = b.filter_prepare(o);
    if (n == 0) return;
    if (n > CHAIN_WARP_N) {                       // deferred: the warp kernel finishes this read
        heavy_list[atomicAdd(heavy_cnt, 1)] = r;
        cnt_chains[r] = n;
        return;
    }
    const int m = b.filter_finish(o, n, b.filter_scan(o, n));
    int64_t ns = 0;
    for (int i = 0; i < m; ++i) ns += b.s.ch[b.s.ord[i]].n;
    cnt_chains[r] = m; cnt_seeds[r] = ns;
}

// One warp per deferred read: the scan of chain i against the kept list runs 32 entries at a time.  Lane l always
// owns entries l, l + 32, ..., so its own flag updates are in program order; the entry appended by lane 0 is made
// visible by __syncwarp().  Exactly the sequential semantics: entries up to and including the first one that drops
// the chain are visited, each overlapping one gets `first` set once, large_ovlp = any overlap among them.
__global__ void __launch_bounds__(128) chain_filter_warp_kernel(ChainOpt o, ChainIn in, bwagpu_seed *sd, WChain *ch, BtNode *nodes,
                                                                int32_t *ord, int32_t *lst, int64_t *cnt_chains, int64_t *cnt_seeds,
                                                                const int32_t *__restrict__ heavy_list, int n_heavy)
{
    const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (wid >= n_heavy) return;
    const int r = heavy_list[wid];
    const int n = (int)cnt_chains[r];
    ChainBuilder b;
    b.s = chain_scratch(sd, ch, nodes, ord, lst, in.coord_off[in.smem_off[r]], r);
    int32_t *a = b.s.ord;
    ChainBuilder::Kept *lv = b.kept_list();
    const int msl2 = o.min_seed_len << 1;
    int nl = 1;
    for (int i = 1; i < n; ++i) {
        const WChain &ci = b.s.ch[a[i]];
        const int bi = b.chn_beg(ci), ei = b.chn_end(ci), wi = ci.w, alt_i = ci.is_alt;
        bool large = false, broke = false;
        for (int k0 = 0; k0 < nl && !broke; k0 += 32) {
            const int k = k0 + lane;
            ChainBuilder::Kept kv; kv.beg = kv.end = kv.w = kv.flags = 0;
            int t = 0;
            if (k < nl) { kv = lv[k]; t = b.flt_test(o, kv, bi, ei, wi, alt_i, msl2); }
            const unsigned mb = __ballot_sync(0xffffffffu, t & 2);
            const int stop = mb ? __ffs(mb) - 1 : 31;                 // last lane the sequential scan would visit
            const bool mine = t && lane <= stop;
            if (mine && !(kv.flags & 2)) { b.s.ch[a[b.s.lst[k]]].first = i; lv[k].flags = kv.flags | 2; }
            large |= __any_sync(0xffffffffu, mine);
            broke = mb != 0;
        }
        if (!broke) {
            if (lane == 0) {
                lv[nl].beg = bi; lv[nl].end = ei; lv[nl].w = wi; lv[nl].flags = alt_i ? 1 : 0;
                b.s.lst[nl] = i; b.s.ch[a[i]].kept = large ? 2 : 3;
            }
            ++nl;
        }
        __syncwarp();
    }
    if (lane == 0) {
        const int m = b.filter_finish(o, n, nl);
        int64_t ns = 0;
        for (int i = 0; i < m; ++i) ns += b.s.ch[a[i]].n;
        cnt_chains[r] = m; cnt_seeds[r] = ns;
    }
}

__global__ void chain_emit_kernel(ChainIn in, const bwagpu_seed *sd, const WChain *ch, const int32_t *ord,
                                  const int64_t *chain_offs, const int64_t *seed_offs, const float *frac_rep,
                                  bwagpu_chain *out_chains, bwagpu_seed *out_seeds)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= in.nseq) return;
    const int m = (int)(chain_offs[r + 1] - chain_offs[r]);
    if (m == 0) return;
    const int64_t c0 = in.coord_off[in.smem_off[r]];
    int64_t so = seed_offs[r];
    for (int i = 0; i < m; ++i) {
        const WChain &c = ch[c0 + ord[c0 + i]];
        bwagpu_chain oc;
        oc.seqid = r; oc.rid = c.rid; oc.n = c.n; oc.w = c.w; oc.kept = c.kept; oc.first = c.first; oc.is_alt = c.is_alt;
        oc.frac_rep = frac_rep[r]; oc.pos = c.pos; oc.seed_off = so;
        out_chains[chain_offs[r] + i] = oc;
        for (int si = c.head; si >= 0; si = sd[c0 + si].pad_) {
            bwagpu_seed s = sd[c0 + si];
            s.pad_ = 0;
            out_seeds[so++] = s;
        }
    }
}

// per-SMEM coordinate counts -> offsets when the caller only has the flat coordinate list
__global__ void chain_rows_kernel(const bwagpu_smem *__restrict__ sm, int64_t n, int max_occ, int64_t *__restrict__ rows)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t s = sm[i].s, step = s > max_occ ? s / max_occ : 1;
    int64_t c = (s + step - 1) / step;
    rows[i] = c > max_occ ? max_occ : (c < 0 ? 0 : c);
}

// reads' SMEM ranges from the rid column of the sorted SMEM list
__global__ void chain_smem_off_kernel(const bwagpu_smem *__restrict__ sm, int64_t n, int32_t nseq, int64_t *__restrict__ off)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    // off[r] = first SMEM with rid >= r: every r in (rid[i-1], rid[i]] starts at i
    const int64_t lo = i == 0 ? 0 : (int64_t)sm[i - 1].rid + 1;
    const int64_t hi = i == n ? nseq : (int64_t)sm[i].rid;
    for (int64_t r = lo; r <= hi && r <= nseq; ++r) off[r] = i;
}

static int chain_scan(bwagpu_ctx *ctx, const int64_t *in, int64_t *out, int64_t n)
{
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum((void *)nullptr, bytes, in, out, (int)n, ctx->stream);
    void *tmp;
    int rc = bwagpu_scratch(ctx, SCR_CUB, (int64_t)bytes + 256, &tmp);
    if (rc) return rc;
    BWAGPU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, bytes, in, out, (int)n, ctx->stream));
    ctx->launches += 2;
    return BWAGPU_OK;
}

static int chain_check(bwagpu_ctx *ctx, const bwagpu_chain_opt *o)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!o) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain_seeds: options are NULL");
    if (!ctx->d_ref_string) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain_seeds: no reference loaded (bwagpu_index_load_ref)");
    if (o->max_occ < 1 || o->min_seed_len < 1) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain_seeds: max_occ and min_seed_len must be positive");
    return BWAGPU_OK;
}

static int chain_device(bwagpu_ctx *ctx, const bwagpu_chain_opt *opt, const bwagpu_smem *d_smems,
                        int64_t n_smem, const int64_t *d_smem_off, const int64_t *d_coords,
                        int64_t n_coords, const int64_t *d_coord_off, const int32_t *d_len,
                        int32_t nseq, int32_t max_len, bwagpu_chain_result *res, int per_read)
{
    int rc = chain_check(ctx, opt);
    if (rc) return rc;
    if (!res || nseq < 0 || n_smem < 0 || n_coords < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain_seeds: bad argument");
    // mem_flt_chained_seeds would re-score seeds with ksw_align2 for reads this long (:509-512)
    if (max_len > 0 && !(5.5f * logf((float)max_len) > 0.05f * max_len) && opt->min_chain_weight == 0)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "chain_seeds: reads of %d bases need mem_flt_chained_seeds, which this stage does not cover", max_len);
    memset(res, 0, sizeof *res);
    if (nseq == 0 || n_smem == 0 || n_coords == 0) return BWAGPU_OK;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    ChainOpt o = { opt->w, opt->max_chain_gap, opt->max_occ, opt->min_seed_len, opt->min_chain_weight,
                   opt->max_chain_extend, opt->mask_level, opt->drop_ratio };
    RefView rv = { ctx->l_pac, ctx->n_contigs, (const int64_t *)ctx->d_contig_off, (const int32_t *)ctx->d_contig_len,
                   (const uint8_t *)ctx->d_contig_alt };
    ChainIn in = { d_smems, d_smem_off, d_coords, d_coord_off, d_len, nseq, n_smem, per_read, ctx->chain_batch_reads };
    bwagpu_seed *sd; WChain *ch; BtNode *nodes; int32_t *ord, *lst; int64_t *cnt, *offs; float *frac;
    if ((rc = bwagpu_scratch(ctx, SCR_CH_SD, (n_coords + 1) * (int64_t)sizeof(bwagpu_seed), (void **)&sd))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CH_CH, (n_coords + 1) * (int64_t)sizeof(WChain), (void **)&ch))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CH_NODES, (n_coords / 2 + 4 * (int64_t)nseq + 8) * (int64_t)sizeof(BtNode), (void **)&nodes))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CH_ORD, (n_coords + 1) * 8, (void **)&ord))) return rc;
    lst = ord + n_coords + 1;
    if ((rc = bwagpu_scratch(ctx, SCR_CH_CNT, (int64_t)(nseq + 1) * 16, (void **)&cnt))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CH_OFFS, (int64_t)(nseq + 1) * 16, (void **)&offs))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CH_FRAC, (int64_t)(nseq + 1) * 4, (void **)&frac))) return rc;
    int32_t *heavy;
    if ((rc = bwagpu_scratch(ctx, SCR_CH_HEAVY, (int64_t)(nseq + 2) * 4, (void **)&heavy))) return rc;
    BWAGPU_CUDA(ctx, cudaMemsetAsync(cnt, 0, (size_t)(nseq + 1) * 16, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(heavy, 0, 4, ctx->stream));
    int32_t *perm = nullptr;
    BWAGPU_KERNEL_BEGIN(ctx);
#if CHAIN_SORTED
    if (nseq > 64) {                               // reads in order of decreasing seed count
        uint32_t *keys;
        if ((rc = bwagpu_scratch(ctx, SCR_CH_PERM, (int64_t)nseq * 16, (void **)&keys))) return rc;
        uint32_t *keys_out = keys + nseq; int32_t *idx = (int32_t *)(keys + 2 * (int64_t)nseq); perm = idx + nseq;
        chain_work_key_kernel<<<(nseq + 255) / 256, 256, 0, ctx->stream>>>(in, keys, idx);
        size_t bytes = 0;
        cub::DeviceRadixSort::SortPairsDescending((void *)nullptr, bytes, keys, keys_out, idx, perm, nseq, 0, 16, ctx->stream);
        void *tmp;
        if ((rc = bwagpu_scratch(ctx, SCR_CUB, (int64_t)bytes + 256, &tmp))) return rc;
        BWAGPU_CUDA(ctx, cub::DeviceRadixSort::SortPairsDescending(tmp, bytes, keys, keys_out, idx, perm, nseq, 0, 16, ctx->stream));
        ctx->launches += 4;
#if CHAIN_SORTED == 2
        {
            static const int hl = getenv("BWAGPU_CHAIN_HEAVY_LANES") ? atoi(getenv("BWAGPU_CHAIN_HEAVY_LANES")) : 8;
            int32_t *n_heavy_d = (int32_t *)keys;                  // the unsorted keys are dead now
            int32_t *mixed = idx;                                   // and so are the identity indices
            chain_heavy_count_kernel<<<1, 32, 0, ctx->stream>>>(keys_out, nseq, n_heavy_d);
            chain_mix_perm_kernel<<<(nseq + 255) / 256, 256, 0, ctx->stream>>>(perm, nseq, n_heavy_d, hl < 1 ? 1 : (hl > 32 ? 32 : hl), mixed);
            perm = mixed;
            ctx->launches += 2;
        }
#endif
    }
#endif
    chain_build_kernel<<<(nseq + CHAIN_TPB - 1) / CHAIN_TPB, CHAIN_TPB, 0, ctx->stream>>>(o, rv, in, sd, ch, nodes, ord, lst, cnt, cnt + nseq + 1, frac,
                                                                                          heavy + 1, heavy, perm);
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_CHAIN);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    int32_t n_heavy = 0;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&n_heavy, heavy, 4, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    if (n_heavy > 0) {
        BWAGPU_KERNEL_BEGIN(ctx);
        chain_filter_warp_kernel<<<(unsigned)(((int64_t)n_heavy * 32 + 127) / 128), 128, 0, ctx->stream>>>(
            o, in, sd, ch, nodes, ord, lst, cnt, cnt + nseq + 1, heavy + 1, n_heavy);
        BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_CHAIN);
        BWAGPU_CUDA(ctx, cudaGetLastError());
    }
    if ((rc = chain_scan(ctx, cnt, offs, nseq + 1))) return rc;
    if ((rc = chain_scan(ctx, cnt + nseq + 1, offs + nseq + 1, nseq + 1))) return rc;
    int64_t tot[2];
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&tot[0], offs + nseq, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&tot[1], offs + 2 * nseq + 1, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    bwagpu_chain *out_c; bwagpu_seed *out_s;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_CHAINS, (tot[0] + 1) * (int64_t)sizeof(bwagpu_chain), (void **)&out_c))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_SEEDS, (tot[1] + 1) * (int64_t)sizeof(bwagpu_seed), (void **)&out_s))) return rc;
    BWAGPU_KERNEL_BEGIN(ctx);
    chain_emit_kernel<<<(nseq + 63) / 64, 64, 0, ctx->stream>>>(in, sd, ch, ord, offs, offs + nseq + 1, frac, out_c, out_s);
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_CHAIN);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    res->d_chains = out_c; res->n_chains = tot[0]; res->d_seeds = out_s; res->n_seeds = tot[1];
    return BWAGPU_OK;
}

extern "C" int bwagpu_chain_seeds_device(bwagpu_ctx *ctx, const bwagpu_chain_opt *opt, const bwagpu_smem *d_smems,
                                         int64_t n_smem, const int64_t *d_smem_off, const int64_t *d_coords,
                                         int64_t n_coords, const int64_t *d_coord_off, const int32_t *d_len,
                                         int32_t nseq, int32_t max_len, bwagpu_chain_result *res)
{
    return chain_device(ctx, opt, d_smems, n_smem, d_smem_off, d_coords, n_coords, d_coord_off, d_len, nseq, max_len, res, 0);
}

extern "C" int bwagpu_chain_mems_device(bwagpu_ctx *ctx, const bwagpu_chain_opt *opt, const bwagpu_smem *d_smems,
                                        int64_t n_smem, const int64_t *d_smem_off, const int64_t *d_coords,
                                        int64_t n_coords, const int64_t *d_coord_off, const int32_t *d_len,
                                        int32_t nseq, int32_t max_len, bwagpu_chain_result *res)
{
    return chain_device(ctx, opt, d_smems, n_smem, d_smem_off, d_coords, n_coords, d_coord_off, d_len, nseq, max_len, res, 1);
}

extern "C" int bwagpu_chain_seeds(bwagpu_ctx *ctx, const bwagpu_chain_opt *opt, const bwagpu_smem *smems, int64_t n_smem,
                                  const int64_t *coords, int64_t n_coords, const int32_t *len, int32_t nseq,
                                  bwagpu_chain *chains, int64_t chains_cap, bwagpu_seed *seeds, int64_t seeds_cap,
                                  int64_t *n_chains, int64_t *n_seeds)
{
    int rc = chain_check(ctx, opt);
    if (rc) return rc;
    if (!n_chains || !n_seeds || nseq < 0 || n_smem < 0 || n_coords < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain_seeds: bad argument");
    *n_chains = *n_seeds = 0;
    if (nseq == 0 || n_smem == 0) return BWAGPU_OK;
    if (!smems || !len || (n_coords > 0 && !coords)) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain_seeds: NULL buffer");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    int max_len = 0;
    for (int i = 0; i < nseq; ++i) if (len[i] > max_len) max_len = len[i];
    for (int64_t i = 0; i < n_smem; ++i)
        if ((int64_t)smems[i].rid >= nseq || (i > 0 && smems[i].rid < smems[i - 1].rid))
            BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain_seeds: SMEM %lld breaks the (rid, m, n) order / rid range", (long long)i);
    bwagpu_smem *d_sm; int64_t *d_coords, *d_rows, *d_coff, *d_soff; int32_t *d_len;
    if ((rc = bwagpu_scratch(ctx, SCR_SA_IN, n_smem * (int64_t)sizeof(bwagpu_smem), (void **)&d_sm))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SA_COORD, (n_coords + 1) * 8, (void **)&d_coords))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SA_CNT, (n_smem + 1) * 8, (void **)&d_rows))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SA_OFF, (n_smem + 1) * 8, (void **)&d_coff))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_SCAN, (int64_t)(nseq + 1) * 8, (void **)&d_soff))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_LEN, (int64_t)nseq * 4, (void **)&d_len))) return rc;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_sm, smems, (size_t)n_smem * sizeof(bwagpu_smem), cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_coords, coords, (size_t)n_coords * 8, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_len, len, (size_t)nseq * 4, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_rows + n_smem, 0, 8, ctx->stream));
    chain_rows_kernel<<<(unsigned)((n_smem + 255) / 256), 256, 0, ctx->stream>>>(d_sm, n_smem, opt->max_occ, d_rows);
    chain_smem_off_kernel<<<(unsigned)((n_smem + 256) / 256), 256, 0, ctx->stream>>>(d_sm, n_smem, nseq, d_soff);
    ctx->launches += 2;
    if ((rc = chain_scan(ctx, d_rows, d_coff, n_smem + 1))) return rc;
    int64_t tot = 0;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&tot, d_coff + n_smem, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    if (tot != n_coords)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain_seeds: the SMEMs call for %lld SA coordinates, %lld were passed", (long long)tot, (long long)n_coords);
    bwagpu_chain_result res;
    if ((rc = bwagpu_chain_seeds_device(ctx, opt, d_sm, n_smem, d_soff, d_coords, n_coords, d_coff, d_len, nseq, max_len, &res))) return rc;
    *n_chains = res.n_chains; *n_seeds = res.n_seeds;
    if (res.n_chains > chains_cap || res.n_seeds > seeds_cap)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "chain_seeds: %lld chains / %lld seeds do not fit the output buffers", (long long)res.n_chains, (long long)res.n_seeds);
    if (res.n_chains > 0) {
        if (!chains || !seeds) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain_seeds: output buffer is NULL");
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(chains, res.d_chains, (size_t)res.n_chains * sizeof(bwagpu_chain), cudaMemcpyDeviceToHost, ctx->stream));
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(seeds, res.d_seeds, (size_t)res.n_seeds * sizeof(bwagpu_seed), cudaMemcpyDeviceToHost, ctx->stream));
    }
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}
