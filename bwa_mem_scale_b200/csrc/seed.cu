// S1 seeding seam of the C ABI (include/bwagpu.h): FM-index replica upload, SMEM collection
// (three seeding rounds + canonical order) and batched suffix-array look-up.
// Replaces mem_collect_smem (reference src/bwamem.cpp:648-786) and
// FMI_search::get_sa_entries_prefetch (reference src/FMI_search.cpp:2261-2379).
//
// Kernels:
//   seed_collect_kernel  one read per thread, persistent threads; every loop iteration performs
//                        exactly one backward extension (two 64-byte CP_OCC records, one when
//                        k and k+s share a block), so a warp's 32 reads meet at one load site.
//   seed_sort_kernel     per read: order its SMEMs by (m, n) and copy them to their final slot.
//   sa_count_kernel      rows each SMEM contributes (min(max_occ, ceil(s/step))).
//   sa_walk_kernel       one thread per output coordinate: LF walk to the next sampled row.
#include <cub/device/device_scan.cuh>
#include <stdlib.h>
#include "ctx.cuh"
#include "fmi_core.cuh"
#include <type_traits>

#define SEED_TPB 128
#ifndef SLOW_BATCH
#define SLOW_BATCH 8
#endif

// prev list of one thread.  The first SEED_PREV_K entries pushed for a pivot live in shared memory
// (4 words each: k, l, s low halves + a word of high bits and the span end n; word w of entry i of
// thread t at sm[(i*4 + w) * SEED_TPB + t], bank = t: conflict-free); deeper entries spill to the
// interleaved global scratch (word w of entry e of thread t at buf[(e*3 + w) * stride + t]).
// Entries are addressed from the END of a len-sized array downwards (see SeedMachine), so slot
// i = len - 1 - e is the push order.
#ifndef SEED_PREV_K
#define SEED_PREV_K 0
#endif
#define SEED_SMEM_BASES 256      // reads up to this long are staged in shared memory (4-bit codes)

// SEED_PREV16 = 1: an entry is ONE 16-byte word pair in a region of the scratch that belongs to the thread alone
// (k, l, s in 38 bits each — the text is shorter than 2^38 — and the span end n in 14): one 128-bit load or store per entry,
// and the 32-byte sectors it touches hold this thread's entries only.  The interleaved 3 x 8-byte layout put word w of
// entry e of four neighbouring threads into one sector, which lanes at different depths used an eighth of: the prev-list
// load was the kernel's largest single stall (17.7 % of the samples) and its partial-sector write-backs a good part of the
// 57 GB written / 285 GB read per 2.5 M reads against 82 GB of records asked for (profiles/r02_ncu_seed_collect_v2_lines.txt).
// Reads of 16384 bases or more keep the interleaved layout (n does not fit).
// MEASURED (scripts/gpu_r2_ac.sh, ms per 5 M reads at config-3 scale): 259.0 with it against 245.7 with the interleaved
// layout — slower: a thread-private region means a 128-byte line serves one thread, the interleaved lines serve sixteen,
// and the hot entries of a warp no longer share L1 lines.  OFF; kept for the record.
#ifndef SEED_PREV16
#define SEED_PREV16 0
#endif
struct DevPrev {
    uint32_t *sm;
    int64_t *buf;
    ulonglong2 *own;             // this thread's 16-byte entries (SEED_PREV16), NULL: interleaved layout
    int stride, top;
    __device__ __forceinline__ void reset(int len) { top = len - 1; }
    __device__ __forceinline__ void put(int e, const Biv &v, int n)
    {
        if (SEED_PREV16 && own) {
            const uint64_t s = (uint64_t)v.s;
            ulonglong2 w;
            w.x = (uint64_t)v.k | (s << 38);
            w.y = (uint64_t)v.l | ((s >> 26) << 38) | ((uint64_t)n << 50);
            own[e] = w;
            return;
        }
        const int i = top - e;
        if (SEED_PREV_K > 0 && i < SEED_PREV_K) {
            uint32_t *p = sm + i * 4 * SEED_TPB;
            p[0] = (uint32_t)v.k; p[SEED_TPB] = (uint32_t)v.l; p[2 * SEED_TPB] = (uint32_t)v.s;
            p[3 * SEED_TPB] = (uint32_t)(v.k >> 32) | ((uint32_t)(v.l >> 32) << 6) | ((uint32_t)(v.s >> 32) << 12) |
                              ((uint32_t)n << 18);
        } else {
            int64_t *p = buf + (int64_t)(e * 3) * stride;
            p[0] = v.k | ((int64_t)n << 40); p[stride] = v.l; p[2 * stride] = v.s;
        }
    }
    __device__ __forceinline__ void get(int e, Biv &v, int &n) const
    {
        if (SEED_PREV16 && own) {
            const ulonglong2 w = own[e];
            const uint64_t m38 = (1ULL << 38) - 1;
            v.k = (int64_t)(w.x & m38); v.l = (int64_t)(w.y & m38);
            v.s = (int64_t)((w.x >> 38) | (((w.y >> 38) & 0xFFFULL) << 26));
            n = (int)(w.y >> 50);
            return;
        }
        const int i = top - e;
        if (SEED_PREV_K > 0 && i < SEED_PREV_K) {
            const uint32_t *p = sm + i * 4 * SEED_TPB;
            const uint32_t hi = p[3 * SEED_TPB];
            v.k = (int64_t)(((uint64_t)(hi & 63u) << 32) | p[0]);
            v.l = (int64_t)(((uint64_t)((hi >> 6) & 63u) << 32) | p[SEED_TPB]);
            v.s = (int64_t)(((uint64_t)((hi >> 12) & 63u) << 32) | p[2 * SEED_TPB]);
            n = (int)(hi >> 18);
        } else {
            const int64_t *p = buf + (int64_t)(e * 3) * stride;
            const int64_t w0 = p[0];
            v.k = w0 & ((1LL << 40) - 1); n = (int)(w0 >> 40); v.l = p[stride]; v.s = p[2 * stride];
        }
    }
};

// a read staged in shared memory as 4-bit codes, 8 per word (word i of thread t at w[i*SEED_TPB + t])
struct SmemRead {
    const uint32_t *w;
    __device__ __forceinline__ int operator[](int j) const { return (int)((w[(j >> 3) * SEED_TPB] >> ((j & 7) * 4)) & 15u); }
};

__device__ __forceinline__ void seed_stage_read(uint32_t *w, const uint8_t *__restrict__ src, int len)
{
    for (int i = 0; i * 8 < len; ++i) {
        uint32_t v = 0;
#pragma unroll
        for (int b = 0; b < 8; ++b)
            if (i * 8 + b < len) { const uint32_t c = src[i * 8 + b]; v |= (c > 4u ? 4u : c) << (4 * b); }
        w[i * SEED_TPB] = v;
    }
}

struct SeedStats { unsigned long long n_ext, n_rec, n_lf, max_cnt; unsigned int n_ovf, next; };

// SEED_DYN = 1: a lane that needs a new read takes the next unclaimed one from a counter in global memory instead of
// striding through the batch.  Reads differ widely in work (a repeat read costs many times a unique one, an EMF-answered
// read nothing), so with a fixed stride the lanes of a warp run out of reads at very different times and the extension
// site sees half-empty warps (15.5 of 32 lanes at config-3 scale, profiles/r02_ncu_seed_collect_hot_lines.txt); with the
// counter every lane stays busy until the batch is exhausted.  Which lane runs which read does not change any result.
#ifndef SEED_DYN
#define SEED_DYN 1
#endif

// ---- warp-cooperative backwardExt ------------------------------------------------------------------------------
// SEED_COOP = 1: the reads stay one per thread (so every issued instruction of the state machine still serves all the
// lanes that are in that state), but the CP_OCC records are fetched by the WARP: the pending requests of the ready lanes
// are served eight at a time, four lanes per request — lane b of a group loads the 8-byte count and the 8-byte one-hot
// word of base b from the record at sp and from the record at ep (a group's four lanes read one 32-byte sector per
// load), popcounts, and shuffles assemble (k, l, s) for the requesting lane.  On the 6.2 GB table of the human-scale
// index this access shape moves 2.8x the records per second of one thread fetching whole records with 256-bit loads
// (profiles/r02_dpx_peak_and_bext_probe.json: 2398 vs 863 GB/s), and the two records of a request no longer sit in 32
// registers of every thread.
// MEASURED in the full kernel (profiles/r02_seed_ab_config3.txt, 2 M reads on the 3.1 Gbp index): 211 ms with two rounds
// in flight (118 registers), 237 / 217 / 252 ms with 1 / 3 / 4 rounds, against 195 ms for the per-thread fetch — the
// kernel is not bound by the fetch shape but by the length of each warp iteration (about 500 warp instructions at 16
// cycles each, 4.7 warps per scheduler), so the cooperative fetch stays OFF by default and is kept for the record.
#ifndef SEED_COOP
#define SEED_COOP 0
#endif
__device__ __forceinline__ int64_t seed_shfl64(int64_t v, int src)
{
    const int lo = __shfl_sync(0xffffffffu, (int)(uint32_t)v, src);
    const int hi = __shfl_sync(0xffffffffu, (int)(uint32_t)((uint64_t)v >> 32), src);
    return (int64_t)(((uint64_t)(uint32_t)hi << 32) | (uint32_t)lo);
}

// All 32 lanes call this.  `ready` lanes pass their request (q, qa) and receive its result; returns that result.
// Two phases per pass so that the loads of SEED_COOP_ROUNDS rounds of eight requests are in flight together.
#ifndef SEED_COOP_ROUNDS
#define SEED_COOP_ROUNDS 2
#endif
__device__ __forceinline__ Biv seed_bext_coop(const FmView &f, unsigned m_ready, bool ready, const Biv &q, int qa)
{
    const int lane = threadIdx.x & 31, g = lane >> 2, b = lane & 3;
    const int my_rank = __popc(m_ready & ((1u << lane) - 1u));      // position of this lane among the ready lanes
    const int n_req = __popc(m_ready);
    Biv res; res.k = res.l = res.s = 0;
    const int gb = lane & ~3;
  for (int pass = 0; pass < n_req; pass += 8 * SEED_COOP_ROUNDS) {
    int64_t c0[SEED_COOP_ROUNDS], c1[SEED_COOP_ROUNDS]; uint64_t o0[SEED_COOP_ROUNDS], o1[SEED_COOP_ROUNDS];
    int src[SEED_COOP_ROUNDS], kx[SEED_COOP_ROUNDS];                 // kx: k & 63 | (ep & 63) << 8 | sentinel << 16
#pragma unroll
    for (int i = 0; i < SEED_COOP_ROUNDS; ++i) {
        const int base = pass + i * 8;
        if (base < n_req) {                                          // warp-uniform
            // the request this group serves in round i: the (base + g)-th ready lane
            const bool have = base + g < n_req;
            src[i] = have ? (int)__fns(m_ready, 0, base + g + 1) : 0;
            const int64_t k = seed_shfl64(q.k, src[i]), ep = k + seed_shfl64(q.s, src[i]);
            c0[i] = c1[i] = 0; o0[i] = o1[i] = 0;
            if (have) {
                const int64_t *r0 = f.cp_occ + (k >> 6) * 8, *r1 = f.cp_occ + (ep >> 6) * 8;
                c0[i] = __ldg(r0 + FM_W_CNT(b)); c1[i] = __ldg(r1 + FM_W_CNT(b));
                o0[i] = (uint64_t)__ldg(r0 + FM_W_BITS(b)); o1[i] = (uint64_t)__ldg(r1 + FM_W_BITS(b));
            }
            kx[i] = (int)(k & 63) | ((int)(ep & 63) << 8) | ((k <= f.sentinel_index && ep > f.sentinel_index) ? 1 << 16 : 0);
        }
    }
#pragma unroll
    for (int i = 0; i < SEED_COOP_ROUNDS; ++i) {
        const int base = pass + i * 8;
        if (base < n_req) {
            const int64_t osp = c0[i] + __popcll(o0[i] & fm_prefix_mask(kx[i] & 63));
            const int64_t sz = c1[i] + __popcll(o1[i] & fm_prefix_mask((kx[i] >> 8) & 63)) - osp;
            const int64_t nk = fm_count(f, b) + osp;
            // l[b] = l + sentinel + sum of s[c] over c > b
            const int64_t s1 = seed_shfl64(sz, gb | ((b + 1) & 3)), s2 = seed_shfl64(sz, gb | ((b + 2) & 3)),
                          s3 = seed_shfl64(sz, gb | ((b + 3) & 3));
            const int64_t nl = seed_shfl64(q.l, src[i]) + (kx[i] >> 16) + (b < 3 ? s1 : 0) + (b < 2 ? s2 : 0) + (b < 1 ? s3 : 0);
            // hand the answer for base qa back to the requesting lane: its group in this round is my_rank - base
            const int from = (((my_rank - base) & 7) << 2) | (qa & 3);
            const int64_t rk = seed_shfl64(nk, from), rl = seed_shfl64(nl, from), rs = seed_shfl64(sz, from);
            if (ready && my_rank >= base && my_rank < base + 8) { res.k = rk; res.l = rl; res.s = rs; }
        }
    }
  }
    return res;
}

// SEED_STAGE_COOP = 1 (needs SEED_DYN): new reads are staged into shared memory by the whole warp, see the slow pass.
#ifndef SEED_STAGE_COOP
#define SEED_STAGE_COOP SEED_DYN
#endif
#ifndef SEED_MIN_BLOCKS
#define SEED_MIN_BLOCKS 5      // 5 blocks of 128 per SM = 96 registers; without the bound the paired step takes 98 and drops to 4
#endif
// STAGED: every read of the batch is at most SEED_SMEM_BASES long and is staged in shared memory.
template <bool STAGED>
__global__ void __launch_bounds__(SEED_TPB, SEED_MIN_BLOCKS)
seed_collect_kernel(FmView f_in, SeedOpt o, const uint8_t *__restrict__ reads,
                    const int64_t *__restrict__ off, const int32_t *__restrict__ len,
                    const uint8_t *__restrict__ skip, int nseq, bwagpu_smem *__restrict__ tmp,
                    int cap, int32_t *__restrict__ cnt, int64_t *__restrict__ prev_buf,
                    SeedStats *__restrict__ stats, const int32_t *__restrict__ list, int32_t *__restrict__ ovf_list,
                    int32_t *__restrict__ slab_of, int prev_cap16)
{
    // list == NULL: first pass over all nseq reads, read i writes slab i of `tmp`; a read that finds more than `cap`
    // SMEMs is appended to ovf_list.  list != NULL: second pass over the nseq reads list[0 .. nseq) with the wider private
    // slabs of `tmp` (slab = position in the list, recorded in slab_of).
    const int64_t T = (int64_t)gridDim.x * blockDim.x;
    const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    extern __shared__ uint32_t seed_sm[];
    __shared__ int64_t s_count[5];                     // count[] where fm_count can index it
    if (threadIdx.x < 5) s_count[threadIdx.x] = fm_count(f_in, (int)threadIdx.x);
    __syncthreads();
    FmView f = f_in;
    f.count_tab = s_count;
    DevPrev pv; pv.buf = prev_buf + t0; pv.stride = (int)T; pv.top = 0;
    pv.own = prev_cap16 > 0 ? reinterpret_cast<ulonglong2 *>(prev_buf) + t0 * (int64_t)prev_cap16 : nullptr;
    pv.sm = seed_sm + threadIdx.x;
    uint32_t *rd_sm = seed_sm + SEED_PREV_K * 4 * SEED_TPB + threadIdx.x;
    using Read = typename std::conditional<STAGED, SmemRead, const uint8_t *>::type;
    SeedMachine<DevPrev, Read> m;
    [[maybe_unused]] int64_t next = t0;
    bool active = false, ready = false, finished = false;
    unsigned long long n_ext = 0, n_rec = 0;
    unsigned int max_cnt = 0;
    // Warp-level scheduling: lanes whose next extension request is ready run the extension
    // together; lanes that need the rarer bookkeeping (advance(): phase / pivot / round changes,
    // fetching the next read) wait until SLOW_BATCH of them have gathered, or nobody is ready,
    // and then run it together.  Scheduling only: per-read results do not depend on it.
    for (;;) {
        const unsigned m_ready = __ballot_sync(0xffffffffu, ready);
        const unsigned m_slow = __ballot_sync(0xffffffffu, !ready && !finished);
        if (m_slow && (__popc(m_slow) >= SLOW_BATCH || !m_ready)) {
#if SEED_STAGE_COOP
            // One slow pass = (1) lanes with a read in hand run its bookkeeping up to the next request or to its end,
            // (2) lanes without a read claim the next one, (3) the WARP stages the claimed reads into shared memory, every
            // lane packing 8 bases of the same read per step (coalesced, ~20 warp instructions per read instead of ~1200 on
            // one lane), (4) the new reads run up to their first request.
            const bool mine = !ready && !finished;
            if (mine && active) {
                if (m.advance(f, o)) ready = true;
                else {
                    cnt[m.rid] = m.n_out;
                    max_cnt = max(max_cnt, (unsigned)m.n_out);
                    if (!list && m.n_out > cap) ovf_list[atomicAdd(&stats->n_ovf, 1u)] = m.rid;
                    active = false;
                }
            }
            int it = -1, i = -1;
            if (mine && !ready && !active) {
                for (;;) {
                    const int64_t cand = (int64_t)atomicAdd(&stats->next, 1u);
                    if (cand >= nseq) { finished = true; it = -1; break; }
                    it = (int)cand;
                    i = list ? list[it] : it;
                    if (list) slab_of[i] = it;
                    if (skip && skip[i]) { cnt[i] = 0; continue; }
                    break;
                }
            }
            const bool claimed = it >= 0 && !finished;
            if constexpr (STAGED) {
                unsigned need = __ballot_sync(0xffffffffu, claimed);
                const int lane = threadIdx.x & 31;
                while (need) {
                    const int src = __ffs(need) - 1;
                    need &= need - 1;
                    const int ri = __shfl_sync(0xffffffffu, i, src);
                    const int L = len[ri];
                    const uint8_t *base = reads + off[ri];
                    uint32_t *dst = seed_sm + SEED_PREV_K * 4 * SEED_TPB + (threadIdx.x & ~31) + src;
                    for (int w = lane; w * 8 < L; w += 32) {
                        uint32_t v = 0;
#pragma unroll
                        for (int b = 0; b < 8; ++b)
                            if (w * 8 + b < L) { const uint32_t c = base[w * 8 + b]; v |= (c > 4u ? 4u : c) << (4 * b); }
                        dst[w * SEED_TPB] = v;
                    }
                }
                __syncwarp();
            }
            if (claimed) {
                if constexpr (STAGED) {
                    SmemRead r; r.w = rd_sm;
                    m.init(r, len[i], i, pv, tmp + (int64_t)it * cap, cap);
                } else {
                    m.init(reads + off[i], len[i], i, pv, tmp + (int64_t)it * cap, cap);
                }
                m.pf_cp = f.cp_occ;
                active = true;
                if (m.advance(f, o)) ready = true;
                else {                                   // a read with nothing to search (empty, all N)
                    cnt[m.rid] = m.n_out;
                    max_cnt = max(max_cnt, (unsigned)m.n_out);
                    if (!list && m.n_out > cap) ovf_list[atomicAdd(&stats->n_ovf, 1u)] = m.rid;
                    active = false;
                }
            }
#else
            if (!ready && !finished) {
                for (;;) {
                    if (!active) {
                        bool got = false;
                        for (;;) {
#if SEED_DYN
                            const int64_t cand = (int64_t)atomicAdd(&stats->next, 1u);
                            if (cand >= nseq) break;
                            const int it = (int)cand;
#else
                            if (next >= nseq) break;
                            const int it = (int)next; next += T;
#endif
                            const int i = list ? list[it] : it;
                            if (list) slab_of[i] = it;
                            if (skip && skip[i]) { cnt[i] = 0; continue; }
                            if constexpr (STAGED) {
                                seed_stage_read(rd_sm, reads + off[i], len[i]);
                                SmemRead r; r.w = rd_sm;
                                m.init(r, len[i], i, pv, tmp + (int64_t)it * cap, cap);
                            } else {
                                m.init(reads + off[i], len[i], i, pv, tmp + (int64_t)it * cap, cap);
                            }
                            m.pf_cp = f.cp_occ;
                            got = true;
                            break;
                        }
                        if (!got) { finished = true; break; }
                        active = true;
                    }
                    if (m.advance(f, o)) { ready = true; break; }
                    cnt[m.rid] = m.n_out;
                    max_cnt = max(max_cnt, (unsigned)m.n_out);
                    if (!list && m.n_out > cap) ovf_list[atomicAdd(&stats->n_ovf, 1u)] = m.rid;
                    active = false;
                }
            }
#endif
            continue;
        }
        if (!m_ready) break;
#if SEED_COOP
        const Biv r = seed_bext_coop(f, m_ready, ready, m.req, m.req_a);
        if (ready) {
            n_rec += ((m.req.k >> 6) == ((m.req.k + m.req.s) >> 6)) ? 1 : 2;
            ++n_ext;
            ready = m.step_fast(r, o);
        }
#else
        if (ready) {
            n_rec += ((m.req.k >> 6) == ((m.req.k + m.req.s) >> 6)) ? 1 : 2;
            ++n_ext;
#if SEED_PAIR2
            if (m.has2) { n_rec += ((m.req2.k >> 6) == ((m.req2.k + m.req2.s) >> 6)) ? 1 : 2; ++n_ext; }
            Biv r2;
            const Biv r = fm_bext2(f, m.req, m.req2, m.has2, m.req_a, r2);
            ready = m.step_fast(r, o, &r2);
#else
            const Biv r = fm_bext(f, m.req, m.req_a);
            ready = m.step_fast(r, o);
#endif
        }
#endif
    }
    // per-warp then global statistics
    for (int d = 16; d; d >>= 1) {
        n_ext += __shfl_down_sync(0xffffffffu, n_ext, d);
        n_rec += __shfl_down_sync(0xffffffffu, n_rec, d);
        max_cnt = max(max_cnt, __shfl_down_sync(0xffffffffu, max_cnt, d));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&stats->n_ext, n_ext);
        atomicAdd(&stats->n_rec, n_rec);
        atomicMax(&stats->max_cnt, (unsigned long long)max_cnt);
    }
}

// Per read: rank its SMEMs by (m, n) and write them to out[offs[i] ..].  Equal keys are
// identical entries (same substring => same interval), so any order among them is the reference's.
__global__ void seed_sort_kernel(const bwagpu_smem *__restrict__ tmp, int cap, const bwagpu_smem *__restrict__ tmp2, int cap2,
                                 const int32_t *__restrict__ slab_of, const int32_t *__restrict__ cnt,
                                 const int64_t *__restrict__ offs, int nseq, bwagpu_smem *__restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nseq) return;
    const int n = cnt[i];
    const int so = slab_of[i];                     // >= 0: the read was re-run with a private slab of cap2 entries
    const bwagpu_smem *src = so >= 0 ? tmp2 + (int64_t)so * cap2 : tmp + (int64_t)i * cap;
    bwagpu_smem *dst = out + offs[i];
    for (int e = 0; e < n; ++e) {
        const uint64_t key = ((uint64_t)src[e].m << 32) | src[e].n;
        int rank = 0;
        for (int g = 0; g < n; ++g) {
            const uint64_t kg = ((uint64_t)src[g].m << 32) | src[g].n;
            rank += (kg < key) || (kg == key && g < e);
        }
        dst[rank] = src[e];
    }
}

__global__ void sa_count_kernel(const bwagpu_smem *__restrict__ sm, int64_t n, int max_occ,
                                int64_t *__restrict__ rows)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) rows[i] = fm_sa_rows(sm[i].s, max_occ, nullptr);
}

__global__ void sa_walk_kernel(FmView f, const bwagpu_smem *__restrict__ sm, int64_t n_smem,
                               const int64_t *__restrict__ offs, int64_t n_coords, int max_occ,
                               int64_t *__restrict__ coords, SeedStats *__restrict__ stats)
{
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long lf = 0;
    if (c < n_coords) {
        // the SMEM this coordinate belongs to: last i with offs[i] <= c
        int64_t lo = 0, hi = n_smem - 1;
        while (lo < hi) {
            const int64_t mid = (lo + hi + 1) >> 1;
            if (offs[mid] <= c) lo = mid; else hi = mid - 1;
        }
        const int64_t s = sm[lo].s, k = sm[lo].k;
        const int64_t step = s > max_occ ? s / max_occ : 1;
        int steps = 0;
        coords[c] = fm_sa_walk(f, k + (c - offs[lo]) * step, &steps);
        lf = (unsigned long long)steps;
    }
    for (int d = 16; d; d >>= 1) lf += __shfl_down_sync(0xffffffffu, lf, d);
    if ((threadIdx.x & 31) == 0 && lf) atomicAdd(&stats->n_lf, lf);
}

// ---------------------------------------------------------------------------------------------
static FmView fm_view(const bwagpu_ctx *ctx)
{
    FmView f;
    f.count_tab = nullptr;
    f.cp_occ = (const int64_t *)ctx->d_cp_occ;
    f.sa_ms_byte = (const int8_t *)ctx->d_sa_ms;
    f.sa_ls_word = (const uint32_t *)ctx->d_sa_ls;
    for (int i = 0; i < 5; ++i) f.count[i] = ctx->fm_count[i];
    f.sentinel_index = ctx->fm_sentinel;
    f.ref_seq_len = ctx->fm_ref_seq_len;
    f.fma_all = ctx->fma_enabled ? (const uint32_t *)ctx->d_fma_all : nullptr;
    f.fma_last = ctx->fma_enabled ? (const uint32_t *)ctx->d_fma_last : nullptr;
    return f;
}

// ---- FMA tables -------------------------------------------------------------------------------
// One thread per k-mer (rank = the k-mer read as a base-4 number, first base most significant,
// reference __seq_next src/FMI_search.cpp:63-76): walk it forward from its first base.
// all_smem_t entry (src/FMI_search.cpp:78-120): for step i (base i of the k-mer, 1..10)
// k32 = k - previous k, l32 = l - count[3 - base], s32 = s; the walk stops after storing the
// first step with s == 0; last_avail = last step with s > 0; the rest of the entry stays 0.
__global__ void fma_build_all_kernel(FmView f, uint32_t *__restrict__ tab)
{
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (1u << (2 * FMA_ALL_BP))) return;
    uint32_t e[32];
    fma_entry_all(f, idx, e);
    uint4 *dst = reinterpret_cast<uint4 *>(tab + (size_t)idx * 32);
#pragma unroll
    for (int i = 0; i < 8; ++i) dst[i] = make_uint4(e[4 * i], e[4 * i + 1], e[4 * i + 2], e[4 * i + 3]);
}

__global__ void fma_build_last_kernel(FmView f, uint32_t *__restrict__ tab)
{
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (1u << (2 * FMA_LAST_BP))) return;
    uint32_t e[4];
    fma_entry_last(f, idx, e);
    reinterpret_cast<uint4 *>(tab)[idx] = make_uint4(e[0], e[1], e[2], e[3]);
}

static int fma_alloc(bwagpu_ctx *ctx, bool all, bool last)
{
    if (all && !ctx->d_fma_all) BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_fma_all, (size_t)BWAGPU_FMA_ALL_BYTES));
    if (last && !ctx->d_fma_last) BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_fma_last, (size_t)BWAGPU_FMA_LAST_BYTES));
    return BWAGPU_OK;
}

extern "C" int bwagpu_index_build_fma(bwagpu_ctx *ctx)
{
    BWAGPU_OWNS_INDEX(ctx);
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!ctx->has_fm) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_build_fma: no FM-index loaded");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = fma_alloc(ctx, true, true);
    if (rc) return rc;
    FmView f = fm_view(ctx);
    BWAGPU_KERNEL_BEGIN(ctx);
    fma_build_all_kernel<<<(1u << (2 * FMA_ALL_BP)) / 128, 128, 0, ctx->stream>>>(f, (uint32_t *)ctx->d_fma_all);
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_FMA_BUILD);
    BWAGPU_KERNEL_BEGIN(ctx);
    fma_build_last_kernel<<<(1u << (2 * FMA_LAST_BP)) / 128, 128, 0, ctx->stream>>>(f, (uint32_t *)ctx->d_fma_last);
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_FMA_BUILD);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    ctx->fma_enabled = 1;
    return BWAGPU_OK;
}

extern "C" int bwagpu_index_load_fma(bwagpu_ctx *ctx, const void *all_tab, const void *last_tab)
{
    BWAGPU_OWNS_INDEX(ctx);
    if (!ctx) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    if (!all_tab && ctx->d_fma_all) { cudaFree(ctx->d_fma_all); ctx->d_fma_all = nullptr; }
    if (!last_tab && ctx->d_fma_last) { cudaFree(ctx->d_fma_last); ctx->d_fma_last = nullptr; }
    int rc = fma_alloc(ctx, all_tab != nullptr, last_tab != nullptr);
    if (rc) return rc;
    if (all_tab) BWAGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_fma_all, all_tab, (size_t)BWAGPU_FMA_ALL_BYTES, cudaMemcpyHostToDevice, ctx->stream));
    if (last_tab) BWAGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_fma_last, last_tab, (size_t)BWAGPU_FMA_LAST_BYTES, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    ctx->fma_enabled = (all_tab || last_tab) ? 1 : 0;
    return BWAGPU_OK;
}

extern "C" int bwagpu_index_get_fma(bwagpu_ctx *ctx, void *all_tab, void *last_tab)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    if ((all_tab && !ctx->d_fma_all) || (last_tab && !ctx->d_fma_last))
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_get_fma: table not resident");
    if (all_tab) BWAGPU_CUDA(ctx, cudaMemcpyAsync(all_tab, ctx->d_fma_all, (size_t)BWAGPU_FMA_ALL_BYTES, cudaMemcpyDeviceToHost, ctx->stream));
    if (last_tab) BWAGPU_CUDA(ctx, cudaMemcpyAsync(last_tab, ctx->d_fma_last, (size_t)BWAGPU_FMA_LAST_BYTES, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}

extern "C" int bwagpu_index_use_fma(bwagpu_ctx *ctx, int on)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (on && !ctx->d_fma_all && !ctx->d_fma_last)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_use_fma: no FMA table resident (bwagpu_index_build_fma / _load_fma)");
    ctx->fma_enabled = on ? 1 : 0;
    return BWAGPU_OK;
}

// CP_OCC records between the reference's word order and the paired order of the replica (fmi_core.cuh, FM_PAIRED);
// src == dst converts in place (every thread owns one record).
__global__ void fm_layout_kernel(const int64_t *src, int64_t *dst, int64_t n_cp, int to_paired)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_cp) return;
    int64_t w[8], o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) w[j] = src[i * 8 + j];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        if (to_paired) { o[2 * b] = w[b]; o[2 * b + 1] = w[4 + b]; }
        else { o[b] = w[2 * b]; o[4 + b] = w[2 * b + 1]; }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) dst[i * 8 + j] = o[j];
}

int bwagpu_fm_convert_layout(bwagpu_ctx *ctx, const void *src, void *dst, int64_t n_cp, int to_paired)
{
#if FM_PAIRED
    fm_layout_kernel<<<(unsigned)((n_cp + 255) / 256), 256, 0, ctx->stream>>>((const int64_t *)src, (int64_t *)dst, n_cp, to_paired);
    BWAGPU_CUDA(ctx, cudaGetLastError());
#else
    if (src != dst) BWAGPU_CUDA(ctx, cudaMemcpyAsync(dst, src, (size_t)n_cp * 64, cudaMemcpyDeviceToDevice, ctx->stream));
#endif
    return BWAGPU_OK;
}

extern "C" int bwagpu_index_load_fm(bwagpu_ctx *ctx, const bwagpu_fm_index_desc *d)
{
    BWAGPU_OWNS_INDEX(ctx);
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!d || !d->cp_occ || !d->sa_ms_byte || !d->sa_ls_word || d->reference_seq_len <= 0)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_load_fm: incomplete index descriptor");
    if (d->reference_seq_len >= (1LL << 38))
        BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "index_load_fm: text longer than 2^38 symbols is not supported");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    if (ctx->d_cp_occ) { cudaFree(ctx->d_cp_occ); cudaFree(ctx->d_sa_ms); cudaFree(ctx->d_sa_ls); }
    ctx->d_cp_occ = ctx->d_sa_ms = ctx->d_sa_ls = nullptr;
    ctx->has_fm = false;
    if (ctx->d_fma_all) { cudaFree(ctx->d_fma_all); ctx->d_fma_all = nullptr; }
    if (ctx->d_fma_last) { cudaFree(ctx->d_fma_last); ctx->d_fma_last = nullptr; }
    ctx->fma_enabled = 0;
    const int64_t n_cp = (d->reference_seq_len >> 6) + 1, n_sa = (d->reference_seq_len >> 3) + 1;
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_cp_occ, (size_t)n_cp * 64));
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_sa_ms, (size_t)n_sa));
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_sa_ls, (size_t)n_sa * 4));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_cp_occ, d->cp_occ, (size_t)n_cp * 64, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_sa_ms, d->sa_ms_byte, (size_t)n_sa, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_sa_ls, d->sa_ls_word, (size_t)n_sa * 4, cudaMemcpyHostToDevice, ctx->stream));
    { const int rc = bwagpu_fm_convert_layout(ctx, ctx->d_cp_occ, ctx->d_cp_occ, n_cp, 1); if (rc) return rc; }
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    for (int i = 0; i < 5; ++i) ctx->fm_count[i] = d->count[i];
    ctx->fm_sentinel = d->sentinel_index;
    ctx->fm_ref_seq_len = d->reference_seq_len;
    ctx->has_fm = true;
    return BWAGPU_OK;
}

static int seed_check(bwagpu_ctx *ctx, const bwagpu_seed_opt *opt)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!opt) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "seed: options are NULL");
    if (!ctx->has_fm) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "seed: no FM-index loaded (bwagpu_index_load_fm)");
    if (opt->min_seed_len < 1 || opt->max_occ < 1)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "seed: min_seed_len and max_occ must be positive");
    return BWAGPU_OK;
}

static int scan_i32_to_i64(bwagpu_ctx *ctx, const int32_t *d_in, int64_t *d_out, int64_t n)
{
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveScan((void *)nullptr, bytes, d_in, d_out, cub::Sum(), (int64_t)0, (int)n, ctx->stream);
    void *tmp;
    int rc = bwagpu_scratch(ctx, SCR_CUB, (int64_t)bytes + 256, &tmp);
    if (rc) return rc;
    BWAGPU_CUDA(ctx, cub::DeviceScan::ExclusiveScan(tmp, bytes, d_in, d_out, cub::Sum(), (int64_t)0, (int)n, ctx->stream));
    ctx->launches += 2;      // CUB's scan is two kernels (tile state init + scan)
    return BWAGPU_OK;
}

static int scan_i64(bwagpu_ctx *ctx, const int64_t *d_in, int64_t *d_out, int64_t n)
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

// SMEM collection with device-resident reads; *d_out is context-owned (SCR_SEED_OUT).
static int seed_collect_dev(bwagpu_ctx *ctx, const bwagpu_seed_opt *opt, const uint8_t *d_reads,
                            const int64_t *d_off, const int32_t *d_len, const uint8_t *d_skip,
                            int32_t nseq, int32_t max_len, bwagpu_smem **d_out, int64_t *tot,
                            SeedStats *h_stats)
{
    *tot = 0; *d_out = nullptr;
    memset(h_stats, 0, sizeof *h_stats);
    if (nseq == 0) return BWAGPU_OK;
    SeedOpt so;
    so.min_seed_len = opt->min_seed_len;
    so.split_len = (int)(opt->min_seed_len * opt->split_factor + .499);
    so.split_width = opt->split_width;
    so.max_mem_intv = opt->max_mem_intv;
    const bool staged = max_len <= SEED_SMEM_BASES;
    const size_t smem = (size_t)(SEED_PREV_K * 4 + (staged ? (max_len + 7) / 8 : 0)) * SEED_TPB * 4;
    static BwagpuSmemBound bound_staged, bound_plain;
    if (staged) BWAGPU_CUDA(ctx, bwagpu_raise_dyn_smem(bound_staged, ctx->device, seed_collect_kernel<true>, (int)smem));
    else BWAGPU_CUDA(ctx, bwagpu_raise_dyn_smem(bound_plain, ctx->device, seed_collect_kernel<false>, (int)smem));
    int per_sm = 0;
    if (staged) BWAGPU_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, seed_collect_kernel<true>, SEED_TPB, smem));
    else BWAGPU_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, seed_collect_kernel<false>, SEED_TPB, smem));
    if (per_sm < 1) per_sm = 1;
    if (getenv("BWAGPU_SEED_BLOCKS_PER_SM")) {                 // experiments: fewer resident blocks than fit
        const int v = atoi(getenv("BWAGPU_SEED_BLOCKS_PER_SM"));
        if (v >= 1 && v < per_sm) per_sm = v;
    }
    int blocks = ctx->sm_count * per_sm;
    const int need = (nseq + SEED_TPB - 1) / SEED_TPB;
    if (blocks > need) blocks = need;
    const int64_t T = (int64_t)blocks * SEED_TPB;
    // Common slab of up to 128 entries per read (only what a read finds is ever written; per call, never sticky); the few reads that find more (repeats) are collected in a list and
    // re-run once with private slabs sized for the largest of them, so one repeat-heavy read costs one small launch, not a
    // second pass over the batch, and the common slab never grows.
    int cap = 128;                                            // widest slab that keeps the scratch under 6 GiB
    while (cap > 32 && (int64_t)nseq * cap * (int64_t)sizeof(bwagpu_smem) > (6LL << 30)) cap >>= 1;
    if (getenv("BWAGPU_SEED_CAP")) cap = atoi(getenv("BWAGPU_SEED_CAP")) > 0 ? atoi(getenv("BWAGPU_SEED_CAP")) : cap;   // tests: force the overflow path
    int32_t *d_cnt, *d_ovf, *d_slab_of; int64_t *d_offs, *d_prev; SeedStats *d_stats; bwagpu_smem *d_tmp, *d_tmp2 = nullptr;
    int rc, cap2 = cap;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_CNT, (int64_t)(nseq + 1) * 4, (void **)&d_cnt))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_SCAN, (int64_t)(nseq + 1) * 8, (void **)&d_offs))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_PREV, T * 3 * (int64_t)(max_len + 1) * 8, (void **)&d_prev))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_STATS, sizeof(SeedStats), (void **)&d_stats))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_OVF, (int64_t)(nseq + 1) * 8, (void **)&d_ovf))) return rc;
    d_slab_of = d_ovf + nseq + 1;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_TMP, (int64_t)nseq * cap * sizeof(bwagpu_smem), (void **)&d_tmp))) return rc;
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_stats, 0, sizeof(SeedStats), ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_cnt + nseq, 0, 4, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_slab_of, 0xFF, (size_t)nseq * 4, ctx->stream));
    // prev-list entries per thread in the 16-byte layout (0 = interleaved 24-byte layout: reads too long for 14 bits of n)
    const int prev_cap16 = (SEED_PREV16 && max_len < 16384) ? max_len + 1 : 0;
    auto launch = [&](int n_items, bwagpu_smem *slab, int slab_cap, const int32_t *list) {
        cudaMemsetAsync(&d_stats->next, 0, 4, ctx->stream);      // the read counter of SEED_DYN
        int b = blocks;
        const int nd = (n_items + SEED_TPB - 1) / SEED_TPB;
        if (b > nd) b = nd;
        BWAGPU_KERNEL_BEGIN(ctx);
        if (staged)
            seed_collect_kernel<true><<<b, SEED_TPB, smem, ctx->stream>>>(fm_view(ctx), so, d_reads, d_off, d_len, d_skip, n_items,
                                                                         slab, slab_cap, d_cnt, d_prev, d_stats, list, d_ovf, d_slab_of, prev_cap16);
        else
            seed_collect_kernel<false><<<b, SEED_TPB, smem, ctx->stream>>>(fm_view(ctx), so, d_reads, d_off, d_len, d_skip, n_items,
                                                                          slab, slab_cap, d_cnt, d_prev, d_stats, list, d_ovf, d_slab_of, prev_cap16);
        BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_SEED_COLLECT);
    };
    launch(nseq, d_tmp, cap, nullptr);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(h_stats, d_stats, sizeof(SeedStats), cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    if (h_stats->n_ovf > 0) {
        const int n_ovf = (int)h_stats->n_ovf;
        if (getenv("BWAGPU_VERBOSE")) fprintf(stderr, "[bwagpu] seed: %d of %d reads outgrow the %d-entry slab (max %llu)\n", n_ovf, nseq, cap, h_stats->max_cnt);
        // a truncated slab also truncates the read's own round-2 input, so the count it reports can still be short of
        // what it finds with room: repeat until every listed read fits
        for (int pass = 0;; ++pass) {
            while (cap2 < (int64_t)h_stats->max_cnt) cap2 *= 2;
            if (pass > 8 || (int64_t)n_ovf * cap2 * (int64_t)sizeof(bwagpu_smem) > (8LL << 30))
                BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "seed: %d reads need up to %d SMEM slots each; their scratch would exceed 8 GiB", n_ovf, cap2);
            if ((rc = bwagpu_scratch(ctx, SCR_SEED_TMP2, (int64_t)n_ovf * cap2 * sizeof(bwagpu_smem), (void **)&d_tmp2))) return rc;
            BWAGPU_CUDA(ctx, cudaMemsetAsync(&d_stats->max_cnt, 0, 8, ctx->stream));
            launch(n_ovf, d_tmp2, cap2, d_ovf);          // its extensions add to the statistics: they are work done
            BWAGPU_CUDA(ctx, cudaGetLastError());
            BWAGPU_CUDA(ctx, cudaMemcpyAsync(h_stats, d_stats, sizeof(SeedStats), cudaMemcpyDeviceToHost, ctx->stream));
            BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
            if ((int64_t)h_stats->max_cnt <= cap2) break;
        }
    }
    if ((rc = scan_i32_to_i64(ctx, d_cnt, d_offs, nseq + 1))) return rc;
    int64_t total = 0;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&total, d_offs + nseq, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    bwagpu_smem *out;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_OUT, (total + 1) * (int64_t)sizeof(bwagpu_smem), (void **)&out))) return rc;
    BWAGPU_KERNEL_BEGIN(ctx);
    seed_sort_kernel<<<(nseq + 127) / 128, 128, 0, ctx->stream>>>(d_tmp, cap, d_tmp2, cap2, d_slab_of, d_cnt, d_offs, nseq, out);
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_SEED_SORT);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    *d_out = out; *tot = total;
    return BWAGPU_OK;
}

// SA look-up with device-resident SMEMs; outputs context-owned (SCR_SA_COORD / SCR_SA_OFF).
static int sa_lookup_dev(bwagpu_ctx *ctx, const bwagpu_smem *d_sm, int64_t n, int32_t max_occ,
                         int64_t **d_coords, int64_t **d_coord_off, int64_t *n_coords, SeedStats *h_stats)
{
    *n_coords = 0; *d_coords = nullptr; *d_coord_off = nullptr;
    int64_t *rows, *offs; SeedStats *d_stats;
    int rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SA_CNT, (n + 1) * 8, (void **)&rows))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SA_OFF, (n + 1) * 8, (void **)&offs))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_STATS, sizeof(SeedStats), (void **)&d_stats))) return rc;
    *d_coord_off = offs;
    BWAGPU_CUDA(ctx, cudaMemsetAsync(rows + n, 0, 8, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(&d_stats->n_lf, 0, 8, ctx->stream));
    if (n > 0) {
        sa_count_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(d_sm, n, max_occ, rows);
        ctx->launches++;
    }
    if ((rc = scan_i64(ctx, rows, offs, n + 1))) return rc;
    int64_t total = 0;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&total, offs + n, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    int64_t *coords;
    if ((rc = bwagpu_scratch(ctx, SCR_SA_COORD, (total + 1) * 8, (void **)&coords))) return rc;
    if (total > 0) {
        BWAGPU_KERNEL_BEGIN(ctx);
        sa_walk_kernel<<<(unsigned)((total + 255) / 256), 256, 0, ctx->stream>>>(fm_view(ctx), d_sm, n, offs, total,
                                                                               max_occ, coords, d_stats);
        BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_SA_WALK);
        BWAGPU_CUDA(ctx, cudaGetLastError());
    }
    if (h_stats) {
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(&h_stats->n_lf, &d_stats->n_lf, 8, cudaMemcpyDeviceToHost, ctx->stream));
        BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    }
    *d_coords = coords; *n_coords = total;
    return BWAGPU_OK;
}

extern "C" int bwagpu_seed_fm_device(bwagpu_ctx *ctx, const bwagpu_seed_opt *opt, const uint8_t *d_reads,
                                     const int64_t *d_off, const int32_t *d_len, const uint8_t *d_skip,
                                     int32_t nseq, int32_t max_len, bwagpu_seed_result *res)
{
    int rc = seed_check(ctx, opt);
    if (rc) return rc;
    if (!res || nseq < 0 || max_len < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "seed_fm_device: bad argument");
    memset(res, 0, sizeof *res);
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    bwagpu_smem *d_sm; int64_t tot;
    SeedStats st;
    if ((rc = seed_collect_dev(ctx, opt, d_reads, d_off, d_len, d_skip, nseq, max_len, &d_sm, &tot, &st))) return rc;
    int64_t *d_coords, *d_coff, nc;
    SeedStats st2; memset(&st2, 0, sizeof st2);
    if (tot > 0) {
        if ((rc = sa_lookup_dev(ctx, d_sm, tot, opt->max_occ, &d_coords, &d_coff, &nc, &st2))) return rc;
        res->d_coords = d_coords; res->n_coords = nc; res->d_coord_off = d_coff;
    }
    res->d_smems = d_sm; res->n_smem = tot;
    res->d_smem_off = (const int64_t *)ctx->scratch[SCR_SEED_SCAN];
    res->n_ext = (int64_t)st.n_ext; res->n_cp_records = (int64_t)st.n_rec; res->n_lf_steps = (int64_t)st2.n_lf;
    return BWAGPU_OK;
}

extern "C" int bwagpu_collect_smem(bwagpu_ctx *ctx, const bwagpu_seed_opt *opt, const uint8_t *enc_reads,
                                   const int64_t *off, const int32_t *len, const uint8_t *skip, int32_t nseq,
                                   bwagpu_smem *out, int64_t out_cap, int64_t *tot_smem)
{
    int rc = seed_check(ctx, opt);
    if (rc) return rc;
    if (!tot_smem || nseq < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "collect_smem: bad argument");
    *tot_smem = 0;
    if (nseq == 0) return BWAGPU_OK;
    if (!enc_reads || !off || !len) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "collect_smem: NULL buffer");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    int64_t bytes = 0; int max_len = 0;
    for (int i = 0; i < nseq; ++i) {
        if (len[i] < 0 || off[i] < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "collect_smem: read %d has a negative length/offset", i);
        if (len[i] > 32767) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "collect_smem: read %d longer than 32767 (int16 query positions upstream)", i);
        if (off[i] + len[i] > bytes) bytes = off[i] + len[i];
        if (len[i] > max_len) max_len = len[i];
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
    bwagpu_smem *d_sm; int64_t tot; SeedStats st;
    if ((rc = seed_collect_dev(ctx, opt, d_reads, d_off, d_len, d_skip, nseq, max_len, &d_sm, &tot, &st))) return rc;
    *tot_smem = tot;
    if (tot > out_cap) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "collect_smem: %lld SMEMs do not fit out_cap=%lld", (long long)tot, (long long)out_cap);
    if (tot > 0) {
        if (!out) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "collect_smem: out is NULL");
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(out, d_sm, (size_t)tot * sizeof(bwagpu_smem), cudaMemcpyDeviceToHost, ctx->stream));
    }
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}

extern "C" int bwagpu_get_sa_entries(bwagpu_ctx *ctx, const bwagpu_smem *smems, int64_t n_smem, int32_t max_occ,
                                     int64_t *coords, int64_t coords_cap, int64_t *n_coords)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!ctx->has_fm) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "get_sa_entries: no FM-index loaded");
    if (!n_coords || n_smem < 0 || max_occ < 1) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "get_sa_entries: bad argument");
    *n_coords = 0;
    if (n_smem == 0) return BWAGPU_OK;
    if (!smems) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "get_sa_entries: smems is NULL");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    bwagpu_smem *d_sm;
    int rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SA_IN, n_smem * (int64_t)sizeof(bwagpu_smem), (void **)&d_sm))) return rc;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_sm, smems, (size_t)n_smem * sizeof(bwagpu_smem), cudaMemcpyHostToDevice, ctx->stream));
    int64_t *d_coords, *d_coff, nc;
    if ((rc = sa_lookup_dev(ctx, d_sm, n_smem, max_occ, &d_coords, &d_coff, &nc, nullptr))) return rc;
    *n_coords = nc;
    if (nc > coords_cap) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "get_sa_entries: %lld coordinates do not fit coords_cap=%lld", (long long)nc, (long long)coords_cap);
    if (nc > 0) {
        if (!coords) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "get_sa_entries: coords is NULL");
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(coords, d_coords, (size_t)nc * 8, cudaMemcpyDeviceToHost, ctx->stream));
    }
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}
