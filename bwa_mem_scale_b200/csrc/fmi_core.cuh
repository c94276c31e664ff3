// FM-index seeding core: backward extension over CP_OCC, the per-read SMEM state machine
// (seeding rounds 1-3) and the suffix-array walk.  Everything here is per-thread code shared by
// the CUDA kernels in seed.cu; it is also compilable as plain host C++ so that tests can single-step
// the exact same logic without a GPU (tests/hostsim), which is test scaffolding, not a product path.
//
// Reference behaviour reproduced:
//   fm_bext              FMI_search::backwardExt + GET_OCC     src/FMI_search.cpp:2029-2056, .h:76-83
//   SeedMachine          getSMEMsOnePosOneThread (rounds 1, 2)  src/FMI_search.cpp:1372-1606
//                        bwtSeedStrategyAllPosOneThread (3)     src/FMI_search.cpp:1662-1816
//                        round orchestration of mem_collect_smem src/bwamem.cpp:648-786
//   fm_sa_walk           call_one_step / get_sa_entries_prefetch src/FMI_search.cpp:2206-2379
#pragma once
#include <stdint.h>
#include "../../include/bwagpu.h"

#if defined(__CUDACC__)
#define BWAGPU_HD __host__ __device__ __forceinline__
#else
#define BWAGPU_HD inline
#endif

// Layout of a 64-byte CP_OCC record.  The reference's is cp_count[4] then one_hot_bwt_str[4] (src/FMI_search.h:76-83).  The
// replica in HBM is PAIRED: count and one-hot word of a base side by side, bases 0,1 in the first 32-byte sector and bases
// 2,3 in the second.  A backward extension by base a needs the occurrences of at most two bases (fm_bext), and with this
// order they always share a sector: one 32-byte load per record instead of the whole record — half the bytes, half the
// registers.  The loaders (bwagpu_index_load_fm, bwagpu_index_build_fm) convert once; bwagpu_index_get_fm converts back.
// Host builds of this header (tests/hostsim single-stepping the same logic on reference arrays) keep the reference order.
#ifndef FM_PAIRED
#if defined(__CUDACC__)
#define FM_PAIRED 1
#else
#define FM_PAIRED 0
#endif
#endif
#define FM_W_CNT(b) (FM_PAIRED ? 2 * (b) : (b))           // word index of cp_count[b] inside a record
#define FM_W_BITS(b) (FM_PAIRED ? 2 * (b) + 1 : 4 + (b))  // word index of one_hot_bwt_str[b]

// Device view of the FM-index replica held in HBM.  count[] is the loaded form (file value + 1).
struct FmView {
    const int64_t *cp_occ;       // 8 x 64-bit words per record, see FM_PAIRED
    const int8_t *sa_ms_byte;
    const uint32_t *sa_ls_word;
    int64_t count[5];
    int64_t sentinel_index;
    int64_t ref_seq_len;
    // FMA tables (reference all_smem_t / last_smem_t, src/FMI_search.h:99-135); NULL = not loaded
    const uint32_t *fma_all;     // 32 words per 11-mer: last_avail, 10 x {k32,l32,s32}, pad
    const uint32_t *fma_last;    // 4 words per 13-mer: bp|kms<<8|lms<<16|sms<<24, kls, lls, sls
    const int64_t *count_tab;    // NULL, or an indexable copy of count[5] the kernel made (fm_count)
};

// SEED_PAIR2 = 1: in the backward rows of the SMEM search two entries of the row are extended per step (fm_bext2): they
// are independent until their results are applied, which still happens in list order.
#ifndef SEED_PAIR2
#define SEED_PAIR2 1
#endif
// SEED_ROW_FWD (SeedMachine::step_fast): MEASURED 247.3 ms per 5 M reads with it against 239.0 without (scripts/gpu_r2_af.sh,
// bit-exact either way) — the load it saves was hidden behind other warps, the longer step is not.  OFF; kept for the record.
#ifndef SEED_ROW_FWD
#define SEED_ROW_FWD 0
#endif

#define FMA_ALL_BP 11
#define FMA_LAST_BP 13

// 64 bytes (16 words) of a table entry
BWAGPU_HD void fm_load_words16(const uint32_t *p, uint32_t e[16])
{
#if defined(__CUDA_ARCH__)
    uint64_t a0, a1, a2, a3, b0, b1, b2, b3;
    asm("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(a0), "=l"(a1), "=l"(a2), "=l"(a3) : "l"(p));
    asm("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(b0), "=l"(b1), "=l"(b2), "=l"(b3) : "l"(p + 8));
    e[0] = (uint32_t)a0; e[1] = (uint32_t)(a0 >> 32); e[2] = (uint32_t)a1; e[3] = (uint32_t)(a1 >> 32);
    e[4] = (uint32_t)a2; e[5] = (uint32_t)(a2 >> 32); e[6] = (uint32_t)a3; e[7] = (uint32_t)(a3 >> 32);
    e[8] = (uint32_t)b0; e[9] = (uint32_t)(b0 >> 32); e[10] = (uint32_t)b1; e[11] = (uint32_t)(b1 >> 32);
    e[12] = (uint32_t)b2; e[13] = (uint32_t)(b2 >> 32); e[14] = (uint32_t)b3; e[15] = (uint32_t)(b3 >> 32);
#else
    for (int i = 0; i < 16; ++i) e[i] = p[i];
#endif
}

BWAGPU_HD void fm_load_words4(const uint32_t *p, uint32_t e[4])
{
#if defined(__CUDA_ARCH__)
    const uint4 v = __ldg(reinterpret_cast<const uint4 *>(p));
    e[0] = v.x; e[1] = v.y; e[2] = v.z; e[3] = v.w;
#else
    for (int i = 0; i < 4; ++i) e[i] = p[i];
#endif
}

struct FmRec { int64_t cnt[4]; uint64_t bits[4]; };

// SEED_PREFETCH: a bi-interval that survives a backward row (or is pushed by the forward walk) WILL be extended again
// one row later, and the two CP_OCC records that extension reads are known now: ask L2 for them (no register is held),
// so that the dependent load a row later finds them on chip.  0 = off, 1 = one prefetch per record, 2 = one per sector.
#ifndef SEED_PREFETCH
#define SEED_PREFETCH 0
#endif
BWAGPU_HD void fm_prefetch_rec(const int64_t *cp_occ, int64_t block)
{
#if defined(__CUDA_ARCH__) && SEED_PREFETCH > 0
    const int64_t *p = cp_occ + block * 8;
    asm volatile("prefetch.global.L2 [%0];" :: "l"(p));
#if SEED_PREFETCH > 1
    asm volatile("prefetch.global.L2 [%0];" :: "l"(p + 4));
#endif
#else
    (void)cp_occ; (void)block;
#endif
}
// records a backward extension of (k, s) will read
BWAGPU_HD void fm_prefetch_bext(const int64_t *cp_occ, int64_t k, int64_t s)
{
    const int64_t b0 = k >> 6, b1 = (k + s) >> 6;
    fm_prefetch_rec(cp_occ, b0);
    if (b1 != b0) fm_prefetch_rec(cp_occ, b1);
}


// `deep`: the interval is narrow, i.e. this record is a random touch with no reuse inside the SM
BWAGPU_HD void fm_load_rec(const FmView &f, int64_t block, FmRec &r, bool deep = false)
{
#if defined(__CUDA_ARCH__)
    const int64_t *p = f.cp_occ + block * 8;
    (void)deep;
    // one 64-byte record = two 256-bit loads (sm_100: LDG.256), one per 32-byte sector
    uint64_t w0, w1, w2, w3, w4, w5, w6, w7;
    asm("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(w0), "=l"(w1), "=l"(w2), "=l"(w3) : "l"(p));
    asm("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(w4), "=l"(w5), "=l"(w6), "=l"(w7) : "l"(p + 4));
#if FM_PAIRED
    r.cnt[0] = (int64_t)w0; r.bits[0] = w1; r.cnt[1] = (int64_t)w2; r.bits[1] = w3;
    r.cnt[2] = (int64_t)w4; r.bits[2] = w5; r.cnt[3] = (int64_t)w6; r.bits[3] = w7;
#else
    r.cnt[0] = (int64_t)w0; r.cnt[1] = (int64_t)w1; r.cnt[2] = (int64_t)w2; r.cnt[3] = (int64_t)w3;
    r.bits[0] = w4; r.bits[1] = w5; r.bits[2] = w6; r.bits[3] = w7;
#endif
#else
    (void)deep;
    const int64_t *p = f.cp_occ + block * 8;
    for (int i = 0; i < 4; ++i) { r.cnt[i] = p[FM_W_CNT(i)]; r.bits[i] = (uint64_t)p[FM_W_BITS(i)]; }
#endif
}

// Half a record: counts and one-hot words of bases 2h and 2h + 1 (one 32-byte sector of the paired layout).
struct FmHalf { int64_t cnt[2]; uint64_t bits[2]; };
BWAGPU_HD void fm_load_half(const FmView &f, int64_t block, int h, FmHalf &r)
{
    const int64_t *p = f.cp_occ + block * 8;
#if defined(__CUDA_ARCH__) && FM_PAIRED
    asm("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(r.cnt[0]), "=l"(r.bits[0]), "=l"(r.cnt[1]), "=l"(r.bits[1]) : "l"(p + 4 * h));
#else
    for (int i = 0; i < 2; ++i) { r.cnt[i] = p[FM_W_CNT(2 * h + i)]; r.bits[i] = (uint64_t)p[FM_W_BITS(2 * h + i)]; }
#endif
}

BWAGPU_HD int fm_popc(uint64_t v)
{
#if defined(__CUDA_ARCH__)
    return __popcll(v);
#else
    return __builtin_popcountll(v);
#endif
}

// mask selecting the first y symbols of a block (one-hot words are MSB-first)
BWAGPU_HD uint64_t fm_prefix_mask(int y) { return y ? (~0ULL << (64 - y)) : 0ULL; }

struct Biv { int64_t k, l, s; };

// count[a] without dynamic indexing of the kernel-parameter struct (that would move it to local memory): a select chain,
// or — when the kernel has put a copy of count[] where it can be indexed (count_tab: shared memory) — one load.  The chain
// was 7.8 % of seed_collect_kernel's instructions (profiles/r02_ncu_seed_collect_v2_lines.txt).
BWAGPU_HD int64_t fm_count(const FmView &f, int a)
{
#if defined(__CUDA_ARCH__)
    if (f.count_tab) return f.count_tab[a];
#endif
    return a == 0 ? f.count[0] : a == 1 ? f.count[1] : a == 2 ? f.count[2] : a == 3 ? f.count[3] : f.count[4];
}

// backward extension of (k,l,s) by base a
BWAGPU_HD Biv fm_bext(const FmView &f, const Biv &v, int a)
{
    // The reference computes k[b], s[b] for all four bases and l[b] = l + sentinel + sum of s[c] over c > b, then keeps
    // base a's.  The rows of [sp, ep) split into the four bases and the sentinel row, s = s[0] + s[1] + s[2] + s[3] +
    // sentinel, so l[a] needs the sizes on ONE side of a only:
    //     a = 0, 1:  l[a] = l + s - (s[0] + .. + s[a])        a = 2, 3:  l[a] = l + sentinel + (s[a+1] + .. + s[3])
    // i.e. bases {0,1} or bases {2,3}: one half of each record.
    const int64_t sp = v.k, ep = v.k + v.s;
    const int h = a >> 1, lo = a & 1;
    FmHalf rs, re;
    fm_load_half(f, sp >> 6, h, rs);
    if ((ep >> 6) == (sp >> 6)) re = rs; else fm_load_half(f, ep >> 6, h, re);
    const uint64_t ms = fm_prefix_mask((int)(sp & 63)), me = fm_prefix_mask((int)(ep & 63));
    const int64_t o0 = rs.cnt[0] + fm_popc(rs.bits[0] & ms), o1 = rs.cnt[1] + fm_popc(rs.bits[1] & ms);
    const int64_t z0 = re.cnt[0] + fm_popc(re.bits[0] & me) - o0, z1 = re.cnt[1] + fm_popc(re.bits[1] & me) - o1;
    const int64_t sent = (v.k <= f.sentinel_index && ep > f.sentinel_index) ? 1 : 0;
    Biv r;
    r.k = fm_count(f, a) + (lo ? o1 : o0);
    r.s = lo ? z1 : z0;
    r.l = h == 0 ? v.l + v.s - z0 - (lo ? z1 : 0) : v.l + sent + (lo ? 0 : z1);
    return r;
}

// Two backward extensions by the same base a, loads of both issued before either is consumed (the second only when
// `two`): the entries of one backward row of the SMEM search are independent of each other (SeedMachine, SEED_PAIR2).
BWAGPU_HD Biv fm_bext2(const FmView &f, const Biv &v, const Biv &v2, bool two, int a, Biv &r2)
{
    const int64_t sp = v.k, ep = v.k + v.s, sp2 = v2.k, ep2 = v2.k + v2.s;
    const int h = a >> 1, lo = a & 1;
    FmHalf rs, re, qs, qe;
    fm_load_half(f, sp >> 6, h, rs);
    if ((ep >> 6) == (sp >> 6)) re = rs; else fm_load_half(f, ep >> 6, h, re);
    if (two) {
        fm_load_half(f, sp2 >> 6, h, qs);
        if ((ep2 >> 6) == (sp2 >> 6)) qe = qs; else fm_load_half(f, ep2 >> 6, h, qe);
    } else {
        qs.cnt[0] = qs.cnt[1] = 0; qs.bits[0] = qs.bits[1] = 0; qe = qs;
    }
    Biv r;
    {
        const uint64_t ms = fm_prefix_mask((int)(sp & 63)), me = fm_prefix_mask((int)(ep & 63));
        const int64_t o0 = rs.cnt[0] + fm_popc(rs.bits[0] & ms), o1 = rs.cnt[1] + fm_popc(rs.bits[1] & ms);
        const int64_t z0 = re.cnt[0] + fm_popc(re.bits[0] & me) - o0, z1 = re.cnt[1] + fm_popc(re.bits[1] & me) - o1;
        const int64_t sent = (v.k <= f.sentinel_index && ep > f.sentinel_index) ? 1 : 0;
        r.k = fm_count(f, a) + (lo ? o1 : o0);
        r.s = lo ? z1 : z0;
        r.l = h == 0 ? v.l + v.s - z0 - (lo ? z1 : 0) : v.l + sent + (lo ? 0 : z1);
    }
    {
        const uint64_t ms = fm_prefix_mask((int)(sp2 & 63)), me = fm_prefix_mask((int)(ep2 & 63));
        const int64_t o0 = qs.cnt[0] + fm_popc(qs.bits[0] & ms), o1 = qs.cnt[1] + fm_popc(qs.bits[1] & ms);
        const int64_t z0 = qe.cnt[0] + fm_popc(qe.bits[0] & me) - o0, z1 = qe.cnt[1] + fm_popc(qe.bits[1] & me) - o1;
        const int64_t sent = (v2.k <= f.sentinel_index && ep2 > f.sentinel_index) ? 1 : 0;
        r2.k = fm_count(f, a) + (lo ? o1 : o0);
        r2.s = lo ? z1 : z0;
        r2.l = h == 0 ? v2.l + v2.s - z0 - (lo ? z1 : 0) : v2.l + sent + (lo ? 0 : z1);
    }
    return r;
}

// ---- FMA table entries (what the reference's `smem-table` command computes) -------------------
// idx = rank of the k-mer read as a base-4 number, first base most significant (__seq_next,
// reference src/FMI_search.cpp:63-76); the k-mer is walked forward from its first base.
// all_smem_t (src/FMI_search.cpp:78-120): for step i (1..10) k32 = k - previous k,
// l32 = l - count[3 - base], s32 = s; the walk stops after storing the first step with s == 0;
// last_avail = last step with s > 0; the rest of the 128-byte entry stays 0.
BWAGPU_HD void fma_entry_all(const FmView &f, uint32_t idx, uint32_t e[32])
{
#pragma unroll
    for (int i = 0; i < 32; ++i) e[i] = 0;
    int a = (int)((idx >> (2 * (FMA_ALL_BP - 1))) & 3);
    Biv cur; cur.k = fm_count(f, a); cur.l = fm_count(f, 3 - a); cur.s = fm_count(f, a + 1) - cur.k;
    bool stop = false;
#pragma unroll
    for (int i = 1; i < FMA_ALL_BP; ++i) {
        if (!stop) {
            a = (int)((idx >> (2 * (FMA_ALL_BP - 1 - i))) & 3);
            Biv t; t.k = cur.l; t.l = cur.k; t.s = cur.s;
            const Biv r = fm_bext(f, t, 3 - a);
            Biv nw; nw.k = r.l; nw.l = r.k; nw.s = r.s;
            e[1 + 3 * (i - 1)] = (uint32_t)(nw.k - cur.k);
            e[2 + 3 * (i - 1)] = (uint32_t)(nw.l - fm_count(f, 3 - a));
            e[3 + 3 * (i - 1)] = (uint32_t)nw.s;
            if (nw.s > 0) e[0] = (uint32_t)i; else stop = true;
            cur = nw;
        }
    }
}

// last_smem_t (src/FMI_search.cpp:160-194): the interval after the longest prefix (bp bases, at
// most 13) that still occurs, as 8+32-bit halves: word 0 = bp | kms<<8 | lms<<16 | sms<<24.
BWAGPU_HD void fma_entry_last(const FmView &f, uint32_t idx, uint32_t e[4])
{
    int a = (int)((idx >> (2 * (FMA_LAST_BP - 1))) & 3);
    Biv cur; cur.k = fm_count(f, a); cur.l = fm_count(f, 3 - a); cur.s = fm_count(f, a + 1) - cur.k;
    int i = 1;
    for (; i < FMA_LAST_BP; ++i) {
        a = (int)((idx >> (2 * (FMA_LAST_BP - 1 - i))) & 3);
        Biv t; t.k = cur.l; t.l = cur.k; t.s = cur.s;
        const Biv r = fm_bext(f, t, 3 - a);
        if (r.s == 0) break;
        cur.k = r.l; cur.l = r.k; cur.s = r.s;
    }
    e[0] = (uint32_t)i | (((uint32_t)(cur.k >> 32) & 0xFF) << 8) | (((uint32_t)(cur.l >> 32) & 0xFF) << 16) |
           (((uint32_t)(cur.s >> 32) & 0xFF) << 24);
    e[1] = (uint32_t)cur.k; e[2] = (uint32_t)cur.l; e[3] = (uint32_t)cur.s;
}

// ---------------------------------------------------------------------------------------------
// Per-read seeding state machine.  One bext per step() so that the threads of a warp meet at a
// single load site; all the bookkeeping between two extensions is register/scratch work.
//
// Scratch `prev`: the bi-intervals found while extending forward from a pivot, kept from the END
// of the buffer downwards so that reading upwards from `base` visits them longest-first (the
// reference reverses its prevArray in place).  Entry = 3 words: k | n<<40, l, s.
// Output: up to `cap` SMEMs in `out` (40-byte reference layout); `n_out` keeps counting past cap
// so the launcher can detect overflow and retry with a larger capacity.
struct SeedOpt { int min_seed_len, split_len, split_width; int64_t max_mem_intv; };

// ReadStore: anything indexable as q[j] -> base code (a plain pointer on the host, a shared-memory
// view in the kernel).
template <class PrevStore, class ReadStore = const uint8_t *>
struct SeedMachine {
    enum { ST_PIVOT, ST_FWD, ST_BROW, ST_BITEM, ST_BEND, ST_NEXT, ST_LSTART, ST_LFWD, ST_DONE };
    ReadStore q;
    int len, rid;
    PrevStore prev;
    bwagpu_smem *out;
    int cap, n_out;
    int state, round, x, j, next_x, min_intv;
    int np, base, p, nc, decided, r1_count, r2_idx;
    int32_t last_s;
    Biv cur; int cur_n;
    Biv req; int req_a;          // pending extension request
#if SEED_PAIR2
    Biv req2; int cur_n2;        // backward rows: the next entry of the row rides along (has2)
#endif
    bool has2;
    int64_t n_ext;
    const int64_t *pf_cp;        // CP_OCC base for prefetches (NULL: none)

    BWAGPU_HD void init(ReadStore q_, int len_, int rid_, PrevStore prev_, bwagpu_smem *out_, int cap_)
    {
        q = q_; len = len_; rid = rid_; prev = prev_; out = out_; cap = cap_; n_out = 0;
        prev.reset(len_);
        round = 1; x = 0; min_intv = 1; state = len > 0 ? ST_PIVOT : ST_DONE; n_ext = 0;
        np = base = p = nc = decided = r1_count = r2_idx = 0; last_s = -1; j = next_x = 0;
        cur.k = cur.l = cur.s = 0; cur_n = 0; req = cur; req_a = 0;
        pf_cp = nullptr; has2 = false;
    }

    // backward row: entry p becomes the request; with SEED_PAIR2 entry p + 1 (if the row has one) rides along
    BWAGPU_HD void request_items()
    {
        int n_;
        prev.get(base + p, req, n_);
        cur_n = n_;
        req_a = q[j];
#if SEED_PAIR2
        has2 = p + 1 < np;
        if (has2) { prev.get(base + p + 1, req2, n_); cur_n2 = n_; }
#endif
    }

    BWAGPU_HD void emit(int m, int n, const Biv &v)
    {
        if (n_out < cap) {
            bwagpu_smem e;
            e.rid = (uint32_t)rid; e.m = (uint32_t)m; e.n = (uint32_t)n; e.pad_ = 0;
            e.k = v.k; e.l = v.l; e.s = v.s;
            out[n_out] = e;
        }
        ++n_out;
    }

    BWAGPU_HD void single_base(const FmView &f, int a)
    {
        cur.k = fm_count(f, a); cur.l = fm_count(f, 3 - a); cur.s = fm_count(f, a + 1) - cur.k;
    }

    BWAGPU_HD void push_prev()                 // forward phase: store below the previous entry
    {
        prev.put(len - 1 - np, cur, cur_n);
        if (pf_cp && x > 0) fm_prefetch_bext(pf_cp, cur.k, cur.s);      // the backward phase starts from these
    }

    BWAGPU_HD void request_fwd(int a)          // forward ext = backward ext of the complement on (l,k,s)
    {
        req.k = cur.l; req.l = cur.k; req.s = cur.s; req_a = 3 - a;
    }

    // One memoised forward step of the all-SMEM table (reference src/FMI_search.cpp:1435-1456):
    // returns true when the interval fell below min_intv (the walk of this pivot is over).
    BWAGPU_HD bool fma_step(const FmView &f, uint32_t k32, uint32_t l32, uint32_t s32, int jj)
    {
        const int a = q[jj];
        next_x = jj + 1;
        Biv nw;
        nw.k = cur.k + (int64_t)k32; nw.l = fm_count(f, 3 - a) + (int64_t)l32; nw.s = (int64_t)s32;
        if (nw.s != cur.s) { push_prev(); ++np; }
        if (nw.s < min_intv) { next_x = jj; return true; }
        cur = nw; cur_n = jj;
        return false;
    }

    // Forward phase through the all-SMEM table: the first (up to 9) extensions of a pivot come
    // from one 128-byte entry instead of CP_OCC.  Follows the reference literally, including what
    // it does when an N lies inside the 11-mer (the walk then ends at the table's last usable
    // step and next_x is set past it, src/FMI_search.cpp:1458-1461).
    BWAGPU_HD void fma_forward(const FmView &f)
    {
        uint32_t idx = 0;
        int kv = 0;
        for (; kv < FMA_ALL_BP; ++kv) {
            const uint32_t a = q[x + kv];
            if (a >= 4) break;
            idx |= a << ((FMA_ALL_BP - 1 - kv) * 2);
        }
        const bool with_n = kv < FMA_ALL_BP;
        const uint32_t *ent = f.fma_all + (size_t)idx * 32;
        uint32_t e[16];
        fm_load_words16(ent, e);
        const int last_avail = (int)e[0];
        const int last_idx = (kv > last_avail ? last_avail : kv) - 1;
        bool broke = false;
        int kk = 0;
#pragma unroll
        for (int u = 0; u < 5; ++u)
            if (!broke && kk < last_idx) { broke = fma_step(f, e[1 + 3 * u], e[2 + 3 * u], e[3 + 3 * u], x + 1 + kk); ++kk; }
        if (!broke && kk < last_idx) {
            fm_load_words16(ent + 16, e);
#pragma unroll
            for (int u = 0; u < 5; ++u)
                if (!broke && kk < last_idx) { broke = fma_step(f, e[3 * u], e[1 + 3 * u], e[2 + 3 * u], x + 1 + kk); ++kk; }
        }
        j = broke ? len : x + 1 + (last_idx > 0 ? last_idx : 0);
        if (with_n) { next_x = j + 1; j = len; }
        if (broke) { base = len - np; j = x - 1; state = ST_BROW; }
    }

    // Round 3 through the last-SMEM table (reference src/FMI_search.cpp:1708-1752): jump over the
    // first bp (<= 13) bases of a forward walk with one 16-byte entry.
    BWAGPU_HD void fma_last_jump(const FmView &f, const SeedOpt &o)
    {
        uint32_t idx = 0, with_n = 0;
        for (int kv = 0; kv < FMA_LAST_BP; ++kv) {
            const uint32_t a = q[x + kv];
            idx |= a << ((FMA_LAST_BP - 1 - kv) * 2);
            with_n += a >> 2;
        }
        if (with_n) return;
        uint32_t e[4];
        fm_load_words4(f.fma_last + (size_t)idx * 4, e);
        const int bp = (int)(e[0] & 0xFF);
        j = x + bp; next_x = j;
        cur.k = (int64_t)(((uint64_t)(int64_t)(int8_t)(e[0] >> 8) << 32) | e[1]);
        cur.l = (int64_t)(((uint64_t)(int64_t)(int8_t)(e[0] >> 16) << 32) | e[2]);
        cur.s = (int64_t)(((uint64_t)(int64_t)(int8_t)(e[0] >> 24) << 32) | e[3]);
        cur_n = j - 1;
        if (cur.s < o.max_mem_intv && cur_n - x + 1 >= o.min_seed_len + 1 && cur.s > 0) emit(x, cur_n, cur);
    }

    // Runs bookkeeping until an extension is needed (returns true, request in req/req_a) or the
    // read is finished (returns false).
    BWAGPU_HD bool advance(const FmView &f, const SeedOpt &o)
    {
        for (;;) {
            switch (state) {
            case ST_PIVOT: {                     // start of a pivot search (rounds 1 and 2)
                next_x = x + 1;
                const int a = q[x];
                if (a > 3) { state = ST_NEXT; break; }
                single_base(f, a); cur_n = x; np = 0; j = x + 1;
                state = ST_FWD;
                if (f.fma_all && len - x >= FMA_ALL_BP) fma_forward(f);
                break;
            }
            case ST_FWD: {
                if (j < len) {
                    next_x = j + 1;
                    const int a = q[j];
                    if (a < 4) { request_fwd(a); return true; }
                }
                // forward phase over: the last interval joins the list if still wide enough
                if (cur.s >= min_intv) { push_prev(); ++np; }
                base = len - np; j = x - 1;
                state = ST_BROW;
                break;
            }
            case ST_BROW: {
                if (j < 0 || q[j] > 3) { state = ST_BEND; break; }
                nc = 0; decided = 0; last_s = -1; p = 0;
                state = ST_BITEM;
                break;
            }
            case ST_BITEM: {
                if (p < np) { request_items(); return true; }      // cur_n = span end of the entry being extended
                np = nc;
                if (np == 0) { state = ST_NEXT; break; }
                --j;
                state = ST_BROW;
                break;
            }
            case ST_BEND: {                      // ran off the read start (or hit N) with survivors
                if (np != 0) {
                    Biv v; int n_;
                    prev.get(base, v, n_);
                    if (n_ - (j + 1) + 1 >= o.min_seed_len) emit(j + 1, n_, v);
                }
                state = ST_NEXT;
                break;
            }
            case ST_NEXT: {                      // pivot finished: choose the next piece of work
                if (round == 1) {
                    x = next_x;
                    if (x < len) { state = ST_PIVOT; break; }
                    round = 2; r1_count = n_out < cap ? n_out : cap; r2_idx = 0;
                }
                if (round == 2) {
                    bool found = false;
                    while (r2_idx < r1_count) {
                        const bwagpu_smem e = out[r2_idx++];
                        const int start = (int)e.m, end = (int)e.n + 1;
                        if (end - start < o.split_len || e.s > o.split_width) continue;
                        x = (end + start) >> 1; min_intv = (int)(e.s + 1);
                        found = true;
                        break;
                    }
                    if (found) { state = ST_PIVOT; break; }
                    round = 3; x = 0;
                    if (o.max_mem_intv <= 0) { state = ST_DONE; break; }
                    state = ST_LSTART;
                    break;
                }
                state = ST_DONE;
                break;
            }
            case ST_LSTART: {                    // round 3: forward-only, restart after each seed
                if (x >= len) { state = ST_DONE; break; }
                next_x = x + 1;
                const int a = q[x];
                if (a > 3) { x = next_x; break; }
                single_base(f, a); cur_n = x; j = x + 1;
                state = ST_LFWD;
                if (f.fma_last && len - x >= FMA_LAST_BP) fma_last_jump(f, o);
                break;
            }
            case ST_LFWD: {
                if (j < len) {
                    next_x = j + 1;
                    const int a = q[j];
                    if (a < 4) { request_fwd(a); return true; }
                }
                x = next_x;
                state = ST_LSTART;
                break;
            }
            default:
                return false;
            }
        }
    }

    // backward row: r extends entry `src` (span end src_n) to the left by q[j]; entries of this row start at m = j + 1
    // Returns true when r became an entry of the next row (stored at base + nc).
    BWAGPU_HD bool consume_item(const Biv &r, const Biv &src, int src_n, const SeedOpt &o)
    {
        bool kept = false;
        if (r.s < min_intv) {
            if (!decided && src_n - (j + 1) + 1 >= o.min_seed_len) {
                emit(j + 1, src_n, src);
                decided = 1;
            }
        } else if (r.s != (int64_t)last_s) {
            last_s = (int32_t)r.s;
            prev.put(base + nc, r, src_n);
            if (pf_cp && j > 0) fm_prefetch_bext(pf_cp, r.k, r.s);           // extended again in the next row
            ++nc;
            decided = 1;
            kept = true;
        }
        ++p;
        return kept;
    }

#if SEED_PAIR2
    // the second result of a paired request (has2): entry p of the same row, after consume() handled the first
    BWAGPU_HD void consume_second(const Biv &r2, const SeedOpt &o)
    {
        ++n_ext;
        consume_item(r2, req2, cur_n2, o);
    }
#endif

    // Consumes the result of the requested extension.
    BWAGPU_HD void consume(const Biv &r, const SeedOpt &o)
    {
        ++n_ext;
        if (state == ST_FWD) {
            Biv nw; nw.k = r.l; nw.l = r.k; nw.s = r.s;
            if (nw.s != cur.s) { push_prev(); ++np; }
            if (nw.s < min_intv) {
                next_x = j;
                // The reference stores `cur` a second time here (its post-loop push sees the
                // un-replaced interval).  The twin can never change an outcome in the backward
                // phase (it dies, or survives with the same s, exactly as its original does), so
                // it is not materialised: one backward extension fewer per pivot.
                base = len - np; j = x - 1;
                state = ST_BROW;
                return;
            }
            cur = nw; cur_n = j; ++j;
        } else if (state == ST_BITEM) {
            consume_item(r, req, cur_n, o);
        } else {                                 // ST_LFWD
            cur.k = r.l; cur.l = r.k; cur.s = r.s; cur_n = j;
            if (cur.s < o.max_mem_intv && cur_n - x + 1 >= o.min_seed_len + 1) {
                if (cur.s > 0) emit(x, cur_n, cur);
                x = next_x;
                state = ST_LSTART;
            } else {
                ++j;
            }
        }
    }

    // consume() followed by the common-case transition to the next request, as straight-line
    // code: forward step -> forward step, backward item -> next item, backward row -> next row.
    // Returns true when the next request is ready; false when the rarer bookkeeping of advance()
    // is needed (phase changes, pivot ends, round changes, read end).
    BWAGPU_HD bool step_fast(const Biv &r, const SeedOpt &o, const Biv *r2 = nullptr)
    {
#if SEED_PAIR2 && SEED_ROW_FWD
        // A backward row of one or two entries is handled by ONE paired step, and the entries of the next row are then
        // exactly the results of this step that were kept: they go on as the next request straight from registers.  The
        // stores into the list still happen (the row's end states read it), the load of what was just stored does not —
        // that load, an L2 round trip in front of every row's DRAM round trip, was the kernel's largest single stall
        // (17.7 % of the samples, profiles/r02_ncu_seed_collect_v2_lines.txt).
        if (state == ST_BITEM && p == 0 && np <= (has2 ? 2 : 1)) {
            ++n_ext;
            const bool k1 = consume_item(r, req, cur_n, o);
            bool k2 = false;
            if (has2) { has2 = false; ++n_ext; k2 = consume_item(*r2, req2, cur_n2, o); }
            np = nc;                             // the row is finished (p == np)
            if (np == 0) { state = ST_NEXT; return false; }
            --j;
            if (j < 0 || q[j] > 3) { state = ST_BEND; return false; }
            nc = 0; decided = 0; last_s = -1; p = 0;
            if (k1) { req = r; if (k2) { req2 = *r2; has2 = true; } }        // cur_n / cur_n2 stay: span ends do not change
            else { req = *r2; cur_n = cur_n2; }
            req_a = q[j];
            return true;
        }
#endif
        consume(r, o);
#if SEED_PAIR2
        if (has2) { has2 = false; consume_second(*r2, o); }
#else
        (void)r2;
#endif
        if (state == ST_FWD || state == ST_LFWD) {
            if (j < len) {
                const int a = q[j];
                if (a < 4) { next_x = j + 1; request_fwd(a); return true; }
            }
            return false;
        }
        if (state == ST_BITEM) {
            if (p >= np) {                       // row finished
                np = nc;
                if (np == 0) { state = ST_NEXT; return false; }
                --j;
                if (j < 0 || q[j] > 3) { state = ST_BEND; return false; }
                nc = 0; decided = 0; last_s = -1; p = 0;
            }
            request_items();
            return true;
        }
        return false;
    }
};

// SA look-up of one BWT row: LF-walk to the next sampled row (every 8th), add the steps walked.
// A walk that lands on the sentinel row yields 0 (the steps are discarded), as call_one_step does.
BWAGPU_HD int64_t fm_sa_walk(const FmView &f, int64_t pos, int *steps)
{
    int64_t walked = 0;
    for (;;) {
        if ((pos & 7) == 0) {
            if (steps) *steps = (int)walked;
            return ((int64_t)f.sa_ms_byte[pos >> 3] << 32) + (int64_t)f.sa_ls_word[pos >> 3] + walked;
        }
        FmRec r;
        fm_load_rec(f, pos >> 6, r);
        const int sh = 63 - (int)(pos & 63);
        int b = 4;
        if ((r.bits[0] >> sh) & 1) b = 0;
        else if ((r.bits[1] >> sh) & 1) b = 1;
        else if ((r.bits[2] >> sh) & 1) b = 2;
        else if ((r.bits[3] >> sh) & 1) b = 3;
        if (b == 4) { if (steps) *steps = (int)walked; return 0; }
        const uint64_t m = fm_prefix_mask((int)(pos & 63));
        pos = fm_count(f, b) + (b == 0 ? r.cnt[0] + fm_popc(r.bits[0] & m) : b == 1 ? r.cnt[1] + fm_popc(r.bits[1] & m)
                                  : b == 2 ? r.cnt[2] + fm_popc(r.bits[2] & m) : r.cnt[3] + fm_popc(r.bits[3] & m));
        ++walked;
    }
}

// rows an SMEM contributes: k, k+step, ... below k+s, at most max_occ of them
BWAGPU_HD int fm_sa_rows(int64_t s, int max_occ, int64_t *step_out)
{
    const int64_t step = s > max_occ ? s / max_occ : 1;
    if (step_out) *step_out = step;
    int64_t c = (s + step - 1) / step;
    if (c > max_occ) c = max_occ;
    if (c < 0) c = 0;
    return (int)c;
}
