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

// Device view of the FM-index replica held in HBM.  count[] is the loaded form (file value + 1).
struct FmView {
    const int64_t *cp_occ;       // 8 x 64-bit words per record: cp_count[4], one_hot_bwt_str[4]
    const int8_t *sa_ms_byte;
    const uint32_t *sa_ls_word;
    int64_t count[5];
    int64_t sentinel_index;
    int64_t ref_seq_len;
    // FMA tables (reference all_smem_t / last_smem_t, src/FMI_search.h:99-135); NULL = not loaded
    const uint32_t *fma_all;     // 32 words per 11-mer: last_avail, 10 x {k32,l32,s32}, pad
    const uint32_t *fma_last;    // 4 words per 13-mer: bp|kms<<8|lms<<16|sms<<24, kls, lls, sls
};

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

#ifndef SEED_LD_MODE
#define SEED_LD_MODE 1      // 0: 4x128-bit __ldg; 1: 2x256-bit cached; 2: 256-bit, L1 no_allocate for narrow intervals
#endif

// `deep`: the interval is narrow, i.e. this record is a random touch with no reuse inside the SM
BWAGPU_HD void fm_load_rec(const FmView &f, int64_t block, FmRec &r, bool deep = false)
{
#if defined(__CUDA_ARCH__)
    const int64_t *p = f.cp_occ + block * 8;
#if SEED_LD_MODE == 0
    const uint4 *p4 = reinterpret_cast<const uint4 *>(p);
    uint4 a = __ldg(p4), b = __ldg(p4 + 1), c = __ldg(p4 + 2), d = __ldg(p4 + 3);
    r.cnt[0] = (int64_t)(((uint64_t)a.y << 32) | a.x); r.cnt[1] = (int64_t)(((uint64_t)a.w << 32) | a.z);
    r.cnt[2] = (int64_t)(((uint64_t)b.y << 32) | b.x); r.cnt[3] = (int64_t)(((uint64_t)b.w << 32) | b.z);
    r.bits[0] = ((uint64_t)c.y << 32) | c.x; r.bits[1] = ((uint64_t)c.w << 32) | c.z;
    r.bits[2] = ((uint64_t)d.y << 32) | d.x; r.bits[3] = ((uint64_t)d.w << 32) | d.z;
#else
    // one 64-byte record = two 256-bit loads (sm_100: LDG.256), one per 32-byte sector
    if (SEED_LD_MODE == 2 && deep) {
        asm("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];"
            : "=l"(r.cnt[0]), "=l"(r.cnt[1]), "=l"(r.cnt[2]), "=l"(r.cnt[3]) : "l"(p));
        asm("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];"
            : "=l"(r.bits[0]), "=l"(r.bits[1]), "=l"(r.bits[2]), "=l"(r.bits[3]) : "l"(p + 4));
    } else {
        asm("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];"
            : "=l"(r.cnt[0]), "=l"(r.cnt[1]), "=l"(r.cnt[2]), "=l"(r.cnt[3]) : "l"(p));
        asm("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];"
            : "=l"(r.bits[0]), "=l"(r.bits[1]), "=l"(r.bits[2]), "=l"(r.bits[3]) : "l"(p + 4));
    }
#endif
#else
    (void)deep;
    const int64_t *p = f.cp_occ + block * 8;
    for (int i = 0; i < 4; ++i) { r.cnt[i] = p[i]; r.bits[i] = (uint64_t)p[4 + i]; }
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

// count[a] without dynamic indexing of the kernel-parameter struct (keeps it out of local memory)
BWAGPU_HD int64_t fm_count(const FmView &f, int a)
{
    return a == 0 ? f.count[0] : a == 1 ? f.count[1] : a == 2 ? f.count[2] : a == 3 ? f.count[3] : f.count[4];
}

// backward extension of (k,l,s) by base a
BWAGPU_HD Biv fm_bext(const FmView &f, const Biv &v, int a)
{
    const int64_t sp = v.k, ep = v.k + v.s;
    FmRec rs, re;
    const bool deep = v.s < (1 << 14);
    fm_load_rec(f, sp >> 6, rs, deep);
    if ((ep >> 6) == (sp >> 6)) re = rs; else fm_load_rec(f, ep >> 6, re, deep);
    const uint64_t ms = fm_prefix_mask((int)(sp & 63)), me = fm_prefix_mask((int)(ep & 63));
    int64_t osp[4], sz[4];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        osp[b] = rs.cnt[b] + fm_popc(rs.bits[b] & ms);
        sz[b] = re.cnt[b] + fm_popc(re.bits[b] & me) - osp[b];
    }
    const int64_t sent = (v.k <= f.sentinel_index && ep > f.sentinel_index) ? 1 : 0;
    int64_t l3 = v.l + sent, l2 = l3 + sz[3], l1 = l2 + sz[2], l0 = l1 + sz[1];
    Biv r;
    r.k = fm_count(f, a) + (a == 0 ? osp[0] : a == 1 ? osp[1] : a == 2 ? osp[2] : osp[3]);
    r.s = a == 0 ? sz[0] : a == 1 ? sz[1] : a == 2 ? sz[2] : sz[3];
    r.l = a == 0 ? l0 : a == 1 ? l1 : a == 2 ? l2 : l3;
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
    int64_t n_ext;
    const int64_t *pf_cp;        // CP_OCC base for prefetches (NULL: none)

    BWAGPU_HD void init(ReadStore q_, int len_, int rid_, PrevStore prev_, bwagpu_smem *out_, int cap_)
    {
        q = q_; len = len_; rid = rid_; prev = prev_; out = out_; cap = cap_; n_out = 0;
        prev.reset(len_);
        round = 1; x = 0; min_intv = 1; state = len > 0 ? ST_PIVOT : ST_DONE; n_ext = 0;
        np = base = p = nc = decided = r1_count = r2_idx = 0; last_s = -1; j = next_x = 0;
        cur.k = cur.l = cur.s = 0; cur_n = 0; req = cur; req_a = 0;
        pf_cp = nullptr;
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
                if (p < np) {
                    int n_;
                    prev.get(base + p, req, n_);
                    cur_n = n_;                  // span end of the entry being extended
                    req_a = q[j];
                    return true;
                }
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
            // r extends entry p to the left by q[j]; entries of this row start at m = j + 1
            if (r.s < min_intv) {
                if (!decided && cur_n - (j + 1) + 1 >= o.min_seed_len) {
                    emit(j + 1, cur_n, req);
                    decided = 1;
                }
            } else if (r.s != (int64_t)last_s) {
                last_s = (int32_t)r.s;
                prev.put(base + nc, r, cur_n);
                if (pf_cp && j > 0) fm_prefetch_bext(pf_cp, r.k, r.s);           // extended again in the next row
                ++nc;
                decided = 1;
            }
            ++p;
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
    BWAGPU_HD bool step_fast(const Biv &r, const SeedOpt &o)
    {
        consume(r, o);
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
            int n_;
            prev.get(base + p, req, n_);
            cur_n = n_;
            req_a = q[j];
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
