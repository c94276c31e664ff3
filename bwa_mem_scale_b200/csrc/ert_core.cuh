// ERT (enumerated radix tree) seeding core: the three seeding rounds of mem_kernel1_core_ert as
// per-thread code shared by the CUDA kernels in ert.cu; also compilable as plain host C++ so that
// tests can run the exact same logic without a GPU (tests/hostsim — test scaffolding, not a product
// path).
//
// Reference behaviour reproduced (reference src/ertseeding.cpp):
//   ErtWalker::seed_read        per-read driver of mem_kernel1_core_ert       src/bwamem.cpp:1124-1193
//   get_seeds / reseed / last   get_seeds[_prefix] :2925/:3062, reseed[_prefix] :3200/:3315, last :3425
//   right_extend*               rightExtend :2142, _wlimit :2317, _last :2500
//   left_extend                 leftExtend :1288, _wlimit :1392
//   walk_fwd / walk_last / walk_bwd / walk_fetch*   getNextByteIdx* :609-1287, :1505-1853
//   right_fetch*                rightExtend_fetch_leaves* :1854-2141
//   gather                      leaf_gather / getNextByteIdx_dfs :521-597 (recursion -> explicit stack)
//   check_add*                  check_and_add_smem* :2639-2924
//   child_off / leaf_off        the 768-row code_table / leaf_table :25-430 as closed forms
//
// The index is the reference's own: kmer_offsets[4^15] (8-byte entries: type, 14-bit LEP mask, hit
// count for reseeding, pointer width, byte offset) and mlt_table (per k-mer: 4-byte multi-hit base,
// optional 256-entry 4-mer table, tree nodes, multi-hit lists), both byte for byte as `index -a ert`
// writes them.  Reads are at most 151 bases (the reference asserts len <= READ_LEN, bwamem.cpp:1135).
#pragma once
#include <stdint.h>
#include "../../include/bwagpu.h"

#if defined(__CUDACC__)
#define ERT_HD __host__ __device__
#define ERT_HDI __host__ __device__ __forceinline__
// One out-of-line copy of every building block on the device: fully inlined, the walker is 55 k SASS instructions
// (880 KB) and the kernel stalls on instruction fetch 87 % of the time (profiles/r01_ncu_ert_seed_summary.txt).
#define ERT_FN __host__ __device__ __noinline__
#else
#define ERT_HD
#define ERT_HDI inline
#define ERT_FN
#endif

// test scaffolding may define this to log which k-mer entries a walk touches
#ifndef ERT_KMER_LOAD
#define ERT_KMER_LOAD(ix, h) ((ix).kmer_offsets[h])
#endif

#define ERT_KMER 15
#define ERT_XMER 4
#define ERT_MAX_READ 151
enum { ERT_EMPTY = 0, ERT_LEAF = 1, ERT_UNIFORM = 2, ERT_DIVERGE = 3 };           // src/ertindex.h:12
enum { ERT_INVALID = 0, ERT_SINGLE_HIT_LEAF = 1, ERT_INFREQUENT = 2, ERT_FREQUENT = 3 };   // src/macro.h:216-219

struct ErtView {
    const uint64_t *kmer_offsets;
    const uint8_t *mlt;
    const uint8_t *ref;             // .0123 text, forward + reverse complement
    int64_t l_pac;
    int64_t mlt_bytes;              // size of the blob (bounds checks of the ERT_GUARD build)
};

struct ErtOpt { int min_seed_len, split_len, split_width, max_mem_intv; };

// what the walker keeps of a pushed mem_t (src/ertseeding.h:62-77)
struct ErtMemRec {
    int32_t start, end, hitbeg, hitcount;
    int32_t corr;                   // -1: hits are forward coordinates; >= 0: end_correction of a backward MEM
    int32_t c_pivot, p_pivot, pp_pivot;
};

struct ErtMem {
    int start, end, rc_start, rc_end, hitbeg, hitcount, end_correction;
    int forward, skip_ref_fetch, fetch_leaves, is_multi_hit;
    int c_pivot, p_pivot, pp_pivot;
    ERT_HDI void zero() { start = end = rc_start = rc_end = hitbeg = hitcount = end_correction = 0;
                          forward = skip_ref_fetch = fetch_leaves = is_multi_hit = 0; c_pivot = p_pivot = pp_pivot = 0; }
};

struct ErtHelper { int prevMemStart, prevMemEnd, curr_pivot, prev_pivot, prev_prev_pivot, stop_be, mem_end_limit; };

ERT_HDI int ert_type(uint8_t code, int c) { return (code >> (c << 1)) & 3; }
// children are laid out after the code byte as: pointers of the DIVERGE children (c = 3..0), then the
// 5-byte records of the LEAF children (c = 3..0)
ERT_HDI int ert_child_off(int ptr_width, uint8_t code, int c)
{
    int n = 0;
    for (int d = c + 1; d < 4; ++d) n += ert_type(code, d) == ERT_DIVERGE;
    return ert_type(code, c) == ERT_DIVERGE ? n * ptr_width : 0;
}
ERT_HDI int ert_leaf_off(int ptr_width, uint8_t code, int c)
{
    if (ert_type(code, c) != ERT_LEAF) return 0;
    int nd = 0, nl = 0;
    for (int d = 0; d < 4; ++d) nd += ert_type(code, d) == ERT_DIVERGE;
    for (int d = c + 1; d < 4; ++d) nl += ert_type(code, d) == ERT_LEAF;
    return nd * ptr_width + 5 * nl;
}

struct ErtWalker {
    ErtView ix;
    const uint8_t *fwd;             // the read, one code per byte (>= 4 = N)
    int l_seq, min_seed_len;
    bool rc;                        // read_buf = reverse-complemented read
    int ptr_width, num_hits, limit;
    uint64_t lep[5];
    uint64_t next_lep;
    uint64_t mh_start;
    ErtMemRec *mems; int mem_cap, n_mem;
    int64_t *hits; int hit_cap, n_hit, hit_peak;
    int watch; int64_t watch_val;
    uint32_t last_visited;          // top of the reference's `visited` path (only ever popped once, then the walk ends)
    int64_t n_bytes;                // index + text bytes touched (algorithmic traffic)

    // ---- read access -------------------------------------------------------------------------------
    ERT_HDI int rd(int i) const
    {
        if (i < 0 || i >= l_seq) return 0;                 // the NUL that ends the reference's seq string
        if (!rc) { const int b = fwd[i]; return b < 4 ? b : 4; }
        const int b = fwd[l_seq - 1 - i];
        return b < 4 ? 3 - b : 4;
    }
    ERT_HDI int rdf(int i) const { if (i < 0 || i >= l_seq) return 0; const int b = fwd[i]; return b < 4 ? b : 4; }

    // ---- index access ------------------------------------------------------------------------------
#ifdef ERT_GUARD
    // debug build: every index / text access is bounds-checked; the first violation is recorded, the access skipped
    int bad_site; int64_t bad_val; uint64_t bad_u, bad_v;
    ERT_HDI bool ok_mlt(const uint8_t *p, int n, int site)
    {
        const int64_t o = (int64_t)(p - ix.mlt);
        if (o >= 0 && o + n <= ix.mlt_bytes + 16) return true;
        if (!bad_site) { bad_site = site; bad_val = o; }
        return false;
    }
#else
    ERT_HDI bool ok_mlt(const uint8_t *, int, int) { return true; }
#endif
    ERT_HDI uint64_t ld(const uint8_t *p, int n)
    {
        // byte loads on purpose: an aligned two-word variant returned wrong MEMs / faulted on the device in some builds
        // (bit-exact on the host); not understood, not kept (profiles/r01_ert_experiments.txt)
        n_bytes += n;
        if (!ok_mlt(p, n, 1)) return 0;
        uint64_t v = 0;
        for (int b = 0; b < n; ++b) v |= (uint64_t)p[b] << (8 * b);
#if defined(ERT_LD_CHECK) && defined(ERT_GUARD) && defined(__CUDA_ARCH__)
        {   // experiment: the aligned two-word form next to the byte form; first disagreement is recorded (site 9)
            const unsigned long long a = (unsigned long long)p;
            const uint64_t *q = (const uint64_t *)(a & ~7ull);
            const int sh = (int)(a & 7) * 8;
            uint64_t u = *q >> sh;
            if (sh + 8 * n > 64) u |= *(q + 1) << (64 - sh);
            if (n < 8) u &= (1ull << (8 * n)) - 1;
            if (u != v && !bad_site) { bad_site = 9; bad_val = (int64_t)(p - ix.mlt) | ((int64_t)n << 56); bad_u = u; bad_v = v; }
        }
#endif
        return v;
    }
    ERT_HDI uint8_t ldb(const uint8_t *p) { ++n_bytes; return ok_mlt(p, 1, 2) ? *p : 0; }
    ERT_HDI int refb(int64_t pos)
    {
        ++n_bytes;
#ifdef ERT_GUARD
        if (pos < 0 || pos >= 2 * ix.l_pac) { if (!bad_site) { bad_site = 3; bad_val = pos; } return 0; }
#endif
        return ix.ref[pos];
    }

    ERT_HDI void push_hit(int64_t pos)
    {
        if (n_hit == watch) watch_val = pos;
        if ((unsigned)n_hit < (unsigned)hit_cap) hits[n_hit] = pos;
        ++n_hit;
        if (n_hit > hit_peak) hit_peak = n_hit;
    }
    // The walk only ever reads back the FIRST hit of the MEM it is working on.  That value is kept in a register
    // (watch / watch_val, armed by mark() wherever the reference sets hitbeg = hits->n), so the control flow never
    // depends on what fits the scratch slab: a read that outgrows it still counts its exact peak.
    ERT_HDI int mark() { watch = n_hit; return n_hit; }
    ERT_HDI int64_t hit_at(int i) const { return i == watch ? watch_val : ((unsigned)i < (unsigned)hit_cap ? hits[i] : 0); }
    ERT_HDI void push_mem(const ErtMem &m)
    {
        if (n_mem < mem_cap) {
            ErtMemRec &r = mems[n_mem];
            r.start = m.start; r.end = m.end; r.hitbeg = m.hitbeg; r.hitcount = m.hitcount;
            r.corr = (m.forward || m.fetch_leaves) ? -1 : m.end_correction;
            r.c_pivot = m.c_pivot; r.p_pivot = m.p_pivot; r.pp_pivot = m.pp_pivot;
        }
        ++n_mem;
    }
    ERT_HDI void set_lep(uint64_t bit) { if (bit < 320) lep[bit >> 6] |= 1ULL << (bit & 63); }
    ERT_HDI void clear_lep() { lep[0] = lep[1] = lep[2] = lep[3] = lep[4] = 0; }

    // get_seq (:455-473): window [beg, end) of the text, empty when it bridges the strand boundary
    ERT_HDI int64_t get_seq(int64_t beg, int64_t end, int64_t *len) const
    {
        if (end < beg) { const int64_t t = beg; beg = end; end = t; }
        if (end > (ix.l_pac << 1)) end = ix.l_pac << 1;
        if (beg < 0) beg = 0;
        if (beg >= ix.l_pac || end <= ix.l_pac) *len = end - beg; else *len = 0;
        return beg;
    }

    // getHashKey (:435-453)
    ERT_FN uint32_t hash_key(int index, int keysize, int *end_flag, int *first_n) const
    {
        uint32_t key = 0;
        int len = keysize;
        if (index + keysize > l_seq) { if (end_flag) *end_flag = 1; len = l_seq - index; }
        for (int i = 0; i < len; ++i) {
            const int b = rd(index + i);
            if (b != 4) key |= (uint32_t)b << (i << 1);
            else { *first_n = i; break; }
        }
        return key;
    }

    // getOffsetToChildNode (:485-496): nb is just past the code byte on entry
    ERT_FN void child(const uint8_t *t, uint8_t code, int c, uint64_t &nb)
    {
        const uint64_t start = nb - 1;
        const uint32_t d = (uint32_t)ld(t + nb + ert_child_off(ptr_width, code, c), ptr_width);
        num_hits = (int)(d & 0x3F);
        nb = start + (d >> 6);
    }

    // the LEAF branch shared by most walkers: record at nb (already advanced past the code byte)
    ERT_FN void leaf_all(const uint8_t *t, uint64_t nb, uint8_t code, int c, ErtMem &m)
    {
        nb += ert_leaf_off(ptr_width, code, c);
        const uint64_t leaf = ld(t + nb, 5);
        if (leaf & 1) {
            nb = mh_start + (leaf >> 1);
            num_hits = (int)ld(t + nb, 2);
            nb += 2;
            m.hitcount += num_hits;
            for (int k = 0; k < num_hits; ++k, nb += 5) push_hit((int64_t)(ld(t + nb, 5) >> 1));
        } else {
            num_hits = 1;
            m.hitcount += 1;
            push_hit((int64_t)(leaf >> 1));
        }
    }
    // the variant used under a hit limit: count first, then push only when num_hits >= limit
    ERT_FN bool leaf_limited(const uint8_t *t, uint64_t nb, uint8_t code, int c, ErtMem &m)
    {
        nb += ert_leaf_off(ptr_width, code, c);
        const uint64_t leaf = ld(t + nb, 5);
        if (leaf & 1) { nb = mh_start + (leaf >> 1); num_hits = (int)ld(t + nb, 2); nb += 2; }
        else num_hits = 1;
        if (num_hits < limit) return false;
        m.hitcount += num_hits;
        for (int k = 0; k < num_hits; ++k, nb += 5) push_hit((int64_t)(ld(t + nb, 5) >> 1));
        return true;
    }

    // UNIFORM run at nb (past the code byte): returns the number of leading bases matching the read at i
    ERT_FN int uniform(const uint8_t *t, uint64_t &nb, int i, int *count)
    {
        const int cnt = ldb(t + nb); ++nb;
        const int nbytes = (cnt * 2 + 7) >> 3;
        int j;
        for (j = 0; j < cnt; ++j) {
            if (i + j >= l_seq) break;
            const int b = rd(i + j);
            if (b == 4) break;
            const int u = ((ok_mlt(t + nb + (j >> 2), 1, 4) ? t[nb + (j >> 2)] : 0) >> ((~j & 3) << 1)) & 3;
            if (3 - b != u) break;
        }
        n_bytes += nbytes;
        nb += nbytes;
        *count = cnt;
        return j;
    }

    // leaf_gather + getNextByteIdx_dfs (:521-597)
    ERT_FN void gather(const uint8_t *t, uint64_t start, ErtMem &m)
    {
        uint32_t st[176];
        int top = 0;
        st[top++] = (uint32_t)start << 3;
        while (top > 0) {
            const uint32_t fr = st[top - 1];
            const int k = fr & 7;
            if (k == 4) { --top; continue; }
            st[top - 1] = fr + 1;
            uint64_t nb = fr >> 3;
            const int c = 3 - k;
            m.skip_ref_fetch = 1;
            const uint8_t code = ldb(t + nb); ++nb;
            const int cc = ert_type(code, c);
            if (cc == ERT_LEAF) leaf_all(t, nb, code, c, m);
            else if (cc == ERT_UNIFORM) {
                const int cnt = ldb(t + nb); ++nb;
                nb += (cnt * 2 + 7) >> 3;
                if (top < 176) st[top++] = (uint32_t)nb << 3;
            } else if (cc == ERT_DIVERGE) {
                child(t, code, c, nb);
                if (top < 176) st[top++] = (uint32_t)nb << 3;
            }
        }
    }

    // ---- forward tree walk: getNextByteIdx (:836) and getNextByteIdx_wlimit (:989) --------------------
    ERT_FN void walk_fwd(bool WL, const uint8_t *t, uint64_t nb, int &i, ErtMem &m)
    {
        for (;;) {
            const uint64_t parent = nb;
            int c = 0, cc = ERT_EMPTY;
            uint8_t code = 0;
            if (rd(i) != 4) { c = 3 - rd(i); code = ldb(t + nb); ++nb; cc = ert_type(code, c); }
            if (cc == ERT_EMPTY) {
                if (m.start == 0 && i >= min_seed_len) gather(t, parent, m);
                set_lep(next_lep); next_lep += 1;
                return;
            }
            if (cc == ERT_LEAF) {
                if (!WL) { leaf_all(t, nb, code, c, m); set_lep(next_lep); next_lep += 1; i += 1; return; }
                if (leaf_limited(t, nb, code, c, m)) i += 1;
                else if (m.start == 0 && i >= min_seed_len) gather(t, last_visited, m);
                set_lep(next_lep); next_lep += 1;
                return;
            }
            if (cc == ERT_UNIFORM) {
                int cnt;
                const int j = uniform(t, nb, i, &cnt);
                next_lep += j; i += j;
                if (j == cnt) {
                    if (i == l_seq) { if (m.start == 0) gather(t, nb, m); set_lep(next_lep); return; }
                    if (i < l_seq) continue;
                    return;
                }
                if (m.start == 0 && i >= min_seed_len) gather(t, nb, m);
                set_lep(next_lep);
                return;
            }
            child(t, code, c, nb);
            set_lep(next_lep); next_lep += 1;
            if (!WL || num_hits == 0 || num_hits >= limit) {
                if (WL) last_visited = (uint32_t)nb;
                i += 1;
                if (i < l_seq) continue;
                if (m.start == 0) gather(t, nb, m);
                set_lep(next_lep); next_lep += 1;
                return;
            }
            if (m.start == 0 && i >= min_seed_len) gather(t, last_visited, m);
            return;
        }
    }

    // getNextByteIdx_last (:1168)
    ERT_FN void walk_last(const uint8_t *t, uint64_t nb, int &i, ErtMem &m)
    {
        for (;;) {
            int c = 0, cc = ERT_EMPTY;
            uint8_t code = 0;
            if (rd(i) != 4) { c = 3 - rd(i); code = ldb(t + nb); ++nb; cc = ert_type(code, c); }
            if (cc == ERT_EMPTY) { i += 1; return; }
            if (cc == ERT_LEAF) { leaf_all(t, nb, code, c, m); i += 1; return; }
            if (cc == ERT_UNIFORM) {
                int cnt;
                const int j = uniform(t, nb, i, &cnt);
                i += j;
                const int len = i - m.start;
                if (num_hits > 0 && num_hits < limit && len >= min_seed_len + 1) {
                    gather(t, nb, m);
                    i = m.start + (min_seed_len + 1);
                    return;
                }
                if (j == cnt) { if (i < l_seq) continue; return; }
                i += 1;
                return;
            }
            child(t, code, c, nb);
            i += 1;
            const int len = i - m.start;
            if (num_hits > 0 && num_hits < limit && len >= min_seed_len + 1) { gather(t, nb, m); return; }
            if (i < l_seq) continue;
            return;
        }
    }

    // getNextByteIdx_backward (:609) and _backward_wlimit (:718)
    ERT_FN void walk_bwd(bool WL, const uint8_t *t, uint64_t nb, int &i, ErtMem &m)
    {
        for (;;) {
            int c = 0, cc = ERT_EMPTY;
            uint8_t code = 0;
            if (rd(i) != 4) { c = 3 - rd(i); code = ldb(t + nb); ++nb; cc = ert_type(code, c); }
            if (cc == ERT_EMPTY) { m.rc_end = i; m.fetch_leaves = 1; return; }
            if (cc == ERT_LEAF) {
                if (!WL) {
                    i += 1; m.rc_end = i;
                    uint64_t lb = nb + ert_leaf_off(ptr_width, code, c);
                    const uint64_t leaf = ld(t + lb, 5);
                    if (leaf & 1) {
                        lb = mh_start + (leaf >> 1);
                        num_hits = (int)ld(t + lb, 2); lb += 2;
                        m.hitcount += num_hits;
                        for (int k = 0; k < num_hits; ++k, lb += 5) push_hit((int64_t)(ld(t + lb, 5) >> 1));
                        m.fetch_leaves = 1;          // re-fetched later in forward order
                    } else { m.hitcount += 1; push_hit((int64_t)(leaf >> 1)); }
                    return;
                }
                uint64_t lb = nb + ert_leaf_off(ptr_width, code, c);
                const uint64_t leaf = ld(t + lb, 5);
                if (leaf & 1) {
                    lb = mh_start + (leaf >> 1);
                    num_hits = (int)ld(t + lb, 2); lb += 2;
                    if (num_hits >= limit) {
                        m.hitcount += num_hits;
                        for (int k = 0; k < num_hits; ++k, lb += 5) push_hit((int64_t)(ld(t + lb, 5) >> 1));
                        i += 1;
                    }
                }
                m.fetch_leaves = 1; m.rc_end = i;
                return;
            }
            if (cc == ERT_UNIFORM) {
                int cnt;
                const int j = uniform(t, nb, i, &cnt);
                i += j;
                if (j == cnt) {
                    if (i < l_seq) continue;
                    m.rc_end = i; if (WL) m.fetch_leaves = 1;
                    return;
                }
                m.rc_end = i; m.fetch_leaves = 1;
                return;
            }
            if (WL) num_hits = 0;
            child(t, code, c, nb);
            if (!WL || num_hits == 0 || num_hits >= limit) {
                i += 1;
                if (i < l_seq) continue;
                m.rc_end = i; if (WL) m.fetch_leaves = 1;
                return;
            }
            m.rc_end = i; m.fetch_leaves = 1;
            return;
        }
    }

    // getNextByteIdx_fetch_leaves_prefix (:1644) and _prefix_reseed (:1505)
    ERT_FN void walk_fetch_prefix(bool RS, const uint8_t *t, uint64_t nb, int i, ErtMem &m)
    {
        for (;;) {
            const uint64_t parent = nb;
            const int c = 3 - rd(i);
            const uint8_t code = ldb(t + nb); ++nb;
            const int cc = ert_type(code, c & 3);
            if (cc == ERT_EMPTY) {
                m.end = i;
                if (m.end - m.start >= min_seed_len) gather(t, parent, m);
                return;
            }
            if (cc == ERT_LEAF) {
                if (!RS) { leaf_all(t, nb, code, c, m); i += 1; m.end = i; return; }
                if (leaf_limited(t, nb, code, c, m)) { i += 1; m.end = i; m.is_multi_hit = 1; }
                else { m.end = i; if (m.end - m.start >= min_seed_len) gather(t, last_visited, m); }
                return;
            }
            if (cc == ERT_UNIFORM) {
                int cnt;
                const int j = uniform(t, nb, i, &cnt);
                i += j;
                if (j == cnt && i < l_seq) continue;
                m.end = i;
                if (m.end - m.start >= min_seed_len) gather(t, nb, m);
                return;
            }
            num_hits = 0;
            child(t, code, c, nb);
            if (!RS || num_hits == 0 || num_hits >= limit) {
                if (RS) last_visited = (uint32_t)nb;
                i += 1;
                if (i < l_seq) continue;
                m.end = i;
                if (m.end - m.start >= min_seed_len) gather(t, nb, m);
                return;
            }
            m.end = i;
            if (m.end - m.start >= min_seed_len) gather(t, last_visited, m);
            return;
        }
    }

    // getNextByteIdx_fetch_leaves (:1759): re-walk of a backward MEM, bounded by mem->end
    ERT_FN void walk_fetch(const uint8_t *t, uint64_t nb, int i, ErtMem &m)
    {
        for (;;) {
            const int c = 3 - rd(i);
            const uint8_t code = ldb(t + nb); ++nb;
            const int cc = ert_type(code, c & 3);
            if (cc == ERT_LEAF) { leaf_all(t, nb, code, c, m); return; }
            if (cc == ERT_UNIFORM) {
                int cnt;
                const int j = uniform(t, nb, i, &cnt);
                i += j;
                if (j == cnt && i < m.end) continue;
                gather(t, nb, m);
                return;
            }
            if (cc == ERT_DIVERGE) {
                num_hits = 0;
                child(t, code, c, nb);
                i += 1;
                if (i < m.end) continue;
                gather(t, nb, m);
            }
            return;
        }
    }

    // ---- k-mer look-up shared by the extenders -------------------------------------------------------
    ERT_HDI void or_lep(uint64_t lep_data, int i)
    {
        const int w = i >> 6, b = i & 63;
        if (w < 5) lep[w] |= lep_data << b;
        if (b > 64 - ERT_KMER && w + 1 < 5) lep[w + 1] |= lep_data >> (64 - b);
    }
    ERT_HDI int width_of(uint64_t e) const { const int w = (int)((e >> 22) & 3); return w == 0 ? 4 : w; }

    // rightExtend (:2142) and rightExtend_wlimit (:2317)
    ERT_FN void right_extend(bool WL, int &i, ErtMem &m)
    {
        int flag = 0, first_n = -1;
        const uint32_t h = hash_key(i, ERT_KMER, &flag, &first_n);
        const uint64_t e = ERT_KMER_LOAD(ix, h); n_bytes += 8;
        int code = (int)(e & 3);
        const uint64_t start_addr = e >> 24;
        mh_start = 0;
        ptr_width = width_of(e);
        if (WL) num_hits = (int)((e >> 17) & 0x1F);
        or_lep((e >> 2) & 0x3FFF, i);
        next_lep = (uint64_t)(i + ERT_KMER - 1);
        if (first_n != -1) {
            if (i != 0) { next_lep = (uint64_t)(i + first_n - 1); set_lep(next_lep); }
            i += first_n;
            return;
        }
        if (flag) { next_lep = (uint64_t)(l_seq - 1); i = l_seq; set_lep(next_lep); return; }
        const uint8_t *t = ix.mlt + start_addr;
        if (code == ERT_INVALID) { i += ERT_KMER + ERT_XMER; return; }
        if (code == ERT_SINGLE_HIT_LEAF) {
            if (WL) { i += ERT_KMER + ERT_XMER; return; }
            m.hitcount += 1;
            push_hit((int64_t)(ld(t + 1, 5) >> 1));
            i += ERT_KMER;
            return;
        }
        if (code == ERT_INFREQUENT) {
            i += ERT_KMER;
            if (WL && !(num_hits == 0 || num_hits >= limit)) return;
            if (i < l_seq) {
                if (WL) last_visited = 0;
                mh_start = ld(t, 4);
                walk_fwd(WL, t, 4, i, m);
            } else { set_lep(next_lep); next_lep += 1; }
            return;
        }
        // FREQUENT: one more look-up in the k-mer's 4-mer table
        i += ERT_KMER;
        flag = 0;
        const uint32_t hx = hash_key(i, ERT_XMER, &flag, &first_n);
        mh_start = ld(t, 4);
        const uint64_t x = ld(t + 4 + ((uint64_t)hx << 3), 8);
        code = (int)(x & 3);
        const uint64_t lep_data = (x >> 2) & 0xF;
        const uint64_t ptr = x >> 24;
        if (WL) num_hits = (int)((x >> 17) & 0x1F);
        const int xlen = (l_seq - i > ERT_XMER) ? ERT_XMER : l_seq - i;
        for (int k = 0; k < xlen; ++k) { if ((lep_data >> k) & 1) set_lep(next_lep); next_lep++; }
        if (first_n != -1) { next_lep = (uint64_t)(i + first_n - 1); set_lep(next_lep); i += first_n; return; }
        if (flag) { next_lep = (uint64_t)(l_seq - 1); i = l_seq; set_lep(next_lep); return; }
        if (code == ERT_INVALID) { i += ERT_XMER; return; }
        if (code == ERT_SINGLE_HIT_LEAF) {
            if (!WL) { m.hitcount += 1; push_hit((int64_t)(ld(t + ptr + 1, 5) >> 1)); }
            i += ERT_XMER;
            return;
        }
        i += ERT_XMER;
        if (WL && !(num_hits == 0 || num_hits >= limit)) return;
        if (i < l_seq) {
            if (WL) last_visited = (uint32_t)ptr;
            walk_fwd(WL, t, ptr, i, m);
        } else { set_lep(next_lep); next_lep += 1; }
    }

    // rightExtend_last (:2500)
    ERT_FN void right_extend_last(int &i, ErtMem &m)
    {
        int flag = 0, first_n = -1;
        const uint32_t h = hash_key(i, ERT_KMER, &flag, &first_n);
        if (first_n != -1) { i += first_n + 1; return; }
        if (flag) { i = l_seq; return; }
        const uint64_t e = ERT_KMER_LOAD(ix, h); n_bytes += 8;
        int code = (int)(e & 3);
        const uint8_t *t = ix.mlt + (e >> 24);
        mh_start = 0;
        ptr_width = width_of(e);
        num_hits = (int)((e >> 17) & 0x1F);
        if (code == ERT_INVALID) { i += ERT_KMER; return; }
        if (code == ERT_SINGLE_HIT_LEAF) {
            m.hitcount += 1;
            push_hit((int64_t)(ld(t + 1, 5) >> 1));
            i += ERT_KMER;
            return;
        }
        if (code == ERT_INFREQUENT) {
            i += ERT_KMER;
            if (i < l_seq) { mh_start = ld(t, 4); walk_last(t, 4, i, m); }
            return;
        }
        i += ERT_KMER;
        flag = 0;
        const uint32_t hx = hash_key(i, ERT_XMER, &flag, &first_n);
        if (first_n != -1) { i += first_n + 1; return; }
        if (flag) { i = l_seq; return; }
        mh_start = ld(t, 4);
        const uint64_t x = ld(t + 4 + ((uint64_t)hx << 3), 8);
        code = (int)(x & 3);
        const uint64_t ptr = x >> 24;
        num_hits = (int)((x >> 17) & 0x1F);
        if (code == ERT_INVALID) { i += ERT_XMER; return; }
        if (code == ERT_SINGLE_HIT_LEAF) {
            m.hitcount += 1;
            push_hit((int64_t)(ld(t + ptr + 1, 5) >> 1));
            i += ERT_XMER;
            return;
        }
        i += ERT_XMER;
        if (num_hits == 0 || num_hits >= limit || (i - m.start) < (min_seed_len + 1)) {
            if (i < l_seq) walk_last(t, ptr, i, m);
        } else gather(t, ptr, m);
    }

    // leftExtend (:1288) and leftExtend_wlimit (:1392); read_buf is the reverse-complemented read
    ERT_FN void left_extend(bool WL, int &i, ErtMem &m)
    {
        int first_n = -1;
        const uint32_t h = hash_key(i, ERT_KMER, nullptr, &first_n);
        if (first_n != -1) { i += ERT_KMER + ERT_XMER; m.rc_end = i; return; }
        const uint64_t e = ERT_KMER_LOAD(ix, h); n_bytes += 8;
        int code = (int)(e & 3);
        const uint8_t *t = ix.mlt + (e >> 24);
        ptr_width = width_of(e);
        if (WL) num_hits = (int)((e >> 17) & 0x1F);
        if (code == ERT_INVALID) { i += ERT_KMER + ERT_XMER; m.rc_end = i; return; }
        if (code == ERT_SINGLE_HIT_LEAF) {
            if (WL) { i += ERT_KMER + ERT_XMER; m.rc_end = i; return; }
            m.hitcount += 1;
            push_hit((int64_t)(ld(t + 1, 5) >> 1));
            i += ERT_KMER; m.rc_end = i;
            return;
        }
        if (code == ERT_INFREQUENT) {
            i += ERT_KMER;
            if (WL && !(num_hits == 0 || num_hits >= limit)) { m.rc_end = i; return; }
            if (i < l_seq) { mh_start = ld(t, 4); walk_bwd(WL, t, 4, i, m); }
            else { m.rc_end = i; if (WL) m.fetch_leaves = 1; }
            return;
        }
        i += ERT_KMER;
        const uint32_t hx = hash_key(i, ERT_XMER, nullptr, &first_n);
        mh_start = ld(t, 4);
        const uint64_t x = ld(t + 4 + ((uint64_t)hx << 3), 8);
        code = (int)(x & 3);
        const uint64_t ptr = x >> 24;
        if (WL) num_hits = (int)((x >> 17) & 0x1F);
        if (first_n != -1) { i += ERT_XMER; m.rc_end = i; return; }
        if (code == ERT_INVALID) { i += ERT_XMER; m.rc_end = i; return; }
        if (code == ERT_SINGLE_HIT_LEAF) {
            if (!WL) { m.hitcount += 1; push_hit((int64_t)(ld(t + ptr + 1, 5) >> 1)); }
            i += ERT_XMER; m.rc_end = i;
            return;
        }
        i += ERT_XMER;
        if (WL && !(num_hits == 0 || num_hits >= limit)) { m.rc_end = i; return; }
        if (i < l_seq) walk_bwd(WL, t, ptr, i, m);
        else { m.rc_end = i; if (WL) m.fetch_leaves = 1; }
    }

    // rightExtend_fetch_leaves_prefix (:1967) and _prefix_reseed (:1854); read_buf is the forward read
    ERT_FN void right_fetch_prefix(bool RS, ErtMem &m)
    {
        int flag = 0, first_n = -1;
        int i = m.start;
        const uint32_t h = hash_key(i, ERT_KMER, &flag, &first_n);
        const uint64_t e = ERT_KMER_LOAD(ix, h); n_bytes += 8;
        int code = (int)(e & 3);
        const uint8_t *t = ix.mlt + (e >> 24);
        mh_start = 0;
        ptr_width = width_of(e);
        if (RS) num_hits = (int)((e >> 17) & 0x1F);
        if (code == ERT_SINGLE_HIT_LEAF) {
            if (RS) { m.end = i; return; }
            m.hitcount += 1;
            push_hit((int64_t)(ld(t + 1, 5) >> 1));
            i += ERT_KMER; m.end = i;
            return;
        }
        if (code == ERT_INFREQUENT) {
            if (RS && !(num_hits == 0 || num_hits >= limit)) { m.end = i; return; }
            i += ERT_KMER;
            mh_start = ld(t, 4);
            if (i < l_seq) { if (RS) last_visited = 4; walk_fetch_prefix(RS, t, 4, i, m); }
            else { m.end = i; if (m.end - m.start >= min_seed_len) gather(t, 4, m); }
            return;
        }
        if (code != ERT_FREQUENT) return;
        const uint32_t hx = hash_key(i + ERT_KMER, ERT_XMER, nullptr, &first_n);
        mh_start = ld(t, 4);
        const uint64_t x = ld(t + 4 + ((uint64_t)hx << 3), 8);
        code = (int)(x & 3);
        const uint64_t ptr = x >> 24;
        if (RS) num_hits = (int)((x >> 17) & 0x1F);
        if (code == ERT_INVALID) { m.end = i; return; }
        if (code == ERT_SINGLE_HIT_LEAF) {
            if (RS) { m.end = i; return; }
            m.hitcount += 1;
            push_hit((int64_t)(ld(t + ptr + 1, 5) >> 1));
            i += ERT_KMER + ERT_XMER; m.end = i;
            return;
        }
        if (RS && !(num_hits == 0 || num_hits >= limit)) { m.end = i; return; }
        i += ERT_KMER + ERT_XMER;
        if (i < l_seq) { if (RS) last_visited = (uint32_t)ptr; walk_fetch_prefix(RS, t, ptr, i, m); }
        else { m.end = i; if (m.end - m.start >= min_seed_len) gather(t, ptr, m); }
    }

    // rightExtend_fetch_leaves (:2071)
    ERT_FN void right_fetch(ErtMem &m)
    {
        int flag = 0, first_n = -1;
        int i = m.start;
        const int end = m.end;
        const uint32_t h = hash_key(i, ERT_KMER, &flag, &first_n);
        const uint64_t e = ERT_KMER_LOAD(ix, h); n_bytes += 8;
        int code = (int)(e & 3);
        const uint8_t *t = ix.mlt + (e >> 24);
        mh_start = 0;
        ptr_width = width_of(e);
        if (code == ERT_INFREQUENT) {
            i += ERT_KMER;
            mh_start = ld(t, 4);
            if (i < end) walk_fetch(t, 4, i, m); else gather(t, 4, m);
        } else if (code == ERT_FREQUENT) {
            i += ERT_KMER;
            const uint32_t hx = hash_key(i, ERT_XMER, nullptr, &first_n);
            mh_start = ld(t, 4);
            const uint64_t x = ld(t + 4 + ((uint64_t)hx << 3), 8);
            const uint64_t ptr = x >> 24;
            i += ERT_XMER;
            if (i < end) walk_fetch(t, ptr, i, m); else gather(t, ptr, m);
        }
    }

    // ---- MEM bookkeeping -----------------------------------------------------------------------------
    ERT_HDI int init_mem(ErtMem &m, int j) const                              // :2609
    {
        const int bit = (int)((lep[j >> 6] >> (j & 63)) & 1);
        m.end = j + 1;
        m.rc_start = l_seq - j - 1;
        m.rc_end = m.rc_start;
        m.skip_ref_fetch = 0; m.forward = 0; m.fetch_leaves = 0;
        m.hitbeg = m.hitcount = 0; m.end_correction = 0; m.is_multi_hit = 0;
        return bit && j >= min_seed_len - 1;
    }

    // extension of a forward MEM against the text (shared head of get_seeds / reseed, :2945-2968)
    ERT_FN void extend_right_on_text(int &i, const ErtMem &rm)
    {
        int64_t len;
        const int64_t h0 = hit_at(rm.hitbeg);
        const int64_t beg = get_seq(h0 + i - rm.start, h0 + l_seq - rm.start, &len);
        int m, n = 0;
        for (m = 0; m < len; ++m) {
            if (refb(beg + m) == rdf(i + m)) ++n;
            else { set_lep((uint64_t)(i + m - 1)); break; }
        }
        if (m == len) set_lep((uint64_t)(i + m - 1));
        i += n;
    }

    // the two text comparisons that open check_and_add_smem_prefix[_reseed] (:2641-2673)
    ERT_FN void lmem_fix_prefix(ErtMem &m)
    {
        m.start = l_seq - m.rc_end;
        const int lmemLen = m.end - m.start;
        if (m.hitcount > 0 && !m.skip_ref_fetch) {
            int64_t len;
            const int64_t h0 = hit_at(m.hitbeg);
            int64_t beg = get_seq(h0 - m.rc_start, h0, &len);
            int n = 0;
            for (int k = 1; k <= len; ++k) {
                if (refb(beg + m.rc_start - k) == rd(m.rc_start - k)) ++n; else break;
            }
            m.end += n; m.end_correction += n;
            const int64_t s2 = h0 + lmemLen;
            beg = get_seq(s2, s2 + m.start, &len);
            n = 0;
            for (int k = 0; k < len; ++k) {
                if (refb(beg + k) == rd(m.rc_end + k)) ++n; else break;
            }
            m.start -= n;
        }
    }

    // check_and_add_smem_prefix (:2756) and _prefix_reseed (:2639)
    ERT_FN int check_add_prefix(bool RS, ErtMem &m, const ErtHelper &sh)
    {
        lmem_fix_prefix(m);
        const int lmemLen = m.end - m.start;
        int next_be = m.end;
        if (m.hitcount == 1) {
            if (lmemLen >= min_seed_len) push_mem(m);
            else next_be += min_seed_len - lmemLen;
        } else if (m.fetch_leaves && m.start <= l_seq - min_seed_len) {
            n_hit -= m.hitcount;
            m.hitbeg = mark();
            m.hitcount = 0;
            rc = false;
            right_fetch_prefix(RS, m);
            rc = true;
            int rmemLen = m.end - m.start;
            next_be = m.end;
            if (m.hitcount > 0) {
                if (!RS || m.is_multi_hit) {
                    int64_t len;
                    const int64_t h0 = hit_at(m.hitbeg);
                    const int64_t beg = get_seq(h0 + rmemLen, h0 + l_seq - m.start, &len);
                    int n = 0;
                    for (int k = 0; k < len; ++k) {
                        if (refb(beg + k) == rdf(m.end + k)) ++n; else break;
                    }
                    m.end += n;
                    rmemLen = m.end - m.start;
                    next_be = m.end;
                }
                if (rmemLen >= min_seed_len && (!RS || m.end <= sh.mem_end_limit)) push_mem(m);
                else next_be += min_seed_len - rmemLen;
            } else next_be += min_seed_len - rmemLen;
        } else {
            if (!RS || lmemLen <= min_seed_len) next_be += min_seed_len - lmemLen;
        }
        return next_be;
    }

    // check_and_add_smem (:2867)
    ERT_FN void check_add(bool WL, ErtMem &m, ErtHelper &sh)
    {
        m.start = l_seq - m.rc_end;
        int lmemLen = m.end - m.start;
        if (m.hitcount > 0 && !m.skip_ref_fetch) {
            int64_t len;
            const int64_t s = hit_at(m.hitbeg) + lmemLen;
            const int64_t beg = get_seq(s, s + m.start, &len);
            int n = 0;
            for (int k = 0; k < len; ++k) {
                if (refb(beg + k) == rd(m.rc_end + k)) ++n; else break;
            }
            m.start -= n;
        }
        lmemLen = m.end - m.start;
        if (lmemLen >= min_seed_len && (m.start < sh.prevMemStart || m.end > sh.prevMemEnd)) {
            if (m.fetch_leaves) {
                n_hit -= m.hitcount;
                m.hitbeg = mark();
                m.hitcount = 0;
                rc = false;
                right_fetch(m);
                rc = true;
            }
            if (m.hitcount > 0) {
                m.c_pivot = sh.curr_pivot; m.p_pivot = sh.prev_pivot; m.pp_pivot = sh.prev_prev_pivot;
                push_mem(m);
                if (m.start <= sh.prev_pivot + 1) sh.stop_be = 1;
            }
            sh.prevMemStart = m.start;
            sh.prevMemEnd = m.end;
        }
    }

    // backward extensions of one forward MEM: the LEP loops of get_seeds[_prefix] / reseed[_prefix]
    ERT_FN void backward_pass(bool PREFIX, bool WL, const ErtMem &rm, ErtHelper &sh, int &i)
    {
        sh.stop_be = 0;
        const int min_j = (rm.start > min_seed_len) ? (rm.start - 1) : (min_seed_len - 1);
        if (PREFIX) {
            const int max_j = rm.end - 1;
            int j = min_j;
            sh.prev_pivot = rm.start;
            if (WL) sh.mem_end_limit = rm.end;
            while (j <= max_j) {
                ErtMem m; m.zero();
                const int valid = init_mem(m, j);
                m.hitbeg = mark();
                int next_j = j + 1;
                if (valid && j + 1 >= min_seed_len) {
                    int rc_i = l_seq - (j + 1);
                    rc = true;
                    left_extend(WL, rc_i, m);
                    next_j = check_add_prefix(WL, m, sh);
                }
                j = next_j;
                if (!WL && m.end > i) i = m.end;
            }
        } else {
            int j = rm.end - 1;
            while (j >= min_j) {
                ErtMem m; m.zero();
                const int valid = init_mem(m, j);
                m.hitbeg = mark();
                if (valid && j + 1 >= min_seed_len) {
                    int rc_i = l_seq - (j + 1);
                    rc = true;
                    left_extend(WL, rc_i, m);
                    check_add(WL, m, sh);
                    if (sh.stop_be) break;
                }
                j -= 1;
            }
        }
    }

    // get_seeds_prefix (:2925) / get_seeds (:3062)
    ERT_FN void get_seeds(bool PREFIX)
    {
        ErtHelper sh = {l_seq, 0, 0, -1, -1, 0, 0};
        int i = 0;
        clear_lep();
        while (i < l_seq) {
            ErtMem rm; rm.zero();
            rm.start = i; rm.forward = 1; rm.hitbeg = mark();
            sh.curr_pivot = rm.start;
            rc = false;
            right_extend(false, i, rm);
            if (rm.hitcount > 0 && !rm.skip_ref_fetch) extend_right_on_text(i, rm);
            rm.end = i;
            if (rm.start == 0) {
                if (rm.end - rm.start >= min_seed_len) {
                    if (rm.hitcount > 0) {
                        rm.c_pivot = sh.curr_pivot; rm.p_pivot = sh.prev_pivot; rm.pp_pivot = sh.prev_prev_pivot;
                        push_mem(rm);
                    }
                } else n_hit -= rm.hitcount;
                clear_lep();
            } else {
                n_hit -= rm.hitcount;
                backward_pass(PREFIX, false, rm, sh, i);
            }
            rc = false;
            while (i < l_seq && rd(i) == 4) ++i;
            while (i < l_seq && (i - rm.start) < min_seed_len) {
                if (rd(i) == 4) { ++i; break; }
                ++i;
            }
            sh.prev_prev_pivot = sh.prev_pivot;
            sh.prev_pivot = rm.start;
            clear_lep();
        }
    }

    // reseed_prefix (:3200) / reseed (:3315)
    ERT_FN void reseed(bool PREFIX, int start, int lim, int c_pivot, int p_pivot, int pp_pivot)
    {
        ErtHelper sh = {l_seq, 0, 0, 0, 0, 0, 0};
        int i = start;
        clear_lep();
        ErtMem rm; rm.zero();
        rm.start = i; rm.forward = 1; rm.hitbeg = mark();
        sh.prev_pivot = (rm.start >= c_pivot) ? p_pivot : pp_pivot;
        rc = false;
        limit = lim;
        right_extend(true, i, rm);
        if (rm.hitcount > 0 && !rm.skip_ref_fetch) extend_right_on_text(i, rm);
        rm.end = i;
        if (rm.start == 0) {
            if (rm.end - rm.start >= min_seed_len) { if (rm.hitcount > 0) push_mem(rm); }
            else n_hit -= rm.hitcount;
            clear_lep();
        } else {
            n_hit -= rm.hitcount;
            backward_pass(PREFIX, true, rm, sh, i);
        }
    }

    // last (:3425)
    ERT_FN void last(int lim)
    {
        int i = 0;
        const int minSeedLen = (uint8_t)(min_seed_len + 1);
        limit = lim;
        while (i < l_seq) {
            ErtMem rm; rm.zero();
            rm.start = i; rm.forward = 1; rm.hitbeg = mark();
            rc = false;
            right_extend_last(i, rm);
            if (rm.hitcount > 0 && !rm.skip_ref_fetch) {
                int64_t len;
                const int64_t h0 = hit_at(rm.hitbeg);
                const int64_t beg = get_seq(h0 + i - rm.start, h0 + l_seq - rm.start, &len);
                int n = 0;
                for (int m = 0; m < len; ++m) {
                    const int seedLen = (i + m) - rm.start;
                    if (!(seedLen < minSeedLen || rm.hitcount >= limit)) break;
                    if (refb(beg + m) == rdf(i + m)) ++n;
                    else { ++i; n_hit -= rm.hitcount; rm.hitcount = 0; break; }
                }
                i += n;
            }
            rm.end = i;
            if (rm.end - rm.start >= minSeedLen && rm.hitcount > 0 && rm.hitcount < limit) push_mem(rm);
            else n_hit -= rm.hitcount;
            if (rd(i - 1) != 4) {
                while (i < l_seq && (i - rm.start) < minSeedLen) {
                    if (rd(i) == 4) { ++i; break; }
                    ++i;
                }
            }
        }
    }

    // the per-read body of mem_kernel1_core_ert (src/bwamem.cpp:1124-1191), before the sort
    ERT_HD void seed_read(const ErtOpt &o)
    {
        n_mem = 0; n_hit = 0; hit_peak = 0; n_bytes = 0; watch = -1; watch_val = 0;
#ifdef ERT_GUARD
        bad_site = 0; bad_val = 0;
#endif
        num_hits = 0; limit = 0; ptr_width = 2; mh_start = 0; next_lep = 0; last_visited = 0;
        min_seed_len = o.min_seed_len;
        bool has_n = false;
        for (int i = 0; i < l_seq; ++i) has_n |= fwd[i] >= 4;
        get_seeds(!has_n);
        const int old_n = n_mem < mem_cap ? n_mem : mem_cap;
        for (int k = 0; k < old_n; ++k) {
            const ErtMemRec r = mems[k];
            if ((r.end - r.start) < o.split_len || r.hitcount > o.split_width) continue;
            reseed(!has_n, (r.start + r.end) >> 1, r.hitcount + 1, r.c_pivot, r.p_pivot, r.pp_pivot);
        }
        last(o.max_mem_intv);
    }
};

// ks_introsort(mem_smem_sort_lt) over the read's MEMs (src/bwamem.cpp:1193; smem_lt_2 :72-74): keys are
// (start << 40 | end << 20 | index); the comparison reads start and end only, the moves are the reference's.
struct ErtSort {
    typedef uint64_t Key;
    ERT_HDI static bool lt(Key a, Key b) { return (a >> 20) < (b >> 20); }
    ERT_HDI static void insertsort(Key *a, int lo, int hi)
    {
        for (int i = lo + 1; i < hi; ++i)
            for (int j = i; j > lo && lt(a[j], a[j - 1]); --j) { const Key t = a[j]; a[j] = a[j - 1]; a[j - 1] = t; }
    }
    ERT_FN static void combsort(Key *a, int n)
    {
        const double shrink = 1.2473309501039786540366528676643;
        int do_swap; size_t gap = (size_t)n;
        do {
            if (gap > 2) { gap = (size_t)(gap / shrink); if (gap == 9 || gap == 10) gap = 11; }
            do_swap = 0;
            for (size_t i = 0; i + gap < (size_t)n; ++i) {
                const size_t j = i + gap;
                if (lt(a[j], a[i])) { const Key t = a[i]; a[i] = a[j]; a[j] = t; do_swap = 1; }
            }
        } while (do_swap || gap > 2);
        if (gap != 1) insertsort(a, 0, n);
    }
    ERT_FN static void introsort(Key *a, int n)
    {
        if (n < 1) return;
        if (n == 2) { if (lt(a[1], a[0])) { const Key t = a[0]; a[0] = a[1]; a[1] = t; } return; }
        int d;
        for (d = 2; (1ul << d) < (unsigned long)n; ++d) {}
        int st_l[64], st_r[64], st_d[64], top = 0;
        int sl = 0, tr = n - 1;
        d <<= 1;
        for (;;) {
            if (sl < tr) {
                if (--d == 0) { combsort(a + sl, tr - sl + 1); tr = sl; continue; }
                int i = sl, j = tr, k = i + ((j - i) >> 1) + 1;
                if (lt(a[k], a[i])) { if (lt(a[k], a[j])) k = j; }
                else k = lt(a[j], a[i]) ? i : j;
                const Key rp = a[k];
                if (k != tr) { const Key t = a[k]; a[k] = a[tr]; a[tr] = t; }
                for (;;) {
                    do ++i; while (lt(a[i], rp));
                    do --j; while (i <= j && lt(rp, a[j]));
                    if (j <= i) break;
                    const Key t = a[i]; a[i] = a[j]; a[j] = t;
                }
                { const Key t = a[i]; a[i] = a[tr]; a[tr] = t; }
                if (i - sl > tr - i) {
                    if (i - sl > 16) { st_l[top] = sl; st_r[top] = i - 1; st_d[top] = d; ++top; }
                    sl = tr - i > 16 ? i + 1 : tr;
                } else {
                    if (tr - i > 16) { st_l[top] = i + 1; st_r[top] = tr; st_d[top] = d; ++top; }
                    tr = i - sl > 16 ? i - 1 : sl;
                }
            } else {
                if (top == 0) { insertsort(a, 0, n); return; }
                --top; sl = st_l[top]; tr = st_r[top]; d = st_d[top];
            }
        }
    }
};

// coordinate of hit k of a MEM in the forward-reverse text, as mem_chain_new derives it (src/bwamem.cpp:991-998)
ERT_HDI int64_t ert_hit_rbeg(const ErtMemRec &r, int64_t hit, int64_t l_pac)
{
    if (r.corr < 0) return hit;
    return (l_pac << 1) - (hit + (r.end - r.start) - r.corr);
}
