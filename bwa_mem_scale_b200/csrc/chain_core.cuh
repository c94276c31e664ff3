// Per-read seed chaining: the reference's mem_chain_seeds (src/bwamem.cpp:789-936) followed by
// mem_chain_flt (:528-646), as per-thread code shared by the CUDA kernels in chain.cu and — compiled
// as plain host C++ — by tests/hostsim (test scaffolding, not a product path).
//
// What has to be reproduced exactly, because it is observable in the chain list:
//   * seeds arrive SMEM by SMEM, SA row by SA row; each is offered to ONE chain only — the closest
//     chain at or before its reference position, found in a B-tree keyed by chain position
//     (kbtree.h, KBTREE_INIT(chn, mem_chain_t, chain_cmp) :63-64).  Several chains may share a
//     position, and which of them the tree returns depends on its shape, so the tree is emulated
//     node for node: order t = 5 (KB_DEFAULT_SIZE + 8 bytes per node, 48-byte keys), the same
//     binary search (__kb_getp_aux), the same split-on-the-way-down insertion (__kb_putp_aux);
//   * test_and_merge (:379-424);
//   * chains come out of the tree in order, get weights (mem_chain_weight :455-474), are ordered by
//     weight with ks_introsort (ksort.h — NOT a stable sort, so its exact steps are replayed:
//     median-of-three quicksort leaving runs <= 16 to a final insertion sort, comb sort when the
//     depth budget runs out) and filtered pairwise.
#pragma once
#include <stdint.h>
#include "../../include/bwagpu.h"

#if defined(__CUDACC__)
#define CHAIN_HD __host__ __device__ __forceinline__
#else
#define CHAIN_HD inline
#endif

// CHAIN_LINEAR_SEARCH (getp_aux): early-exit scan of a node's sorted keys instead of nine predicated compares.  MEASURED
// 117.9 vs 116.2 ms per 5 M reads (scripts/gpu_r2_ai.sh): no gain — the dependent loads cost what the compares saved.  OFF.
#ifndef CHAIN_LINEAR_SEARCH
#define CHAIN_LINEAR_SEARCH 0
#endif
#define CHAIN_BT_T 5                        // kb_init: ((520 - 4 - 8) / (8 + 48) + 1) >> 1
#define CHAIN_BT_MAXKEYS (2 * CHAIN_BT_T - 1)

struct ChainOpt {
    int w, max_chain_gap, max_occ, min_seed_len, min_chain_weight, max_chain_extend;
    float mask_level, drop_ratio;
};

struct RefView {                            // contig table of the reference replica
    int64_t l_pac;
    int n_seqs;
    const int64_t *contig_off;
    const int32_t *contig_len;
    const uint8_t *contig_is_alt;           // may be NULL
};

// working chain = the mem_chain_t key the reference stores in the tree
struct WChain {
    int64_t pos;                            // seeds[0].rbeg
    int64_t last_rbeg;
    int32_t first_qbeg, last_qbeg, last_len;
    int32_t rid, n, head, tail;             // head/tail: seed list (links in bwagpu_seed.pad_)
    int32_t w, kept, first, is_alt;
};

// kpos[a] mirrors ch[key[a]].pos so that a search reads one node instead of chasing every key
struct BtNode { int32_t n, is_internal; int32_t key[CHAIN_BT_MAXKEYS]; int32_t ptr[CHAIN_BT_MAXKEYS + 1]; int64_t kpos[CHAIN_BT_MAXKEYS]; };

CHAIN_HD int chain_pos2rid(const RefView &r, int64_t pos_f)          // bns_pos2rid, src/bntseq.cpp:397-411
{
    if (pos_f >= r.l_pac) return -1;
    int left = 0, mid = 0, right = r.n_seqs;
    while (left < right) {
        mid = (left + right) >> 1;
        if (pos_f >= r.contig_off[mid]) {
            if (mid == r.n_seqs - 1) break;
            if (pos_f < r.contig_off[mid + 1]) break;
            left = mid + 1;
        } else right = mid;
    }
    return mid;
}

CHAIN_HD int chain_intv2rid(const RefView &r, int64_t rb, int64_t re)  // bns_intv2rid, :413-421
{
    if (rb < r.l_pac && re > r.l_pac) return -2;
    const int64_t fb = rb >= r.l_pac ? (r.l_pac << 1) - 1 - rb : rb;
    const int rid_b = chain_pos2rid(r, fb);
    int rid_e = rid_b;
    if (rb < re) {
        const int64_t e = re - 1, fe = e >= r.l_pac ? (r.l_pac << 1) - 1 - e : e;
        // nearly always the interval ends in the contig it starts in: one look at that contig's bounds instead of a second search
        const bool same = rid_b >= 0 && fe >= r.contig_off[rid_b] && (rid_b == r.n_seqs - 1 || fe < r.contig_off[rid_b + 1]);
        rid_e = same ? rid_b : chain_pos2rid(r, fe);
    }
    return rid_b == rid_e ? rid_b : -1;
}

// Scratch of one read, all sized by the number of SA coordinates `cap` the read has:
//   sd[cap]    seeds in arrival order (pad_ = link to the next seed of the same chain)
//   ch[cap]    working chains          nodes[cap/4 + 3]  tree nodes
//   ord[cap], lst[cap]  chain order / kept list of mem_chain_flt
struct ChainScratch { bwagpu_seed *sd; WChain *ch; BtNode *nodes; int32_t *ord, *lst; };

struct ChainBuilder {
    ChainScratch s;
    int n_seeds, n_chains, n_nodes, root;

    CHAIN_HD int cmp_pos(int64_t a, int64_t b) const { return (b < a) - (a < b); }

    // __kb_getp_aux: the binary search's answer is the lower bound, i.e. the number of keys whose pos
    // is < k (a node's keys are sorted by pos); counted over the node's own copy of the positions,
    // nine independent loads instead of a chain of dependent ones
    CHAIN_HD int getp_aux(const BtNode &x, int64_t k, int *r) const
    {
        const int n = x.n;
        if (n == 0) return -1;
        int begin = 0, eq = 0;
#if CHAIN_LINEAR_SEARCH
        // the keys of a node are sorted: stop at the first one that is not below k (on average half the node instead of
        // always nine predicated compares — the kernel runs one or two lanes per warp, so instructions are what counts)
        while (begin < n && x.kpos[begin] < k) ++begin;
        eq = begin < n && x.kpos[begin] == k;
#else
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int a = 0; a < CHAIN_BT_MAXKEYS; ++a) {
            const int64_t p = a < n ? x.kpos[a] : k + 1;
            begin += p < k;
            eq |= p == k;
        }
#endif
        if (begin == n) { *r = 1; return n - 1; }
        *r = eq ? 0 : -1;
        return eq ? begin : begin - 1;
    }

    CHAIN_HD int interval_lower(int64_t k) const                           // kb_intervalp: lower only
    {
        int lower = -1, r = 0, x = root;
        for (;;) {
            const BtNode &nd = s.nodes[x];
            const int i = getp_aux(nd, k, &r);
            if (i >= 0 && r == 0) return nd.key[i];
            if (i >= 0) lower = nd.key[i];
            if (!nd.is_internal) return lower;
            x = nd.ptr[i + 1];
        }
    }

    CHAIN_HD int new_node(int internal)
    {
        BtNode &z = s.nodes[n_nodes];
        z.n = 0; z.is_internal = internal;
        return n_nodes++;
    }

    CHAIN_HD void split(int xi, int i, int yi)                             // __kb_split: x internal, y = x.ptr[i] full
    {
        const int zi = new_node(s.nodes[yi].is_internal);
        BtNode &x = s.nodes[xi], &y = s.nodes[yi], &z = s.nodes[zi];
        z.n = CHAIN_BT_T - 1;
        for (int a = 0; a < CHAIN_BT_T - 1; ++a) { z.key[a] = y.key[CHAIN_BT_T + a]; z.kpos[a] = y.kpos[CHAIN_BT_T + a]; }
        if (y.is_internal) for (int a = 0; a < CHAIN_BT_T; ++a) z.ptr[a] = y.ptr[CHAIN_BT_T + a];
        y.n = CHAIN_BT_T - 1;
        for (int a = x.n; a > i; --a) x.ptr[a + 1] = x.ptr[a];
        x.ptr[i + 1] = zi;
        for (int a = x.n - 1; a >= i; --a) { x.key[a + 1] = x.key[a]; x.kpos[a + 1] = x.kpos[a]; }
        x.key[i] = y.key[CHAIN_BT_T - 1]; x.kpos[i] = y.kpos[CHAIN_BT_T - 1];
        ++x.n;
    }

    CHAIN_HD void put(int ci)                                              // kb_putp
    {
        const int64_t k = s.ch[ci].pos;
        if (s.nodes[root].n == CHAIN_BT_MAXKEYS) {
            const int r0 = root, sn = new_node(1);
            s.nodes[sn].ptr[0] = r0;
            root = sn;
            split(sn, 0, r0);
        }
        int xi = root;
        for (;;) {                                                         // __kb_putp_aux, iteratively
            BtNode &x = s.nodes[xi];
            int r;
            if (!x.is_internal) {
                const int i = getp_aux(x, k, &r);
                for (int a = x.n - 1; a > i; --a) { x.key[a + 1] = x.key[a]; x.kpos[a + 1] = x.kpos[a]; }
                x.key[i + 1] = ci; x.kpos[i + 1] = k;
                ++x.n;
                return;
            }
            int i = getp_aux(x, k, &r) + 1;
            if (s.nodes[x.ptr[i]].n == CHAIN_BT_MAXKEYS) {
                split(xi, i, x.ptr[i]);
                if (cmp_pos(k, s.nodes[xi].kpos[i]) > 0) ++i;
            }
            xi = s.nodes[xi].ptr[i];
        }
    }

    CHAIN_HD void init(const ChainScratch &sc)
    {
        s = sc; n_seeds = n_chains = n_nodes = 0;
        root = new_node(0);
    }

    // test_and_merge (:379-424); returns 1 when the seed was absorbed (appended or contained)
    CHAIN_HD int test_and_merge(const ChainOpt &o, int64_t l_pac, WChain &c, const bwagpu_seed &p, int seed_rid, int si)
    {
        const int64_t qend = c.last_qbeg + c.last_len, rend = c.last_rbeg + c.last_len;
        if (seed_rid != c.rid) return 0;
        if (p.qbeg >= c.first_qbeg && p.qbeg + p.len <= qend && p.rbeg >= c.pos && p.rbeg + p.len <= rend) return 1;
        if ((c.last_rbeg < l_pac || c.pos < l_pac) && p.rbeg >= l_pac) return 0;
        const int64_t x = p.qbeg - c.last_qbeg, y = p.rbeg - c.last_rbeg;
        if (y >= 0 && x - y <= o.w && y - x <= o.w && x - c.last_len < o.max_chain_gap && y - c.last_len < o.max_chain_gap) {
            s.sd[si] = p; s.sd[si].pad_ = -1;
            s.sd[c.tail].pad_ = si;
            c.tail = si; ++c.n;
            c.last_rbeg = p.rbeg; c.last_qbeg = p.qbeg; c.last_len = p.len;
            return 2;
        }
        return 0;
    }

    // one seed in arrival order (:870-934)
    CHAIN_HD void add_seed(const ChainOpt &o, const RefView &rv, const bwagpu_seed &p)
    {
        const int rid = chain_intv2rid(rv, p.rbeg, p.rbeg + p.len);
        if (rid < 0) return;
        bool to_add = true;
        const int si = n_seeds;                       // slot this seed takes if it is kept
        if (n_chains > 0) {
            const int lower = interval_lower(p.rbeg);
            if (lower >= 0) {
                const int m = test_and_merge(o, rv.l_pac, s.ch[lower], p, rid, si);
                if (m) { to_add = false; if (m == 2) ++n_seeds; }
            }
        }
        if (to_add) {
            s.sd[si] = p; s.sd[si].pad_ = -1; ++n_seeds;
            WChain &c = s.ch[n_chains];
            c.pos = p.rbeg; c.first_qbeg = p.qbeg; c.rid = rid; c.n = 1; c.head = c.tail = si;
            c.last_rbeg = p.rbeg; c.last_qbeg = p.qbeg; c.last_len = p.len;
            c.w = 0; c.kept = 0; c.first = -1;
            c.is_alt = rv.contig_is_alt ? (rv.contig_is_alt[rid] != 0) : 0;
            put(n_chains++);
        }
    }

    // chains in tree order -> ord[0..n_chains)   (__kb_traverse)
    CHAIN_HD void traverse()
    {
        int stack_x[24], stack_i[24], sp = 0, out = 0;
        stack_x[0] = root; stack_i[0] = 0;
        for (;;) {
            while (stack_x[sp] >= 0 && stack_i[sp] <= s.nodes[stack_x[sp]].n) {
                const BtNode &x = s.nodes[stack_x[sp]];
                stack_i[sp + 1] = 0;
                stack_x[sp + 1] = x.is_internal ? x.ptr[stack_i[sp]] : -1;
                ++sp;
            }
            --sp;
            if (sp < 0) break;
            if (stack_x[sp] >= 0 && stack_i[sp] < s.nodes[stack_x[sp]].n) s.ord[out++] = s.nodes[stack_x[sp]].key[stack_i[sp]];
            ++stack_i[sp];
        }
    }

    CHAIN_HD int weight(const WChain &c) const                              // mem_chain_weight :455-474
    {
        if (c.n == 1) return c.last_len < (1 << 30) ? c.last_len : (1 << 30) - 1;   // one seed: both covers = its length
        int64_t end = 0; int w = 0;
        for (int si = c.head; si >= 0; si = s.sd[si].pad_) {
            const bwagpu_seed &e = s.sd[si];
            if (e.qbeg >= end) w += e.len;
            else if (e.qbeg + e.len > end) w += (int)(e.qbeg + e.len - end);
            end = end > e.qbeg + e.len ? end : e.qbeg + e.len;
        }
        const int tmp = w; w = 0; end = 0;
        for (int si = c.head; si >= 0; si = s.sd[si].pad_) {
            const bwagpu_seed &e = s.sd[si];
            if (e.rbeg >= end) w += e.len;
            else if (e.rbeg + e.len > end) w += (int)(e.rbeg + e.len - end);
            end = end > e.rbeg + e.len ? end : e.rbeg + e.len;
        }
        w = w < tmp ? w : tmp;
        return w < (1 << 30) ? w : (1 << 30) - 1;
    }

    // ---- ks_introsort(mem_flt) replayed with flt_lt(a,b) = a.w > b.w ---------------------------------
    // The elements are (weight << 32 | chain index): every comparison reads the weight from the element
    // itself instead of chasing the index into the chain array; the moves are those of the reference.
    typedef uint64_t SortKey;
    CHAIN_HD bool lt(SortKey a, SortKey b) const { return (uint32_t)(a >> 32) > (uint32_t)(b >> 32); }
    CHAIN_HD void insertsort(SortKey *a, int lo, int hi)                     // [lo, hi)
    {
        for (int i = lo + 1; i < hi; ++i)
            for (int j = i; j > lo && lt(a[j], a[j - 1]); --j) { const SortKey t = a[j]; a[j] = a[j - 1]; a[j - 1] = t; }
    }
    CHAIN_HD void combsort(SortKey *a, int n)
    {
        const double shrink = 1.2473309501039786540366528676643;
        int do_swap; size_t gap = (size_t)n;
        do {
            if (gap > 2) { gap = (size_t)(gap / shrink); if (gap == 9 || gap == 10) gap = 11; }
            do_swap = 0;
            for (size_t i = 0; i + gap < (size_t)n; ++i) {
                const size_t j = i + gap;
                if (lt(a[j], a[i])) { const SortKey t = a[i]; a[i] = a[j]; a[j] = t; do_swap = 1; }
            }
        } while (do_swap || gap > 2);
        if (gap != 1) insertsort(a, 0, n);
    }
    CHAIN_HD void introsort(SortKey *a, int n)
    {
        if (n < 1) return;
        if (n == 2) { if (lt(a[1], a[0])) { const SortKey t = a[0]; a[0] = a[1]; a[1] = t; } return; }
        int d;
        for (d = 2; (1ul << d) < (unsigned long)n; ++d) {}
        int st_l[72], st_r[72], st_d[72], top = 0;
        int sl = 0, tr = n - 1;
        d <<= 1;
        for (;;) {
            if (sl < tr) {
                if (--d == 0) { combsort(a + sl, tr - sl + 1); tr = sl; continue; }
                int i = sl, j = tr, k = i + ((j - i) >> 1) + 1;
                if (lt(a[k], a[i])) { if (lt(a[k], a[j])) k = j; }
                else k = lt(a[j], a[i]) ? i : j;
                const SortKey rp = a[k];
                if (k != tr) { const SortKey t = a[k]; a[k] = a[tr]; a[tr] = t; }
                for (;;) {
                    do ++i; while (lt(a[i], rp));
                    do --j; while (i <= j && lt(rp, a[j]));
                    if (j <= i) break;
                    const SortKey t = a[i]; a[i] = a[j]; a[j] = t;
                }
                { const SortKey t = a[i]; a[i] = a[tr]; a[tr] = t; }
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

    CHAIN_HD int chn_beg(const WChain &c) const { return c.first_qbeg; }
    CHAIN_HD int chn_end(const WChain &c) const { return c.last_qbeg + c.last_len; }

    // mem_chain_flt for this read's chains (:528-646), in three parts so that the kernels can run the middle one —
    // the scan of every chain against the list of chains kept so far, quadratic for reads from repeats — with a whole
    // warp (chain.cu).  On return of filter() ord[0..result) are the surviving chains in their final order;
    // WChain.w/kept/first are set.
    struct Kept { int32_t beg, end, w, flags; };      // the fields the scan reads, side by side (flags: 1 = is_alt, 2 = `first` set)
    CHAIN_HD Kept *kept_list() const { return reinterpret_cast<Kept *>(s.nodes); }   // node storage is dead after traverse()

    // weights, the weight filter, the replayed introsort, the first kept chain.  Returns n (0: nothing to filter).
    CHAIN_HD int filter_prepare(const ChainOpt &o)
    {
        if (n_chains == 0) return 0;
        int n = 0;
        for (int i = 0; i < n_chains; ++i) {
            WChain &c = s.ch[s.ord[i]];
            c.first = -1; c.kept = 0;
            c.w = weight(c);
            if (c.w >= o.min_chain_weight) s.ord[n++] = s.ord[i];
        }
        if (n == 0) return 0;
        int32_t *a = s.ord;
        {
            SortKey *sk = reinterpret_cast<SortKey *>(s.nodes);
            for (int i = 0; i < n; ++i) sk[i] = ((SortKey)(uint32_t)s.ch[a[i]].w << 32) | (uint32_t)a[i];
            introsort(sk, n);
            for (int i = 0; i < n; ++i) a[i] = (int32_t)(uint32_t)sk[i];
        }
        Kept *lv = kept_list();
        WChain &c0 = s.ch[a[0]];
        c0.kept = 3;
        lv[0].beg = chn_beg(c0); lv[0].end = chn_end(c0); lv[0].w = c0.w; lv[0].flags = c0.is_alt ? 1 : 0;
        s.lst[0] = 0;
        return n;
    }

    // does chain (bi, ei, wi, alt_i) overlap kept entry kv by mask_level (bit 0), and is it dropped by it (bit 1)?
    CHAIN_HD int flt_test(const ChainOpt &o, const Kept &kv, int bi, int ei, int wi, int alt_i, int msl2) const
    {
        const int b_max = kv.beg > bi ? kv.beg : bi;
        const int e_min = kv.end < ei ? kv.end : ei;
        if (e_min > b_max && (!(kv.flags & 1) || alt_i)) {
            const int li = ei - bi, lj = kv.end - kv.beg;
            const int min_l = li < lj ? li : lj;
            if ((float)(e_min - b_max) >= (float)min_l * o.mask_level && min_l < o.max_chain_gap)
                return 1 | (((float)wi < (float)kv.w * o.drop_ratio && kv.w - wi >= msl2) ? 2 : 0);
        }
        return 0;
    }

    // the pairwise scan, one chain after the other (:560-598).  Returns the length of the kept list.
    CHAIN_HD int filter_scan(const ChainOpt &o, int n)
    {
        int32_t *a = s.ord;
        Kept *lv = kept_list();
        int nl = 1;
        const int msl2 = o.min_seed_len << 1;
        for (int i = 1; i < n; ++i) {
            int large_ovlp = 0, k;
            WChain &ci = s.ch[a[i]];
            const int bi = chn_beg(ci), ei = chn_end(ci), wi = ci.w, alt_i = ci.is_alt;
            for (k = 0; k < nl; ++k) {
                const Kept kv = lv[k];
                const int t = flt_test(o, kv, bi, ei, wi, alt_i, msl2);
                if (t) {
                    large_ovlp = 1;
                    if (!(kv.flags & 2)) { s.ch[a[s.lst[k]]].first = i; lv[k].flags = kv.flags | 2; }
                    if (t & 2) break;
                }
            }
            if (k == nl) {
                lv[nl].beg = bi; lv[nl].end = ei; lv[nl].w = wi; lv[nl].flags = alt_i ? 1 : 0;
                s.lst[nl++] = i; ci.kept = large_ovlp ? 2 : 3;
            }
        }
        return nl;
    }

    CHAIN_HD int filter(const ChainOpt &o)
    {
        const int n = filter_prepare(o);
        if (n == 0) return 0;
        return filter_finish(o, n, filter_scan(o, n));
    }

    // kept = 1 marks, the max_chain_extend cut, compaction (:600-646).  Returns the number of surviving chains.
    CHAIN_HD int filter_finish(const ChainOpt &o, int n, int nl)
    {
        int32_t *a = s.ord;
        for (int i = 0; i < nl; ++i) {
            const WChain &c = s.ch[a[s.lst[i]]];
            if (c.first >= 0) s.ch[a[c.first]].kept = 1;
        }
        int i, k;
        for (i = k = 0; i < n; ++i) {
            const int kept = s.ch[a[i]].kept;
            if (kept == 0 || kept == 3) continue;
            if (++k >= o.max_chain_extend) break;
        }
        for (; i < n; ++i) if (s.ch[a[i]].kept < 3) s.ch[a[i]].kept = 0;
        int m = 0;
        for (i = 0; i < n; ++i) if (s.ch[a[i]].kept != 0) a[m++] = a[i];
        return m;
    }
};
