/* TEST INFRASTRUCTURE ONLY — see oracle.h.
 * CPU restatement of the reference's FM-index seeding:
 *   orc_backward_ext  : FMI_search::backwardExt + GET_OCC   (src/FMI_search.cpp:2029-2056,
 *                                                            src/FMI_search.h:76-83)
 *   pivot_search      : one (read, pivot) step of getSMEMsOnePosOneThread (cpp:1372-1606)
 *   last_like_seeds   : bwtSeedStrategyAllPosOneThread for one read     (cpp:1662-1816)
 *   orc_collect_smem  : mem_collect_smem                                 (src/bwamem.cpp:648-786)
 *   orc_sa_lookup     : get_sa_entries_prefetch / call_one_step          (cpp:2206-2379)
 * Written read-by-read from the algorithm (the reference sweeps all reads pivot by pivot; the
 * emission order is not observable after its final sort).
 * With f->fma_all / f->fma_last set, the walks go through the FMA tables exactly as the SMEM_ACCEL
 * build does (cpp:1414-1463, 1708-1752), including what that code does when an N lies inside the
 * table's k-mer (the tables are otherwise a pure memoisation of backwardExt).
 */
#include <stdlib.h>
#include <string.h>
#include "oracle.h"

typedef struct { int64_t k, l, s; int m, n; } biv_t;    /* bi-interval + query span [m, n] */

static inline int64_t occ(const orc_fmi_t *f, int64_t pos, int c)
{
    const int64_t *rec = f->cp_occ + (pos >> 6) * 8;
    int y = (int)(pos & 63);
    uint64_t bits = (uint64_t)rec[4 + c];
    uint64_t top = y ? (~0ULL << (64 - y)) : 0ULL;      /* first y symbols of the block */
    return rec[c] + __builtin_popcountll(bits & top);
}

void orc_backward_ext(const orc_fmi_t *f, int64_t k, int64_t l, int64_t s, int a, int64_t out[3])
{
    int64_t kk[4], ss[4], ll[4];
    for (int b = 0; b < 4; ++b) {
        int64_t o_sp = occ(f, k, b), o_ep = occ(f, k + s, b);
        kk[b] = f->count[b] + o_sp;
        ss[b] = o_ep - o_sp;
    }
    int64_t sent = (k <= f->sentinel_index && k + s > f->sentinel_index) ? 1 : 0;
    ll[3] = l + sent;
    ll[2] = ll[3] + ss[3];
    ll[1] = ll[2] + ss[2];
    ll[0] = ll[1] + ss[1];
    out[0] = kk[a]; out[1] = ll[a]; out[2] = ss[a];
}

static inline biv_t bext(const orc_fmi_t *f, biv_t v, int a, int64_t *n_ext)
{
    int64_t o[3];
    orc_backward_ext(f, v.k, v.l, v.s, a, o);
    if (n_ext) ++*n_ext;
    v.k = o[0]; v.l = o[1]; v.s = o[2];
    return v;
}

/* forward extension = backward extension of the complement on the swapped interval (cpp:1477-1485) */
static inline biv_t fext(const orc_fmi_t *f, biv_t v, int a, int64_t *n_ext)
{
    biv_t t = v;
    t.k = v.l; t.l = v.k;
    t = bext(f, t, 3 - a, n_ext);
    biv_t r = t;
    r.k = t.l; r.l = t.k;
    return r;
}

static inline biv_t single_base(const orc_fmi_t *f, int a, int x)
{
    biv_t v;
    v.k = f->count[a]; v.l = f->count[3 - a]; v.s = f->count[a + 1] - f->count[a];
    v.m = v.n = x;
    return v;
}

typedef struct { orc_smem_t *out; int64_t cap, n; } sink_t;

static inline void emit(sink_t *o, int rid, biv_t v)
{
    if (o->n < o->cap) {
        orc_smem_t *e = &o->out[o->n];
        e->rid = (uint32_t)rid; e->m = (uint32_t)v.m; e->n = (uint32_t)v.n; e->pad_ = 0;
        e->k = v.k; e->l = v.l; e->s = v.s;
    }
    ++o->n;
}

/* SMEMs through pivot x; returns the next pivot */
static int pivot_search(const orc_fmi_t *f, const uint8_t *q, int len, int x, int min_intv,
                        int min_seed_len, int rid, biv_t *prev, sink_t *o, int64_t *n_ext)
{
    if (q[x] > 3) return x + 1;
    biv_t cur = single_base(f, q[x], x);
    int np = 0, next_x = x + 1, j = x + 1;
    if (f->fma_all && len - x >= 11) {                  /* cpp:1414-1463 */
        uint32_t idx = 0;
        int k;
        for (k = 0; k < 11; ++k) {
            if (q[x + k] >= 4) break;
            idx |= (uint32_t)q[x + k] << ((10 - k) * 2);
        }
        const uint32_t *ent = f->fma_all + (size_t)idx * 32;
        const int with_n = k < 11;
        const int last_avail = (int)ent[0];
        const int last_idx = (k > last_avail ? last_avail : k) - 1;
        for (j = x + 1, k = 0; k < last_idx; ++j, ++k) {
            const int a = q[j];
            next_x = j + 1;
            biv_t nw = cur;
            nw.k = cur.k + ent[1 + 3 * k];
            nw.l = f->count[3 - a] + ent[2 + 3 * k];
            nw.s = ent[3 + 3 * k];
            nw.n = j;
            prev[np] = cur;
            if (nw.s != cur.s) ++np;
            if (nw.s < min_intv) { next_x = j; j = len; break; }
            cur = nw;
        }
        if (with_n) { next_x = j + 1; j = len; }
    }
    for (; j < len; ++j) {
        next_x = j + 1;
        if (q[j] > 3) break;
        biv_t nw = fext(f, cur, q[j], n_ext);
        nw.n = j;
        prev[np] = cur;
        if (nw.s != cur.s) ++np;
        if (nw.s < min_intv) { next_x = j; break; }
        cur = nw;
    }
    if (cur.s >= min_intv) prev[np++] = cur;
    for (int p = 0; p < np / 2; ++p) { biv_t t = prev[p]; prev[p] = prev[np - 1 - p]; prev[np - 1 - p] = t; }

    for (j = x - 1; j >= 0; --j) {
        int a = q[j];
        if (a > 3) break;
        int nc = 0, decided = 0;
        int32_t last_s = -1;                 /* an int upstream (curr_s): keep the narrowing */
        for (int p = 0; p < np; ++p) {
            biv_t nw = bext(f, prev[p], a, n_ext);
            nw.m = j;
            if (nw.s < min_intv) {
                if (!decided && prev[p].n - prev[p].m + 1 >= min_seed_len) {
                    emit(o, rid, prev[p]);
                    decided = 1;
                }
            } else if (nw.s != (int64_t)last_s) {
                last_s = (int32_t)nw.s;
                prev[nc++] = nw;
                decided = 1;
            }
        }
        np = nc;
        if (np == 0) break;
    }
    if (np != 0 && prev[0].n - prev[0].m + 1 >= min_seed_len) emit(o, rid, prev[0]);
    return next_x;
}

static void last_like_seeds(const orc_fmi_t *f, const uint8_t *q, int len, int64_t max_intv,
                            int min_seed_len, int rid, sink_t *o, int64_t *n_ext)
{
    int x = 0;
    while (x < len) {
        int next_x = x + 1;
        if (q[x] < 4) {
            biv_t cur = single_base(f, q[x], x);
            int j = x + 1;
            if (f->fma_last && len - x >= 13) {             /* cpp:1708-1752 */
                uint32_t idx = 0, with_n = 0;
                for (int k = 0; k < 13; ++k) {
                    idx |= (uint32_t)q[x + k] << ((12 - k) * 2);
                    with_n += q[x + k] >> 2;
                }
                if (!with_n) {
                    const uint32_t *e = f->fma_last + (size_t)idx * 4;
                    j = x + (int)(e[0] & 0xFF);
                    next_x = j;
                    cur.k = (int64_t)(((uint64_t)(int64_t)(int8_t)(e[0] >> 8) << 32) | e[1]);
                    cur.l = (int64_t)(((uint64_t)(int64_t)(int8_t)(e[0] >> 16) << 32) | e[2]);
                    cur.s = (int64_t)(((uint64_t)(int64_t)(int8_t)(e[0] >> 24) << 32) | e[3]);
                    cur.n = j - 1;
                    if (cur.s < max_intv && cur.n - cur.m + 1 >= min_seed_len && cur.s > 0)
                        emit(o, rid, cur);                  /* no break here upstream */
                }
            }
            for (; j < len; ++j) {
                next_x = j + 1;
                if (q[j] > 3) break;
                cur = fext(f, cur, q[j], n_ext);
                cur.n = j;
                if (cur.s < max_intv && cur.n - cur.m + 1 >= min_seed_len) {
                    if (cur.s > 0) emit(o, rid, cur);
                    break;
                }
            }
        }
        x = next_x;
    }
}

/* ---- FMA table construction (SMEM_ACCEL) ------------------------------------------------------ */
void orc_fma_entry_all(const orc_fmi_t *f, uint32_t idx, uint32_t out[32])
{
    memset(out, 0, 32 * sizeof(uint32_t));
    int a = (idx >> 20) & 3;
    biv_t cur = single_base(f, a, 0);
    for (int i = 1; i < 11; ++i) {
        a = (idx >> (2 * (10 - i))) & 3;
        biv_t nw = fext(f, cur, a, NULL);
        out[1 + 3 * (i - 1)] = (uint32_t)(nw.k - cur.k);
        out[2 + 3 * (i - 1)] = (uint32_t)(nw.l - f->count[3 - a]);
        out[3 + 3 * (i - 1)] = (uint32_t)nw.s;
        if (nw.s > 0) out[0] = (uint32_t)i; else break;
        cur = nw;
    }
}

void orc_fma_entry_last(const orc_fmi_t *f, uint32_t idx, uint32_t out[4])
{
    int a = (idx >> 24) & 3, i;
    biv_t cur = single_base(f, a, 0);
    for (i = 1; i < 13; ++i) {
        a = (idx >> (2 * (12 - i))) & 3;
        biv_t nw = fext(f, cur, a, NULL);
        if (nw.s == 0) break;
        cur = nw;
    }
    out[0] = (uint32_t)i | (((uint32_t)(cur.k >> 32) & 0xFF) << 8) | (((uint32_t)(cur.l >> 32) & 0xFF) << 16) |
             (((uint32_t)(cur.s >> 32) & 0xFF) << 24);
    out[1] = (uint32_t)cur.k; out[2] = (uint32_t)cur.l; out[3] = (uint32_t)cur.s;
}

void orc_fma_build(const orc_fmi_t *f, uint32_t *all_tab, uint32_t *last_tab)
{
    if (all_tab)
        for (uint32_t i = 0; i < (1u << 22); ++i) orc_fma_entry_all(f, i, all_tab + (size_t)i * 32);
    if (last_tab)
        for (uint32_t i = 0; i < (1u << 26); ++i) orc_fma_entry_last(f, i, last_tab + (size_t)i * 4);
}

void orc_fma_fill(const orc_fmi_t *f, int which_last, const uint32_t *idx, int64_t n, uint32_t *tab)
{
    for (int64_t i = 0; i < n; ++i) {
        if (which_last) orc_fma_entry_last(f, idx[i], tab + (size_t)idx[i] * 4);
        else orc_fma_entry_all(f, idx[i], tab + (size_t)idx[i] * 32);
    }
}

static int cmp_smem(const void *a, const void *b)
{
    const orc_smem_t *x = (const orc_smem_t *)a, *y = (const orc_smem_t *)b;
    if (x->rid != y->rid) return x->rid < y->rid ? -1 : 1;
    if (x->m != y->m) return x->m < y->m ? -1 : 1;
    if (x->n != y->n) return x->n < y->n ? -1 : 1;
    return 0;
}

int64_t orc_collect_smem(const orc_fmi_t *f, const orc_seed_opt_t *opt, const uint8_t *reads,
                         const int64_t *off, const int32_t *len, const uint8_t *skip, int32_t nseq,
                         orc_smem_t *out, int64_t cap, int64_t *n_ext)
{
    sink_t o = { out, cap, 0 };
    const int split_len = (int)(opt->min_seed_len * opt->split_factor + .499);
    int maxlen = 1;
    for (int i = 0; i < nseq; ++i) if (len[i] > maxlen) maxlen = len[i];
    biv_t *prev = (biv_t *)malloc(sizeof(biv_t) * (size_t)(maxlen + 1));
    for (int i = 0; i < nseq; ++i) {
        if (skip && skip[i]) continue;
        const uint8_t *q = reads + off[i];
        const int L = len[i];
        const int64_t first = o.n;
        /* round 1: sweep the pivots with min_intv = 1 */
        for (int x = 0; x < L;) x = pivot_search(f, q, L, x, 1, opt->min_seed_len, i, prev, &o, n_ext);
        /* round 2: re-seed long, rare SMEMs from their middle with min_intv = s + 1 */
        const int64_t last = o.n;
        for (int64_t e = first; e < last && e < cap; ++e) {
            const orc_smem_t p = out[e];
            int start = (int)p.m, end = (int)p.n + 1;
            if (end - start < split_len || p.s > opt->split_width) continue;
            pivot_search(f, q, L, (end + start) >> 1, (int)(p.s + 1), opt->min_seed_len, i, prev, &o, n_ext);
        }
        /* round 3: forward-only seeds that become rarer than max_mem_intv */
        if (opt->max_mem_intv > 0)
            last_like_seeds(f, q, L, opt->max_mem_intv, opt->min_seed_len + 1, i, &o, n_ext);
    }
    free(prev);
    qsort(out, (size_t)(o.n < cap ? o.n : cap), sizeof(orc_smem_t), cmp_smem);
    return o.n;
}

int64_t orc_sa_lookup(const orc_fmi_t *f, const orc_smem_t *smems, int64_t n, int32_t max_occ,
                      int64_t *coords, int64_t *n_lf)
{
    int64_t w = 0;
    for (int64_t i = 0; i < n; ++i) {
        const orc_smem_t *p = &smems[i];
        int64_t step = p->s > max_occ ? p->s / max_occ : 1;
        int c = 0;
        for (int64_t row = p->k; row < p->k + p->s && c < max_occ; row += step, ++c) {
            int64_t pos = row, walked = 0, sa;
            for (;;) {
                if ((pos & 7) == 0) {
                    sa = ((int64_t)f->sa_ms_byte[pos >> 3] << 32) + f->sa_ls_word[pos >> 3] + walked;
                    break;
                }
                const int64_t *rec = f->cp_occ + (pos >> 6) * 8;
                int sh = 63 - (int)(pos & 63), b = 4;
                for (int c2 = 0; c2 < 4; ++c2)
                    if (((uint64_t)rec[4 + c2] >> sh) & 1) { b = c2; break; }
                if (b == 4) { sa = 0; break; }          /* sentinel row: steps discarded */
                pos = f->count[b] + occ(f, pos, b);
                ++walked;
                if (n_lf) ++*n_lf;
            }
            coords[w++] = sa;
        }
    }
    return w;
}
