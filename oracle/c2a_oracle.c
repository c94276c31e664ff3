/* TEST INFRASTRUCTURE ONLY — see oracle.h.
 * CPU restatement of mem_chain2aln_across_reads_V2 (reference src/bwamem.cpp:2810-3745; SURVEY.md
 * §8a row a17): for every seed of every chain one alignment region; left and right extension jobs
 * (SeqPair) built from the read and the reference window, run through the banded-SW kernels with
 * the 8-bit / 16-bit / scalar routing of sortPairsLenExt (:2667-2740) and the two-try band
 * doubling (:3218-3618); results applied to the regions; then the containment filter (:3634-3745).
 * Chains arrive as the reference's chain_ar flattened (orc_chain_t / orc_seed_t).
 */
#include <stdlib.h>
#include <string.h>
#include "oracle.h"

#define H0_ (-99)
#define MAX_BAND_TRY 2

static int cal_max_gap(const orc_ext_opt_t *o, int qlen)           /* :94-104 */
{
    int l_del = (int)((double)(qlen * o->a - o->o_del) / o->e_del + 1.);
    int l_ins = (int)((double)(qlen * o->a - o->o_ins) / o->e_ins + 1.);
    int l = l_del > l_ins ? l_del : l_ins;
    l = l > 1 ? l : 1;
    return l < o->w << 1 ? l : o->w << 1;
}

static int cmp_u64(const void *a, const void *b)
{
    uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return x < y ? -1 : x > y;
}

static int pair_kind(const orc_seqpair_t *sp, int a)               /* sortPairsLenExt :2684-2694 */
{
    int mn = sp->len1 < sp->len2 ? sp->len1 : sp->len2;
    int minval = sp->h0 + mn * a;
    if (sp->len1 < 128 && sp->len2 < 128 && minval < 128) return 8;
    if (sp->len1 < 32768 && sp->len2 < 32768 && minval < 32768) return 16;
    return 0;
}

static void seedcov(orc_alnreg_t *a, const orc_chain_t *c, const orc_seed_t *cs)
{
    if (a->rb != H0_ && a->qb != H0_ && a->qe != H0_ && a->re != H0_) {
        a->seedcov = 0;
        for (int i = 0; i < c->n; ++i) {
            const orc_seed_t *t = &cs[i];
            if (t->qbeg >= a->qb && t->qbeg + t->len <= a->qe && t->rbeg >= a->rb && t->rbeg + t->len <= a->re)
                a->seedcov += t->len;
        }
    }
}

/* one side (left = 1 / right = 0): both band tries over the three kernel buckets */
static void run_side(const orc_ext_opt_t *o, int left, orc_seqpair_t *pairs, int n, const uint8_t *rbuf,
                     const uint8_t *qbuf, orc_alnreg_t *regs, const orc_chain_t *chains, const orc_seed_t *seeds,
                     const int32_t *len)
{
    orc_bsw_params_t P = { o->o_del, o->e_del, o->o_ins, o->e_ins, o->zdrop, left ? o->pen_clip5 : o->pen_clip3,
                           o->a, o->b };
    const int kinds[3] = { 0, 16, 8 };
    orc_seqpair_t *cur = malloc(sizeof(orc_seqpair_t) * (size_t)(n + 1)), *nxt = malloc(sizeof(orc_seqpair_t) * (size_t)(n + 1));
    for (int kk = 0; kk < 3; ++kk) {
        int nump = 0;
        for (int i = 0; i < n; ++i) if (pair_kind(&pairs[i], o->a) == kinds[kk]) cur[nump++] = pairs[i];
        for (int t = 0; t < MAX_BAND_TRY; ++t) {
            const int w = o->w << t;
            orc_bsw_batch(&P, kinds[kk], cur, nump, rbuf, qbuf, w);
            int num = 0;
            for (int l = 0; l < nump; ++l) {
                orc_seqpair_t *sp = &cur[l];
                orc_alnreg_t *a = &regs[sp->id];                 /* id = global region index */
                const int prev = a->score;
                a->score = sp->score;
                if (a->score == prev || sp->max_off < (w >> 1) + (w >> 2) || t + 1 == MAX_BAND_TRY) {
                    if (left) {
                        if (sp->gscore <= 0 || sp->gscore <= a->score - o->pen_clip5) {
                            a->qb -= sp->qle; a->rb -= sp->tle; a->truesc = a->score;
                        } else { a->qb = 0; a->rb -= sp->gtle; a->truesc = sp->gscore; }
                    } else {
                        if (sp->gscore <= 0 || sp->gscore <= a->score - o->pen_clip3) {
                            a->qe += sp->qle; a->re += sp->tle; a->truesc += a->score - sp->h0;
                        } else { a->qe = len[sp->seqid]; a->re += sp->gtle; a->truesc += sp->gscore - sp->h0; }
                    }
                    a->w = a->w > w ? a->w : w;
                    seedcov(a, &chains[a->chain], seeds + chains[a->chain].seed_off);
                } else nxt[num++] = *sp;
            }
            nump = num;
            orc_seqpair_t *tmp = cur; cur = nxt; nxt = tmp;
        }
    }
    free(cur); free(nxt);
}

void orc_chain2aln(const orc_ext_opt_t *o, int64_t l_pac, const int64_t *contig_off, const int32_t *contig_len,
                   const uint8_t *ref_string, const uint8_t *reads, const int64_t *off, const int32_t *len,
                   int32_t nseq, const orc_chain_t *chains, int64_t n_chains, const orc_seed_t *seeds,
                   orc_alnreg_t *regs, uint64_t *cells)
{
    int64_t n_seeds = 0;
    for (int64_t c = 0; c < n_chains; ++c) n_seeds += chains[c].n;
    orc_seqpair_t *L = malloc(sizeof(orc_seqpair_t) * (size_t)(n_seeds + 1)), *R = malloc(sizeof(orc_seqpair_t) * (size_t)(n_seeds + 1));
    int64_t lq = 0, lr = 0, rq = 0, rr = 0, nL = 0, nR = 0;
    /* sizes first (exact), then fill */
    uint32_t *srt_all = malloc(sizeof(uint32_t) * (size_t)(n_seeds + 1));
    int64_t *rmax0 = malloc(sizeof(int64_t) * (size_t)(n_chains + 1)), *rmax1 = malloc(sizeof(int64_t) * (size_t)(n_chains + 1));
    for (int64_t ci = 0; ci < n_chains; ++ci) {
        const orc_chain_t *c = &chains[ci];
        const orc_seed_t *cs = seeds + c->seed_off;
        const int l_query = len[c->seqid];
        int64_t r0 = l_pac << 1, r1 = 0;
        for (int i = 0; i < c->n; ++i) {                        /* :2892-2905 */
            const orc_seed_t *t = &cs[i];
            int64_t b = t->rbeg - (t->qbeg + cal_max_gap(o, t->qbeg));
            int64_t e = t->rbeg + t->len + ((l_query - t->qbeg - t->len) + cal_max_gap(o, l_query - t->qbeg - t->len));
            r0 = r0 < b ? r0 : b;
            r1 = r1 > e ? r1 : e;
        }
        r0 = r0 > 0 ? r0 : 0;
        r1 = r1 < l_pac << 1 ? r1 : l_pac << 1;
        if (c->n > 0 && r0 < l_pac && l_pac < r1) { if (cs[0].rbeg < l_pac) r1 = l_pac; else r0 = l_pac; }
        if (c->n > 0) {                                          /* bns_fetch_seq_v2: clip to the contig (bntseq.cpp:484-517) */
            int64_t fb = contig_off[c->rid], fe = fb + contig_len[c->rid];
            if (cs[0].rbeg >= l_pac) { int64_t t = fb; fb = (l_pac << 1) - fe; fe = (l_pac << 1) - t; }
            r0 = r0 > fb ? r0 : fb;
            r1 = r1 < fe ? r1 : fe;
        }
        rmax0[ci] = r0; rmax1[ci] = r1;
        uint64_t *srt = malloc(sizeof(uint64_t) * (size_t)(c->n + 1));
        for (int i = 0; i < c->n; ++i) srt[i] = (uint64_t)cs[i].score << 32 | (uint32_t)i;
        qsort(srt, (size_t)c->n, 8, cmp_u64);                  /* keys are unique: any sort = ks_introsort_64 */
        for (int i = 0; i < c->n; ++i) srt_all[c->seed_off + i] = (uint32_t)srt[i];
        free(srt);
        for (int i = 0; i < c->n; ++i) {
            const orc_seed_t *s = &cs[i];
            if (s->qbeg) { lq += s->qbeg; lr += s->rbeg - r0; }
            if (s->qbeg + s->len != l_query) { rq += l_query - (s->qbeg + s->len); rr += r1 - r0 - (s->rbeg + s->len - r0); }
        }
    }
    uint8_t *LQ = malloc((size_t)lq + 64), *LR = malloc((size_t)lr + 64), *RQ = malloc((size_t)rq + 64), *RR = malloc((size_t)rr + 64);
    lq = lr = rq = rr = 0;
    for (int64_t ci = 0; ci < n_chains; ++ci) {
        const orc_chain_t *c = &chains[ci];
        const orc_seed_t *cs = seeds + c->seed_off;
        const uint8_t *query = reads + off[c->seqid];
        const int l_query = len[c->seqid];
        const int64_t r0 = rmax0[ci], r1 = rmax1[ci];
        const uint8_t *rseq = ref_string + r0;
        for (int k = c->n - 1; k >= 0; --k) {                   /* :2953-3190 */
            const orc_seed_t *s = &cs[srt_all[c->seed_off + k]];
            const int64_t ri = c->seed_off + (c->n - 1 - k);    /* region index: creation order */
            orc_alnreg_t *a = &regs[ri];
            memset(a, 0, sizeof *a);
            a->w = o->w; a->score = a->truesc = -1; a->rid = c->rid; a->frac_rep = c->frac_rep;
            a->seedlen0 = s->len; a->chain = (int32_t)ci;
            a->rb = a->qb = a->re = a->qe = H0_;
            if (s->qbeg) {
                orc_seqpair_t sp; memset(&sp, 0, sizeof sp);
                sp.h0 = s->len * o->a; sp.seqid = c->seqid; sp.id = (int32_t)ri;
                sp.idq = (int32_t)lq; sp.idr = (int32_t)lr;
                for (int i = 0; i < s->qbeg; ++i) LQ[lq + i] = query[s->qbeg - 1 - i];
                const int64_t tmp = s->rbeg - r0;
                for (int64_t i = 0; i < tmp; ++i) LR[lr + i] = rseq[tmp - 1 - i];
                lq += s->qbeg; lr += tmp;
                sp.len2 = s->qbeg; sp.len1 = (int32_t)tmp;
                L[nL++] = sp;
                a->qb = s->qbeg; a->rb = s->rbeg;
            } else { a->score = a->truesc = s->len * o->a; a->qb = 0; a->rb = s->rbeg; }
            if (s->qbeg + s->len != l_query) {
                const int64_t qe = s->qbeg + s->len, re = s->rbeg + s->len - r0;
                orc_seqpair_t sp; memset(&sp, 0, sizeof sp);
                sp.h0 = H0_; sp.seqid = c->seqid; sp.id = (int32_t)ri;
                sp.len2 = (int32_t)(l_query - qe); sp.len1 = (int32_t)(r1 - r0 - re);
                sp.idq = (int32_t)rq; sp.idr = (int32_t)rr;
                for (int i = 0; i < sp.len2; ++i) RQ[rq + i] = query[qe + i];
                for (int i = 0; i < sp.len1; ++i) RR[rr + i] = rseq[re + i];
                rq += sp.len2; rr += sp.len1;
                R[nR++] = sp;
                a->qe = (int32_t)qe; a->re = r0 + re;
            } else {
                a->qe = l_query; a->re = s->rbeg + s->len;
                if (a->rb != H0_ && a->qb != H0_) seedcov(a, c, cs);
            }
        }
    }
    if (cells) {
        orc_bsw_params_t P = { o->o_del, o->e_del, o->o_ins, o->e_ins, o->zdrop, o->pen_clip5, o->a, o->b };
        (void)P;
    }
    run_side(o, 1, L, (int)nL, LR, LQ, regs, chains, seeds, len);
    for (int64_t i = 0; i < nR; ++i) R[i].h0 = regs[R[i].id].score;          /* :3423-3428 */
    run_side(o, 0, R, (int)nR, RR, RQ, regs, chains, seeds, len);

    /* containment filter (:3634-3745): per read, chains in order, seeds by descending score */
    for (int64_t c0 = 0; c0 < n_chains;) {
        int64_t c1 = c0;
        while (c1 < n_chains && chains[c1].seqid == chains[c0].seqid) ++c1;
        const int l_query = len[chains[c0].seqid];
        const int64_t rbase = chains[c0].seed_off;
        int64_t n_regs = 0;
        for (int64_t c = c0; c < c1; ++c) n_regs += chains[c].n;
        orc_alnreg_t *av = regs + rbase;
        int lim = 0;
        uint32_t *purged = calloc((size_t)n_regs + 1, sizeof(uint32_t));     /* srt2[k] == UINT_MAX */
        for (int64_t ci = c0; ci < c1; ++ci) {
            const orc_chain_t *c = &chains[ci];
            const orc_seed_t *cs = seeds + c->seed_off;
            const uint32_t *srt2 = srt_all + c->seed_off;
            uint32_t *pg = purged + (c->seed_off - rbase);
            for (int k = c->n - 1; k >= 0; --k) {
                const orc_seed_t *s = &cs[srt2[k]];
                int i, v = 0;
                for (i = 0; i < n_regs && v < lim; ++i) {
                    const orc_alnreg_t *p = &av[i];
                    if (p->qb == -1 && p->qe == -1) continue;
                    int64_t rd; int qd, w, max_gap;
                    if (s->rbeg < p->rb || s->rbeg + s->len > p->re || s->qbeg < p->qb || s->qbeg + s->len > p->qe) { v++; continue; }
                    if (s->len - p->seedlen0 > .1 * l_query) { v++; continue; }
                    qd = s->qbeg - p->qb; rd = s->rbeg - p->rb;
                    max_gap = cal_max_gap(o, qd < rd ? qd : (int)rd);
                    w = max_gap < p->w ? max_gap : p->w;
                    if (qd - rd < w && rd - qd < w) break;
                    qd = p->qe - (s->qbeg + s->len); rd = p->re - (s->rbeg + s->len);
                    max_gap = cal_max_gap(o, qd < rd ? qd : (int)rd);
                    w = max_gap < p->w ? max_gap : p->w;
                    if (qd - rd < w && rd - qd < w) break;
                    v++;
                }
                if (v < lim) {
                    for (v = k + 1; v < c->n; ++v) {
                        if (pg[v]) continue;
                        const orc_seed_t *t = &cs[srt2[v]];
                        if (t->len < s->len * .95) continue;
                        if (s->qbeg <= t->qbeg && s->qbeg + s->len - t->qbeg >= s->len >> 2 && t->qbeg - s->qbeg != t->rbeg - s->rbeg) break;
                        if (t->qbeg <= s->qbeg && t->qbeg + t->len - s->qbeg >= s->len >> 2 && s->qbeg - t->qbeg != s->rbeg - t->rbeg) break;
                    }
                    if (v == c->n) {
                        orc_alnreg_t *ar = &regs[c->seed_off + (c->n - 1 - k)];
                        ar->qb = ar->qe = -1;
                        pg[k] = 1;
                        continue;
                    }
                }
                lim++;
            }
        }
        free(purged);
        c0 = c1;
    }
    free(L); free(R); free(LQ); free(LR); free(RQ); free(RR); free(srt_all); free(rmax0); free(rmax1);
}
