/* TEST INFRASTRUCTURE ONLY — see oracle.h.
 * CPU restatement of the reference's mate-rescue local Smith-Waterman:
 *   ksw_qinit   src/ksw.cpp:62-109   striped query layout, shift / mdiff / max
 *   ksw_u8      src/ksw.cpp:111-232  16 unsigned saturating byte lanes
 *   ksw_i16     src/ksw.cpp:234-338  8 signed 16-bit lanes
 *   ksw_align2  src/ksw.cpp:347-381  forward pass, then the reversed pass for (tb, qb)
 * The reference is Farrar's striped kernel on 128-bit vectors in EVERY build (ksw.cpp uses __m128i
 * only).  Its result is not plain affine-gap SW: E(i+1,j) is derived from H(i,j) BEFORE the lazy-F
 * correction ("we disallow adjacent insertion and then deletion", :168), and which cells the
 * lazy-F pass corrects depends on the striping.  So this restatement keeps the striped order: P
 * lanes (16 or 8), slen = ceil(qlen/P) segments, query position of (lane l, segment j) = j + l*slen.
 */
#include <stdlib.h>
#include <string.h>
#include "oracle.h"

#define KSW_XBYTE  0x10000
#define KSW_XSTOP  0x20000
#define KSW_XSUBO  0x40000
#define KSW_XSTART 0x80000

typedef struct { int qlen, slen, P, size, shift, mdiff, max; const uint8_t *query; const int8_t *mat; int m; } qprof_t;

static void qinit(qprof_t *q, int size, int qlen, const uint8_t *query, int m, const int8_t *mat)
{
    q->size = size > 1 ? 2 : 1;
    q->P = 8 * (3 - q->size);
    q->slen = (qlen + q->P - 1) / q->P;
    q->qlen = qlen; q->query = query; q->mat = mat; q->m = m;
    int sh = 127, md = 0;
    for (int a = 0; a < m * m; ++a) {                  /* :79-82, with the reference's int8 casts */
        if (mat[a] < (int8_t)sh) sh = mat[a];
        if (mat[a] > (int8_t)md) md = mat[a];
    }
    q->max = (uint8_t)md;
    q->shift = (uint8_t)(256 - sh);
    q->mdiff = (uint8_t)(md + q->shift);
}

static inline int sub_score(const qprof_t *q, int t, int k)   /* profile value before the shift */
{
    return k >= q->qlen ? 0 : q->mat[t * q->m + q->query[k]];
}

static inline int sat_u8(int v) { return v < 0 ? 0 : v > 255 ? 255 : v; }
static inline int sat_i16(int v) { return v < -32768 ? -32768 : v > 32767 ? 32767 : v; }
static inline int subs_u16(int a, int b) { int v = (a & 0xFFFF) - (b & 0xFFFF); return v < 0 ? 0 : v; }

/* one pass of ksw_u8 / ksw_i16 over `target`; r must hold g_defr on entry */
static void ksw_pass(const qprof_t *q, int tlen, const uint8_t *target, int o_del, int e_del, int o_ins,
                     int e_ins, int xtra, orc_kswr_t *r)
{
    const int P = q->P, slen = q->slen, n = P * slen, byte = q->size == 1;
    const int minsc = (xtra & KSW_XSUBO) ? xtra & 0xffff : 0x10000;
    const int endsc = (xtra & KSW_XSTOP) ? xtra & 0xffff : 0x10000;
    const int oe_del = byte ? (uint8_t)(o_del + e_del) : (int16_t)(o_del + e_del);
    const int oe_ins = byte ? (uint8_t)(o_ins + e_ins) : (int16_t)(o_ins + e_ins);
    const int ed = byte ? (uint8_t)e_del : (int16_t)e_del, ei = byte ? (uint8_t)e_ins : (int16_t)e_ins;
    int *H0 = calloc((size_t)n * 4, sizeof(int)), *H1 = H0 + n, *E = H1 + n, *Hmax = E + n;
    int *f = malloc(sizeof(int) * (size_t)P);
    uint64_t *b = NULL; int n_b = 0, m_b = 0;
    int gmax = 0, te = -1;
#define POS(l, j) ((j) + (l) * slen)
    for (int i = 0; i < tlen; ++i) {
        const int t = target[i];
        int imax = 0;
        for (int l = 0; l < P; ++l) {                  /* lanes are independent in the main pass */
            int fl = 0, h = l == 0 ? 0 : H0[POS(l - 1, slen - 1)];
            for (int j = 0; j < slen; ++j) {
                const int k = POS(l, j);
                int e = E[k];
                if (byte) {
                    h = sat_u8(h + ((sub_score(q, t, k) + q->shift) & 0xFF));
                    h = sat_u8(h - q->shift);
                } else {
                    h = sat_i16(h + sub_score(q, t, k));
                }
                if (e > h) h = e;
                if (fl > h) h = fl;
                if (h > imax) imax = h;
                H1[k] = h;
                if (byte) {
                    e = sat_u8(e - ed); int tt = sat_u8(h - oe_del); E[k] = e > tt ? e : tt;
                    fl = sat_u8(fl - ei); tt = sat_u8(h - oe_ins); if (tt > fl) fl = tt;
                } else {                               /* unsigned saturating subtract, signed max */
                    e = (int16_t)subs_u16(e, ed); int tt = (int16_t)subs_u16(h, oe_del); E[k] = e > tt ? e : tt;
                    fl = (int16_t)subs_u16(fl, ei); tt = (int16_t)subs_u16(h, oe_ins); if (tt > fl) fl = tt;
                }
                h = H0[k];
            }
            f[l] = fl;
        }
        /* lazy-F: up to 16 passes; each shifts f by one lane, then sweeps the segments */
        for (int kk = 0, done = 0; kk < 16 && !done; ++kk) {
            for (int l = P - 1; l > 0; --l) f[l] = f[l - 1];
            f[0] = 0;
            for (int j = 0; j < slen && !done; ++j) {
                int any = 0;
                for (int l = 0; l < P; ++l) {
                    const int k = POS(l, j);
                    int h = H1[k];
                    if (f[l] > h) h = f[l];
                    H1[k] = h;
                    if (byte) {
                        h = sat_u8(h - oe_ins); f[l] = sat_u8(f[l] - ei);
                        if (sat_u8(f[l] - h) != 0) any = 1;
                    } else {
                        h = (int16_t)subs_u16(h, oe_ins); f[l] = (int16_t)subs_u16(f[l], ei);
                        if (f[l] > h) any = 1;
                    }
                }
                if (!any) done = 1;
            }
        }
        if (imax >= minsc) {                           /* :190-198 */
            if (n_b == 0 || (int32_t)b[n_b - 1] + 1 != i) {
                if (n_b == m_b) { m_b = m_b ? m_b << 1 : 8; b = realloc(b, 8 * (size_t)m_b); }
                b[n_b++] = (uint64_t)imax << 32 | (uint32_t)i;
            } else if ((int)(b[n_b - 1] >> 32) < imax) b[n_b - 1] = (uint64_t)imax << 32 | (uint32_t)i;
        }
        if (imax > gmax) {
            gmax = imax; te = i;
            memcpy(Hmax, H1, sizeof(int) * (size_t)n);
            if (byte ? (gmax + q->shift >= 255 || gmax >= endsc) : (gmax >= endsc)) break;
        }
        int *S = H1; H1 = H0; H0 = S;
    }
    r->score = byte ? (gmax + q->shift < 255 ? gmax : 255) : gmax;
    r->te = te;
    if (!byte || r->score != 255) {
        int max = -1;
        if (!byte) r->qe = -1;
        for (int i = 0; i < n; ++i) {                  /* memory order of the Hmax vectors: segment-major */
            const int v = Hmax[POS(i % P, i / P)], tmp = i / P + i % P * slen;
            if (v > max) { max = v; r->qe = tmp; }
            else if (v == max && tmp < r->qe) r->qe = tmp;
        }
        if (b) {
            int w = (r->score + q->max - 1) / q->max;
            const int low = te - w, high = te + w;
            for (int i = 0; i < n_b; ++i) {
                const int e = (int32_t)b[i];
                if ((e < low || e > high) && (int)(b[i] >> 32) > r->score2) { r->score2 = (int)(b[i] >> 32); r->te2 = e; }
            }
        }
    }
#undef POS
    /* H0/H1 were swapped an unknown number of times: free the lower of the two */
    free(H0 < H1 ? H0 : H1); free(f); free(b);
}

static void revseq(int l, uint8_t *s)
{
    for (int i = 0; i < l >> 1; ++i) { uint8_t t = s[i]; s[i] = s[l - 1 - i]; s[l - 1 - i] = t; }
}

void orc_ksw_align2(int qlen, const uint8_t *query, int tlen, const uint8_t *target, int m, const int8_t *mat,
                    int o_del, int e_del, int o_ins, int e_ins, int xtra, orc_kswr_t *out)
{
    static const orc_kswr_t defr = { 0, -1, -1, -1, -1, -1, -1 };
    qprof_t q;
    orc_kswr_t r = defr, rr = defr;
    qinit(&q, (xtra & KSW_XBYTE) ? 1 : 2, qlen, query, m, mat);
    ksw_pass(&q, tlen, target, o_del, e_del, o_ins, e_ins, xtra, &r);
    *out = r;
    if ((xtra & KSW_XSTART) == 0 || ((xtra & KSW_XSUBO) && r.score < (xtra & 0xffff))) return;
    /* the reference reverses query[0..qe] and target[0..te] in place, runs the same kernel over the
     * WHOLE target with XSTOP|score, and reverses back (:367-375) */
    uint8_t *qq = malloc((size_t)qlen + 1), *tt = malloc((size_t)tlen + 1);
    memcpy(qq, query, (size_t)qlen); memcpy(tt, target, (size_t)tlen);
    revseq(r.qe + 1, qq); revseq(r.te + 1, tt);
    qinit(&q, q.size, r.qe + 1, qq, m, mat);
    ksw_pass(&q, tlen, tt, o_del, e_del, o_ins, e_ins, KSW_XSTOP | r.score, &rr);
    free(qq); free(tt);
    if (r.score == rr.score) { r.tb = r.te - rr.te; r.qb = r.qe - rr.qe; }
    *out = r;
}

void orc_ksw_batch(const orc_seqpair_t *pairs, int32_t n, const uint8_t *ref_buf, const uint8_t *qer_buf,
                   const int8_t *mat, int o_del, int e_del, int o_ins, int e_ins, orc_kswr_t *aln)
{
    for (int32_t i = 0; i < n; ++i)
        orc_ksw_align2(pairs[i].len2, qer_buf + pairs[i].idq, pairs[i].len1, ref_buf + pairs[i].idr, 5, mat,
                       o_del, e_del, o_ins, e_ins, pairs[i].h0, &aln[i]);
}
