/* TEST INFRASTRUCTURE ONLY — see oracle.h.
 * CPU restatement of the banded Smith-Waterman extension kernels of the reference:
 *   orc_bsw_scalar : scalarBandedSWA                (src/bandedSWA.cpp:116-237)
 *   orc_bsw_lane   : one lane of the SSE4.1 kernels  (src/bandedSWA.cpp:3448-4121 16-bit,
 *                                                     4172-4837 8-bit)
 * Written from the algorithm, not transliterated: plain int arithmetic per pair.
 */
#include <stdlib.h>
#include <string.h>
#include "oracle.h"

static inline int imax(int a, int b) { return a > b ? a : b; }
static inline int imin(int a, int b) { return a < b ? a : b; }

/* substitution score as the 5x5 matrix of bwa_fill_scmat gives it (src/bwa.cpp:368-378):
 * +a on the diagonal of ACGT, -b off it, -1 whenever either symbol is N (>=4). */
static inline int sub_score(const orc_bsw_params_t *p, int t, int q)
{
    if (t > 3 || q > 3) return -1;
    return t == q ? p->w_match : -p->w_mismatch;
}

int orc_bsw_scalar(const orc_bsw_params_t *p, int qlen, const uint8_t *query, int tlen,
                   const uint8_t *target, int w, int h0, int *o_qle, int *o_tle, int *o_gtle,
                   int *o_gscore, int *o_max_off, uint64_t *cells)
{
    const int gapo_d = p->o_del + p->e_del, gapo_i = p->o_ins + p->e_ins;
    int *H = (int *)calloc((size_t)qlen + 2, sizeof(int));   /* H(i-1, j-1) per column */
    int *E = (int *)calloc((size_t)qlen + 2, sizeof(int));   /* E(i, j)              */
    uint64_t ncell = 0;

    /* first row: the seed score decays by an insertion run (bandedSWA.cpp:140-143) */
    H[0] = h0;
    if (qlen >= 1) H[1] = h0 > gapo_i ? h0 - gapo_i : 0;
    for (int j = 2; j <= qlen && H[j - 1] > p->e_ins; ++j) H[j] = H[j - 1] - p->e_ins;

    /* clip the band by the longest gap a full-length match could pay for (:145-154) */
    int best_sub = imax(imax(p->w_match, -p->w_mismatch), 0);   /* max over mat[], incl. -1 */
    {
        int g = (int)((double)(qlen * best_sub + p->end_bonus - p->o_ins) / p->e_ins + 1.);
        w = imin(w, imax(g, 1));
        g = (int)((double)(qlen * best_sub + p->end_bonus - p->o_del) / p->e_del + 1.);
        w = imin(w, imax(g, 1));
    }

    int best = h0, best_i = -1, best_j = -1, g_i = -1, g_sc = -1, off = 0;
    int lo = 0, hi = qlen;
    for (int i = 0; i < tlen; ++i) {
        int t = target[i];
        if (lo < i - w) lo = i - w;
        if (hi > i + w + 1) hi = i + w + 1;
        if (hi > qlen) hi = qlen;
        int left = 0;                                   /* H(i, j-1) */
        if (lo == 0) left = imax(h0 - (p->o_del + p->e_del * (i + 1)), 0);
        int f = 0, rowmax = 0, rowarg = -1, j;
        for (j = lo; j < hi; ++j) {
            int diag = H[j], e = E[j];
            H[j] = left;
            int M = diag ? diag + sub_score(p, t, query[j]) : 0;
            int h = imax(imax(M, e), f);
            left = h;
            if (h >= rowmax) rowarg = j;                /* ties keep the LAST column (:187) */
            rowmax = imax(rowmax, h);
            E[j] = imax(e - p->e_del, imax(M - gapo_d, 0));
            f = imax(f - p->e_ins, imax(M - gapo_i, 0));
            ++ncell;
        }
        H[hi] = left; E[hi] = 0;
        if (j == qlen) {                                /* reached the query end (:202-205) */
            if (!(g_sc > left)) g_i = i;
            g_sc = imax(g_sc, left);
        }
        if (rowmax == 0) break;
        if (rowmax > best) {
            best = rowmax; best_i = i; best_j = rowarg;
            off = imax(off, abs(rowarg - i));
        } else if (p->zdrop > 0) {
            int di = i - best_i, dj = rowarg - best_j;
            if (di > dj) { if (best - rowmax - (di - dj) * p->e_del > p->zdrop) break; }
            else         { if (best - rowmax - (dj - di) * p->e_ins > p->zdrop) break; }
        }
        for (j = lo; j < hi && H[j] == 0 && E[j] == 0; ++j) {}
        lo = j;
        for (j = hi; j >= lo && H[j] == 0 && E[j] == 0; --j) {}
        hi = imin(j + 2, qlen);
    }
    free(H); free(E);
    if (o_qle) *o_qle = best_j + 1;
    if (o_tle) *o_tle = best_i + 1;
    if (o_gtle) *o_gtle = g_i + 1;
    if (o_gscore) *o_gscore = g_sc;
    if (o_max_off) *o_max_off = off;
    if (cells) *cells += ncell;
    return best;
}

/* ------------------------------------------------------------------------------------------
 * One lane of the inter-task SIMD kernels.  State names follow the algorithm:
 *   Hd[j] : value the row above left for the diagonal of column j (H_h upstream)
 *   Fc[j] : per-column gap state carried down the rows, uses o_del/e_del (F upstream)
 *   er    : per-row gap state carried along the row, uses o_ins/e_ins (e11 upstream)
 *   head/tail : this lane's private live column range; lo/hi: the vector-wide loop bounds
 * 8-bit specifics are confined to `wrap` (band arithmetic in uint8, bandedSWA.cpp:4384-4401)
 * and two mask differences noted inline.  The 8-bit model is only meaningful for pairs the
 * host routes there (all values < 128, bwamem.cpp:2689).
 * ------------------------------------------------------------------------------------------ */
int orc_bsw_lane(const orc_bsw_params_t *p, int bits, int len2, const uint8_t *query, int len1,
                 const uint8_t *target, int w, int h0, int nrow_ext, int ncol_ext,
                 int *o_qle, int *o_tle, int *o_gtle, int *o_gscore, int *o_max_off)
{
    const int wrap = bits == 8 ? 0xFF : 0xFFFF;
    const int nrow = imax(nrow_ext, len1), ncol = imax(ncol_ext, len2);
    const int gapo_d = p->o_del + p->e_del, gapo_i = p->o_ins + p->e_ins;
    const int sc_match = p->w_match, sc_mis = -p->w_mismatch, sc_amb = -1;
    const int PAD_T = 99, PAD_Q = 100, NSYM = wrap;     /* DUMMY1/DUMMY2, N -> all-ones */

    int *Hd = (int *)calloc((size_t)ncol + 3, sizeof(int));
    int *Fc = (int *)calloc((size_t)ncol + 3, sizeof(int));
    int *Hv = (int *)calloc((size_t)nrow + 3, sizeof(int));   /* first-column values */

    /* wrapper: first column / first row initialisation (bandedSWA.cpp:3580-3650) */
    Hv[nrow] = PAD_T;
    for (int k = len1; k < nrow; ++k) Hv[k] = PAD_T;
    Hv[0] = h0;
    {
        int t = h0 - p->o_del;
        for (int k = 1; k < nrow; ++k) { t -= p->e_del; Hv[k] = imax(t, 0); }
    }
    Hd[0] = h0;
    Hd[1] = h0 > gapo_i ? h0 - gapo_i : 0;
    for (int k = 2; k < ncol; ++k) Hd[k] = imax(Hd[k - 1] - p->e_ins, 0);

    /* wrapper: per-lane band (bandedSWA.cpp:3652-3679): integer divide, unsigned wrap */
    int best_sub = imax(imax(sc_match, sc_mis), imax(sc_amb, 0));
    int qmax = (len2 * best_sub) & wrap;
    int band;
    {
        int t = (qmax + p->end_bonus - p->o_ins) & wrap;
        band = imin(w, imax(t / p->e_ins + 1, 1));
        t = (qmax + p->end_bonus - p->o_del) & wrap;
        band = imin(band, imax(t / p->e_del + 1, 1));
        band &= wrap;
    }

    int head = 0, tail = len2;
    int mlen = imin((len2 + band) & wrap, len1);
    int best = h0, bx = 0, by = 0, gsc = -1, gi = 0, off = 0, live = 1;
    int lo, hi, nlo = 0, nhi = ncol;

    for (int i = 0; i < nrow; ++i) {
        int t = i < len1 ? (target[i] == 4 ? NSYM : target[i]) : PAD_T;
        lo = nlo; hi = nhi;
        if (lo < i - w) lo = i - w;
        if (hi > i + w + 1) hi = i + w + 1;
        if (hi > ncol) hi = ncol;
        int left = lo == 0 ? Hv[i + 1] : 0;
        int er = 0, rowmax = 0, rowarg = 0;

        int phead = head, ptail = tail;
        head = imax(head, i - band);
        tail = imin(imin(tail, i + 1 + band), len2);
        if (head != phead || tail != ptail) {
            /* trimming pre-pass; alone in the vector the loop stops once l reaches head */
            for (int l = lo; l < hi && head > l; ++l) { Hd[l] = 0; Fc[l] = 0; }
        }
        if (i + 1 > mlen || head >= tail) live = 0;

        int j;
        for (j = lo; j < hi; ++j) {
            int q = j < len2 ? (query[j] == 4 ? NSYM : query[j]) : PAD_Q;
            int diag = Hd[j], fc = Fc[j];
            int s = t == q ? sc_match : sc_mis;
            if (t == NSYM || q == NSYM) s = sc_amb;
            int M = diag ? diag + s : 0;
            if (bits == 8 && M < 0) M = 0;
            int h = imax(imax(M, er), fc);
            er = imax(imax(M - gapo_i, 0), er - p->e_ins);
            int fnew = imax(imax(M - gapo_d, 0), fc - p->e_del);
            int masked = head > j || j > tail;
            if (masked) { left = 0; fnew = 0; }
            /* row maximum, ties -> last column; columns past the lane's tail do not count */
            int skip = bits == 8 ? masked : (j + 1 > tail);
            if (!skip) {
                if (h >= rowmax) rowarg = j + 1;
                rowmax = imax(rowmax, h);
            }
            Fc[j] = fnew; Hd[j] = left; left = h;
            if (j + 1 == len2 && live && !(j + 1 > tail)) {
                if (!(gsc > h)) gi = i + 1;
                gsc = imax(gsc, h);
            }
        }
        Hd[j] = left; Fc[j] = 0;

        if (rowmax == 0) {
            live = 0;
            if (nrow_ext <= len1 && ncol_ext <= len2) break;   /* alone: whole vector is dead */
        }
        int prev = best;
        if (live) best = imax(best, rowmax);
        if (best > prev) { by = rowarg; bx = i + 1; off = imax(off, abs(rowarg - (i + 1))); }
        {   /* z-drop WITHOUT the gap-extension factor (bandedSWA.cpp:3350-3363) */
            int di = (i + 1) - bx, dj = rowarg - by;
            int gap = di > dj ? di - dj : dj - di;
            if ((best - rowmax) - gap > p->zdrop) live = 0;
        }

        /* vector-wide bounds for the next row (single-lane view: this lane's zeros) */
        for (int l = lo; l < hi; ++l) { if ((Fc[l] | Hd[l]) == 0) nlo = l; else break; }
        {
            int l;
            for (l = hi; l >= lo; --l) if ((Fc[l] | Hd[l]) != 0) break;
            nhi = imin(l + 2, ncol);
        }
        if (live) {
            int l;
            for (l = lo; l < hi && (Fc[l] | Hd[l]) == 0; ++l) {}
            if (l > lo) head = l;
            int idx = tail;
            for (l = hi; l >= lo && (Fc[l] | Hd[l]) == 0; --l) idx = l - 1;
            tail = imin(idx + 2, len2);
        }
    }
    free(Hd); free(Fc); free(Hv);
    if (o_qle) *o_qle = by;
    if (o_tle) *o_tle = bx;
    if (o_gtle) *o_gtle = gi;
    if (o_gscore) *o_gscore = gsc;
    if (o_max_off) *o_max_off = off;
    return best;
}

uint64_t orc_bsw_batch(const orc_bsw_params_t *p, int kind, orc_seqpair_t *pairs, int32_t n,
                       const uint8_t *ref_buf, const uint8_t *qer_buf, int32_t w)
{
    uint64_t cells = 0;
    for (int32_t i = 0; i < n; ++i) {
        orc_seqpair_t *sp = &pairs[i];
        const uint8_t *t = ref_buf + sp->idr, *q = qer_buf + sp->idq;
        if (kind == 0) {
            sp->score = orc_bsw_scalar(p, sp->len2, q, sp->len1, t, w, sp->h0, &sp->qle, &sp->tle,
                                       &sp->gtle, &sp->gscore, &sp->max_off, &cells);
        } else {
            sp->score = orc_bsw_lane(p, kind, sp->len2, q, sp->len1, t, w, sp->h0, 0, 0, &sp->qle,
                                     &sp->tle, &sp->gtle, &sp->gscore, &sp->max_off);
        }
    }
    return cells;
}
