// CIGAR / NM / MD generation for a final alignment region: the alignment core of mem_reg2aln (reference
// src/bwamem.cpp:2533-2570: band inference, up to three bwa_gen_cigar2 tries with a doubling band), bwa_gen_cigar2
// (src/bwa.cpp:380-467: strand handling, gap-free shortcut, band choice, NM and MD) and ksw_global2
// (src/ksw.cpp:558-668: banded global alignment with a byte-per-cell backtrack matrix).  SURVEY.md §8f rank 3.
// Per-thread code shared by the CUDA kernel in cigar.cu and, compiled as host C++, by tests/hostsim (test
// scaffolding, not a product path).
#pragma once
#include <stdint.h>
#include "../../include/bwagpu.h"

#if defined(__CUDACC__)
#define CIG_HD __host__ __device__
#else
#define CIG_HD
#endif

#define CIG_MINUS_INF (-0x40000000)

struct CigarOpt { int8_t mat[25]; int32_t a, o_del, e_del, o_ins, e_ins, w; };

struct CigarScratch {
    int32_t *eh;        // 2 x (qlen + 1) words, element k at eh[k * eh_stride]: H and E rows of ksw_global2
    int eh_stride;      // 1 for a private slab; blockDim for the interleaved shared-memory layout (bank = thread)
    int eh_mask;        // rows are a circular window of eh_mask + 1 columns (a power of two >= 2 * band + 2: the DP
                        // only ever touches columns beg..end of a row); 0x7fffffff = full rows of qlen + 1 columns
    uint8_t *z;         // backtrack matrix, n_col x tlen for the widest band tried
    uint32_t *cigar;    // qlen + tlen + 2 entries
    char *md;           // MD text, 3 x tlen + 16 bytes
};

struct CigarResult { int32_t score, n_cigar, nm, l_md; };

CIG_HD inline int cig_infer_bw(int l1, int l2, int score, int a, int q, int r)          // src/bwamem.cpp:2630-2637
{
    if (l1 == l2 && l1 * a - score < (q + r - a) << 1) return 0;
    int w = (int)((double)((l1 < l2 ? l1 : l2) * a - score - q) / r + 2.);
    const int d = l1 > l2 ? l1 - l2 : l2 - l1;
    if (w < d) w = d;
    return w;
}

CIG_HD inline void cig_push(uint32_t *cigar, int *n, int op, int len)                   // push_cigar, src/ksw.cpp:546-556
{
    if (*n == 0 || op != (int)(cigar[*n - 1] & 0xf)) cigar[(*n)++] = (uint32_t)len << 4 | (uint32_t)op;
    else cigar[*n - 1] += (uint32_t)len << 4;
}

CIG_HD inline void cig_putw(char *s, int *l, int c)                                      // kputw, src/kstring.h:92-107
{
    if (c == 0) { s[(*l)++] = '0'; return; }
    char buf[16];
    int n = 0;
    for (int x = c; x > 0; x /= 10) buf[n++] = (char)(x % 10 + '0');
    for (int x = n - 1; x >= 0; --x) s[(*l)++] = buf[x];
}

// query / target accessors: both sequences are read through a direction flag instead of being reversed in place
struct CigarSeqs {
    const uint8_t *q; int lq;          // read bases qb..qe (codes 0-3, >= 4 = N -> 4)
    const uint8_t *t; int lt;          // text window [rb, re)
    bool rev;                          // rb >= l_pac: both are walked backwards (src/bwa.cpp:394-399)
    CIG_HD int qa(int i) const { const int b = q[rev ? lq - 1 - i : i]; return b < 4 ? b : 4; }
    CIG_HD int ta(int i) const { return t[rev ? lt - 1 - i : i]; }
};

// ksw_global2 with m = 5; returns the score, fills cigar / n_cigar
CIG_HD inline int cig_global2(const CigarOpt &o, const CigarSeqs &s, int w, const CigarScratch &sc, int *n_cigar)
{
    const int qlen = s.lq, tlen = s.lt;
    const int oe_del = o.o_del + o.e_del, oe_ins = o.o_ins + o.e_ins;
    // row stride of the backtrack matrix, padded to whole 32-bit words: a row's bytes are stored four at a time
    const int n_col = ((qlen < 2 * w + 1 ? qlen : 2 * w + 1) + 3) & ~3;
    const int st = sc.eh_stride, mk = sc.eh_mask;
    const int rows = mk == 0x7fffffff ? qlen + 1 : mk + 1;
    int32_t *H = sc.eh, *E = sc.eh + (int64_t)rows * st;
#define CIG_IX(j) (((j) & mk) * st)
    H[0] = 0; E[0] = CIG_MINUS_INF;
    int j;
    for (j = 1; j <= qlen && j <= w; ++j) { H[CIG_IX(j)] = -(o.o_ins + o.e_ins * j); E[CIG_IX(j)] = CIG_MINUS_INF; }
    for (; j <= qlen && j < rows; ++j) H[CIG_IX(j)] = E[CIG_IX(j)] = CIG_MINUS_INF;   // beyond the window: written before read
    for (int i = 0; i < tlen; ++i) {
        int32_t f = CIG_MINUS_INF, h1;
        const int8_t *mrow = o.mat + s.ta(i) * 5;
        const int beg = i > w ? i - w : 0;
        const int end = i + w + 1 < qlen ? i + w + 1 : qlen;
        h1 = beg == 0 ? -(o.o_del + o.e_del * (i + 1)) : CIG_MINUS_INF;
        uint32_t *zw = (uint32_t *)(sc.z + (int64_t)i * n_col);
        uint32_t acc = 0;
        int nb = 0;
        for (j = beg; j < end; ++j) {
            int32_t m = H[CIG_IX(j)], e = E[CIG_IX(j)], h, t;
            uint8_t d;
            H[CIG_IX(j)] = h1;
            m += mrow[s.qa(j)];
            d = m >= e ? 0 : 1;
            h = m >= e ? m : e;
            d = h >= f ? d : 2;
            h = h >= f ? h : f;
            h1 = h;
            t = m - oe_del;
            e -= o.e_del;
            d |= e > t ? 1 << 2 : 0;
            e = e > t ? e : t;
            E[CIG_IX(j)] = e;
            t = m - oe_ins;
            f -= o.e_ins;
            d |= f > t ? 2 << 4 : 0;
            f = f > t ? f : t;
            acc |= (uint32_t)d << (8 * nb);
            if (++nb == 4) { *zw++ = acc; acc = 0; nb = 0; }
        }
        if (nb) *zw = acc;
        H[CIG_IX(end)] = h1; E[CIG_IX(end)] = CIG_MINUS_INF;
    }
    const int score = H[CIG_IX(qlen)];
#undef CIG_IX
    int n = 0, which = 0;
    int i = tlen - 1, k = (i + w + 1 < qlen ? i + w + 1 : qlen) - 1;
    while (i >= 0 && k >= 0) {
        which = sc.z[(int64_t)i * n_col + (k - (i > w ? i - w : 0))] >> (which << 1) & 3;
        if (which == 0) { cig_push(sc.cigar, &n, 0, 1); --i; --k; }
        else if (which == 1) { cig_push(sc.cigar, &n, 2, 1); --i; }
        else { cig_push(sc.cigar, &n, 1, 1); --k; }
    }
    if (i >= 0) cig_push(sc.cigar, &n, 2, i + 1);
    if (k >= 0) cig_push(sc.cigar, &n, 1, k + 1);
    for (i = 0; i < n >> 1; ++i) { const uint32_t tmp = sc.cigar[i]; sc.cigar[i] = sc.cigar[n - 1 - i]; sc.cigar[n - 1 - i] = tmp; }
    *n_cigar = n;
    return score;
}

// bwa_gen_cigar2; returns 0 where the reference returns a NULL cigar (empty / strand-bridging / out-of-range window), 1 when
// done, 2 when this try's band needs more row columns than the scratch window holds (the caller re-runs the region with full rows)
CIG_HD inline int cig_gen_cigar2(const CigarOpt &o, int w_, int64_t l_pac, const uint8_t *ref_string, const uint8_t *query,
                                  int l_query, int64_t rb, int64_t re, const CigarScratch &sc, CigarResult *res)
{
    res->n_cigar = 0; res->nm = -1; res->l_md = 0;
    if (l_query <= 0 || rb >= re || (rb < l_pac && re > l_pac)) return 0;
    if (rb < 0 || re > (l_pac << 1)) return 0;                        // bns_get_seq would clip: re - rb != rlen
    CigarSeqs s;
    s.q = query; s.lq = l_query; s.t = ref_string + rb; s.lt = (int)(re - rb); s.rev = rb >= l_pac;
    const int rlen = s.lt;
    int n_cigar = 0;
    if (l_query == rlen && w_ == 0) {
        sc.cigar[0] = (uint32_t)l_query << 4;
        n_cigar = 1;
        int sum = 0;
        for (int i = 0; i < l_query; ++i) sum += o.mat[s.ta(i) * 5 + s.qa(i)];
        res->score = sum;
    } else {
        int max_ins = (int)((double)(((l_query + 1) >> 1) * o.mat[0] - o.o_ins) / o.e_ins + 1.);
        int max_del = (int)((double)(((l_query + 1) >> 1) * o.mat[0] - o.o_del) / o.e_del + 1.);
        int max_gap = max_ins > max_del ? max_ins : max_del;
        max_gap = max_gap > 1 ? max_gap : 1;
        const int diff = rlen > l_query ? rlen - l_query : l_query - rlen;
        int w = (max_gap + diff + 1) >> 1;
        w = w < w_ ? w : w_;
        const int min_w = diff + 3;
        w = w > min_w ? w : min_w;
        if (sc.eh_mask != 0x7fffffff && 2 * w + 2 > sc.eh_mask + 1) return 2;
        res->score = cig_global2(o, s, w, sc, &n_cigar);
    }
    // NM and MD (src/bwa.cpp:429-459)
    int x = 0, y = 0, u = 0, n_mm = 0, n_gap = 0, l = 0;
    const char *int2base = rb < l_pac ? "ACGTN" : "TGCAN";
    for (int k = 0; k < n_cigar; ++k) {
        const int op = sc.cigar[k] & 0xf, len = (int)(sc.cigar[k] >> 4);
        if (op == 0) {
            for (int i = 0; i < len; ++i) {
                if (s.qa(x + i) != s.ta(y + i)) { cig_putw(sc.md, &l, u); sc.md[l++] = int2base[s.ta(y + i)]; ++n_mm; u = 0; }
                else ++u;
            }
            x += len; y += len;
        } else if (op == 2) {
            if (k > 0 && k < n_cigar - 1) {
                cig_putw(sc.md, &l, u); sc.md[l++] = '^';
                for (int i = 0; i < len; ++i) sc.md[l++] = int2base[s.ta(y + i)];
                u = 0; n_gap += len;
            }
            y += len;
        } else if (op == 1) { x += len; n_gap += len; }
    }
    cig_putw(sc.md, &l, u);
    res->n_cigar = n_cigar; res->nm = n_mm + n_gap; res->l_md = l;
    return 1;
}

// the alignment core of mem_reg2aln (src/bwamem.cpp:2551-2570): l_query = qe - qb bases of the read from qb
CIG_HD inline int cig_reg2aln(const CigarOpt &o, int64_t l_pac, const uint8_t *ref_string, const uint8_t *read, int qb, int qe,
                               int64_t rb, int64_t re, int truesc, int reg_w, const CigarScratch &sc, CigarResult *res)
{
    int tmp = cig_infer_bw(qe - qb, (int)(re - rb), truesc, o.a, o.o_del, o.e_del);
    int w2 = cig_infer_bw(qe - qb, (int)(re - rb), truesc, o.a, o.o_ins, o.e_ins);
    w2 = w2 > tmp ? w2 : tmp;
    if (w2 > o.w) w2 = w2 < reg_w ? w2 : reg_w;
    int i = 0, last_sc = -(1 << 30);
    int ok;
    do {
        w2 = w2 < o.w << 2 ? w2 : o.w << 2;
        ok = cig_gen_cigar2(o, w2, l_pac, ref_string, read + qb, qe - qb, rb, re, sc, res);
        if (ok != 1) return ok;
        if (res->score == last_sc || w2 == o.w << 2) break;
        last_sc = res->score;
        w2 <<= 1;
    } while (++i < 3 && res->score < truesc - o.a);
    return 1;
}

// true when cig_reg2aln takes the gap-free shortcut of bwa_gen_cigar2 (first band 0 and equal lengths; a zero band
// stays zero through the retries)
CIG_HD inline bool cig_is_gapfree(const CigarOpt &o, int qb, int qe, int64_t rb, int64_t re, int truesc, int reg_w)
{
    int tmp = cig_infer_bw(qe - qb, (int)(re - rb), truesc, o.a, o.o_del, o.e_del);
    int w2 = cig_infer_bw(qe - qb, (int)(re - rb), truesc, o.a, o.o_ins, o.e_ins);
    w2 = w2 > tmp ? w2 : tmp;
    if (w2 > o.w) w2 = w2 < reg_w ? w2 : reg_w;
    w2 = w2 < o.w << 2 ? w2 : o.w << 2;
    return w2 == 0 && (qe - qb) == (int)(re - rb);
}

// bytes of backtrack matrix the widest try of cig_reg2aln can need for this region: the band of try k is
// min((max_gap + diff + 1) >> 1, w2 << k) raised to diff + 3, k = 0..2, w2 << k capped at 4 * opt->w
CIG_HD inline int cig_band_max(const CigarOpt &o, int qb, int qe, int64_t rb, int64_t re, int truesc, int reg_w)
{
    const int lq = qe - qb, lt = (int)(re - rb);
    if (lq <= 0 || lt <= 0) return 0;
    int tmp = cig_infer_bw(lq, lt, truesc, o.a, o.o_del, o.e_del);
    int w2 = cig_infer_bw(lq, lt, truesc, o.a, o.o_ins, o.e_ins);
    w2 = w2 > tmp ? w2 : tmp;
    if (w2 > o.w) w2 = w2 < reg_w ? w2 : reg_w;
    int64_t wt = (int64_t)w2 << 2;
    if (wt > (o.w << 2)) wt = o.w << 2;
    const int max_ins = (int)((double)(((lq + 1) >> 1) * o.mat[0] - o.o_ins) / o.e_ins + 1.);
    const int max_del = (int)((double)(((lq + 1) >> 1) * o.mat[0] - o.o_del) / o.e_del + 1.);
    int max_gap = max_ins > max_del ? max_ins : max_del;
    max_gap = max_gap > 1 ? max_gap : 1;
    const int diff = lt > lq ? lt - lq : lq - lt;
    int64_t w = (max_gap + diff + 1) >> 1;
    w = w < wt ? w : wt;
    w = w > diff + 3 ? w : diff + 3;
    return (int)w;
}
CIG_HD inline int64_t cig_z_bytes(const CigarOpt &o, int qb, int qe, int64_t rb, int64_t re, int truesc, int reg_w)
{
    const int lq = qe - qb, lt = (int)(re - rb);
    if (lq <= 0 || lt <= 0) return 0;
    const int64_t w = cig_band_max(o, qb, qe, rb, re, truesc, reg_w);
    const int64_t n_col = ((lq < 2 * w + 1 ? lq : 2 * w + 1) + 3) & ~3;     // padded row stride, as cig_global2 lays it out
    return n_col * lt;
}
// window tier of a region, by the band of its FIRST try (later tries, rare, re-run with full rows): the smallest power of
// two W in {16, 32, 64, 128} with W >= 2 * band + 2, or 4 = full rows
CIG_HD inline int cig_tier(const CigarOpt &o, int qb, int qe, int64_t rb, int64_t re, int truesc, int reg_w)
{
    const int lq = qe - qb, lt = (int)(re - rb);
    if (lq <= 0 || lt <= 0) return 0;
    int tmp = cig_infer_bw(lq, lt, truesc, o.a, o.o_del, o.e_del);
    int w2 = cig_infer_bw(lq, lt, truesc, o.a, o.o_ins, o.e_ins);
    w2 = w2 > tmp ? w2 : tmp;
    if (w2 > o.w) w2 = w2 < reg_w ? w2 : reg_w;
    w2 = w2 < o.w << 2 ? w2 : o.w << 2;
    const int max_ins = (int)((double)(((lq + 1) >> 1) * o.mat[0] - o.o_ins) / o.e_ins + 1.);
    const int max_del = (int)((double)(((lq + 1) >> 1) * o.mat[0] - o.o_del) / o.e_del + 1.);
    int max_gap = max_ins > max_del ? max_ins : max_del;
    max_gap = max_gap > 1 ? max_gap : 1;
    const int diff = lt > lq ? lt - lq : lq - lt;
    int w = (max_gap + diff + 1) >> 1;
    w = w < w2 ? w : w2;
    w = w > diff + 3 ? w : diff + 3;
    const int need = 2 * w + 2;
    if (need >= lq + 1) return 4;
    return need <= 16 ? 0 : need <= 32 ? 1 : need <= 64 ? 2 : need <= 128 ? 3 : 4;
}

// the tail of mem_reg2aln (src/bwamem.cpp:2571-2600): position, removal of a leading / trailing deletion, soft clips.
// cigar[] must have room for n_cigar + 2 entries; returns the new n_cigar.
CIG_HD inline int cig_finish(int64_t l_pac, int64_t rb, int64_t re, int qb, int qe, int l_query, uint32_t *cigar, int n_cigar,
                             int64_t *pos_out, int *is_rev_out)
{
    const int64_t p0 = rb < l_pac ? rb : re - 1;
    const int is_rev = p0 >= l_pac;                                        // bns_depos, src/bntseq.h
    int64_t pos = is_rev ? (l_pac << 1) - 1 - p0 : p0;
    if (n_cigar > 0) {
        if ((cigar[0] & 0xf) == 2) {
            pos += cigar[0] >> 4;
            --n_cigar;
            for (int k = 0; k < n_cigar; ++k) cigar[k] = cigar[k + 1];
        } else if ((cigar[n_cigar - 1] & 0xf) == 2) --n_cigar;
    }
    if (qb != 0 || qe != l_query) {
        const int clip5 = is_rev ? l_query - qe : qb, clip3 = is_rev ? qb : l_query - qe;
        if (clip5) {
            for (int k = n_cigar; k > 0; --k) cigar[k] = cigar[k - 1];
            cigar[0] = (uint32_t)clip5 << 4 | 3;
            ++n_cigar;
        }
        if (clip3) cigar[n_cigar++] = (uint32_t)clip3 << 4 | 3;
    }
    *pos_out = pos; *is_rev_out = is_rev;
    return n_cigar;
}
