// Banded SW extension, one pair per thread ("generic" path).
//
// Exact per-pair semantics of the reference kernels:
//   bsw_lane_pair   : one SIMD lane of smithWaterman128_16 / _8 incl. its batch wrapper
//                     (reference src/bandedSWA.cpp:3448-4121, 4172-4837)
//   bsw_scalar_pair : scalarBandedSWA (reference src/bandedSWA.cpp:116-237)
// Column state lives in a caller-provided scratch, addressed through COL(j) so that the threads
// of a warp touch consecutive words when they sit on the same column.  This path handles any
// length; the warp-cooperative kernel (bsw_warp.cuh) covers the common short-read shapes.
#pragma once
#include <stdint.h>
#include "../../include/bwagpu.h"

struct BswOut { int score, qle, tle, gtle, gscore, max_off; int cells; };   // cells = sum over rows of (hi - lo): DP cells computed

__device__ __forceinline__ int bsw_sym(uint8_t c, int nsym) { return c == 4 ? nsym : (int)c; }

// scratch word: low 16 bits = Hd (diagonal carry), high 16 bits = Fc (column gap state)
__device__ __forceinline__ uint32_t bsw_pack(int hd, int fc)
{
    return ((uint32_t)hd & 0xFFFFu) | ((uint32_t)fc << 16);
}
__device__ __forceinline__ int bsw_hd(uint32_t v) { return (int)(int16_t)(v & 0xFFFFu); }
__device__ __forceinline__ int bsw_fc(uint32_t v) { return (int)(int16_t)(v >> 16); }

template <int BITS>
__device__ BswOut bsw_lane_pair(const bwagpu_bsw_params &p, const uint8_t *__restrict__ query,
                                int len2, const uint8_t *__restrict__ target, int len1, int w,
                                int h0, uint32_t *__restrict__ col, int64_t stride, int dir = 1)
{
    // dir: +1 = query[j] / target[i] ascend in memory (sequence arenas); -1 = they descend from the given bytes (a left
    // extension read straight out of the read / the resident reference text, see bsw_pair_view in bsw.cu)
#define COL(j) col[(int64_t)(j) * stride]
    const int wrap = BITS == 8 ? 0xFF : 0xFFFF;
    const int gapo_d = p.o_del + p.e_del, gapo_i = p.o_ins + p.e_ins;
    const int sc_match = p.w_match, sc_mis = -p.w_mismatch, sc_amb = -1;
    const int nrow = len1, ncol = len2;

    // first row: seed score decaying along an insertion run; columns ncol..ncol+2 start at 0
    {
        int v = h0;
        COL(0) = bsw_pack(v, 0);
        v = h0 > gapo_i ? h0 - gapo_i : 0;
        COL(1) = bsw_pack(v, 0);
        for (int k = 2; k < ncol; ++k) { v = max(v - p.e_ins, 0); COL(k) = bsw_pack(v, 0); }
        for (int k = max(ncol, 2); k < ncol + 3; ++k) COL(k) = 0;
    }
    // per-pair band: integer divide on the wrapped unsigned sum, as the SIMD wrapper does
    int band;
    {
        int best_sub = max(max(sc_match, sc_mis), max(sc_amb, 0));
        int qmax = (len2 * best_sub) & wrap;
        int t = (qmax + p.end_bonus - p.o_ins) & wrap;
        band = min(w, max(t / p.e_ins + 1, 1));
        t = (qmax + p.end_bonus - p.o_del) & wrap;
        band = min(band, max(t / p.e_del + 1, 1)) & wrap;
    }
    int head = 0, tail = len2;
    const int mlen = min((len2 + band) & wrap, len1);
    int best = h0, bx = 0, by = 0, gsc = -1, gi = 0, off = 0;
    bool live = true;
    int nlo = 0, nhi = ncol;
    int cells = 0;

    for (int i = 0; i < nrow; ++i) {
        const int t = bsw_sym(target[i * dir], wrap);
        int lo = max(nlo, i - w), hi = min(min(nhi, i + w + 1), ncol);
        cells += max(hi - lo, 0);
        // first-column value H(i,-1); the SIMD code reads a pad value on the very last row,
        // which can only reach state that is never read again
        int left = lo == 0 ? max(h0 - p.o_del - (i + 1) * p.e_del, 0) : 0;
        int er = 0, rowmax = 0, rowarg = 0;

        const int phead = head, ptail = tail;
        head = max(head, i - band);
        tail = min(min(tail, i + 1 + band), len2);
        if (head != phead || tail != ptail)
            for (int l = lo; l < hi && head > l; ++l) COL(l) = 0;
        if (i + 1 > mlen || head >= tail) live = false;

        int j;
        for (j = lo; j < hi; ++j) {
            const int q = bsw_sym(query[j * dir], wrap);
            const uint32_t cv = COL(j);
            const int diag = bsw_hd(cv), fc = bsw_fc(cv);
            int s = t == q ? sc_match : sc_mis;
            if (t == wrap || q == wrap) s = sc_amb;
            int M = diag ? diag + s : 0;
            if (BITS == 8 && M < 0) M = 0;
            const int h = max(max(M, er), fc);
            er = max(max(M - gapo_i, 0), er - p.e_ins);
            int fnew = max(max(M - gapo_d, 0), fc - p.e_del);
            const bool masked = head > j || j > tail;
            if (masked) { left = 0; fnew = 0; }
            const bool skip = BITS == 8 ? masked : (j + 1 > tail);
            if (!skip) {
                if (h >= rowmax) rowarg = j + 1;
                rowmax = max(rowmax, h);
            }
            COL(j) = bsw_pack(left, fnew);
            left = h;
            if (j + 1 == len2 && live && !(j + 1 > tail)) {
                if (!(gsc > h)) gi = i + 1;
                gsc = max(gsc, h);
            }
        }
        COL(j) = bsw_pack(left, 0);

        if (rowmax == 0) break;           // alone in the vector: the whole vector is dead
        const int prev = best;
        if (live) best = max(best, rowmax);
        if (best > prev) { by = rowarg; bx = i + 1; off = max(off, abs(rowarg - (i + 1))); }
        {   // z-drop without the gap-extension factor (vector rule)
            const int di = (i + 1) - bx, dj = rowarg - by;
            const int gap = di > dj ? di - dj : dj - di;
            if ((best - rowmax) - gap > p.zdrop) live = false;
        }
        // loop bounds of the next row: zero prefix / suffix of this row's stored state
        for (int l = lo; l < hi; ++l) { if (COL(l) == 0) nlo = l; else break; }
        {
            int l;
            for (l = hi; l >= lo; --l) if (COL(l) != 0) break;
            nhi = min(l + 2, ncol);
        }
        if (live) {
            int l;
            for (l = lo; l < hi && COL(l) == 0; ++l) {}
            if (l > lo) head = l;
            int idx = tail;
            for (l = hi; l >= lo && COL(l) == 0; --l) idx = l - 1;
            tail = min(idx + 2, len2);
        } else {
            // a dead lane's outputs are frozen; nothing below can change them
            break;
        }
    }
#undef COL
    BswOut o;
    o.score = best; o.qle = by; o.tle = bx; o.gtle = gi; o.gscore = gsc; o.max_off = off; o.cells = cells;
    return o;
}

// scalarBandedSWA: int32 state, two scratch words per column (H carry, E)
__device__ BswOut bsw_scalar_pair(const bwagpu_bsw_params &p, const uint8_t *__restrict__ query,
                                  int qlen, const uint8_t *__restrict__ target, int tlen, int w,
                                  int h0, uint32_t *__restrict__ col, int64_t stride, int dir = 1)
{
#define HC(j) col[(int64_t)(2 * (j)) * stride]
#define EC(j) col[(int64_t)(2 * (j) + 1) * stride]
    const int gapo_d = p.o_del + p.e_del, gapo_i = p.o_ins + p.e_ins;
    const int sc_match = p.w_match, sc_mis = -p.w_mismatch;
    for (int j = 0; j <= qlen + 1; ++j) { HC(j) = 0; EC(j) = 0; }
    HC(0) = (uint32_t)h0;
    if (qlen >= 1) HC(1) = (uint32_t)(h0 > gapo_i ? h0 - gapo_i : 0);
    for (int j = 2; j <= qlen && (int)HC(j - 1) > p.e_ins; ++j) HC(j) = HC(j - 1) - p.e_ins;
    {   // band clip: double division as the scalar code does
        const int best_sub = max(max(sc_match, sc_mis), 0);
        int g = (int)((double)(qlen * best_sub + p.end_bonus - p.o_ins) / p.e_ins + 1.);
        w = min(w, max(g, 1));
        g = (int)((double)(qlen * best_sub + p.end_bonus - p.o_del) / p.e_del + 1.);
        w = min(w, max(g, 1));
    }
    int best = h0, best_i = -1, best_j = -1, g_i = -1, g_sc = -1, off = 0;
    int lo = 0, hi = qlen, cells = 0;
    for (int i = 0; i < tlen; ++i) {
        const int t = target[i * dir];
        lo = max(lo, i - w);
        hi = min(min(hi, i + w + 1), qlen);
        cells += max(hi - lo, 0);
        int left = lo == 0 ? max(h0 - (p.o_del + p.e_del * (i + 1)), 0) : 0;
        int f = 0, rowmax = 0, rowarg = -1, j;
        for (j = lo; j < hi; ++j) {
            const int diag = (int)HC(j), e = (int)EC(j);
            HC(j) = (uint32_t)left;
            const int q = query[j * dir];
            const int s = (t > 3 || q > 3) ? -1 : (t == q ? sc_match : sc_mis);
            const int M = diag ? diag + s : 0;
            const int h = max(max(M, e), f);
            left = h;
            if (h >= rowmax) rowarg = j;
            rowmax = max(rowmax, h);
            EC(j) = (uint32_t)max(e - p.e_del, max(M - gapo_d, 0));
            f = max(f - p.e_ins, max(M - gapo_i, 0));
        }
        HC(hi) = (uint32_t)left; EC(hi) = 0;
        if (j == qlen) {
            if (!(g_sc > left)) g_i = i;
            g_sc = max(g_sc, left);
        }
        if (rowmax == 0) break;
        if (rowmax > best) {
            best = rowmax; best_i = i; best_j = rowarg;
            off = max(off, abs(rowarg - i));
        } else if (p.zdrop > 0) {
            const int di = i - best_i, dj = rowarg - best_j;
            if (di > dj) { if (best - rowmax - (di - dj) * p.e_del > p.zdrop) break; }
            else         { if (best - rowmax - (dj - di) * p.e_ins > p.zdrop) break; }
        }
        for (j = lo; j < hi && HC(j) == 0 && EC(j) == 0; ++j) {}
        lo = j;
        for (j = hi; j >= lo && HC(j) == 0 && EC(j) == 0; --j) {}
        hi = min(j + 2, qlen);
    }
#undef HC
#undef EC
    BswOut o;
    o.score = best; o.qle = best_j + 1; o.tle = best_i + 1; o.gtle = g_i + 1; o.gscore = g_sc;
    o.max_off = off; o.cells = cells;
    return o;
}
