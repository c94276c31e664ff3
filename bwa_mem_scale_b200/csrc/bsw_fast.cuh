// Banded SW extension, one pair per thread, column state in shared memory ("fast" path for
// short-read shapes: len2 small enough that every thread's H/F column fits in shared memory).
//
// Same per-pair semantics as bsw_lane_pair (bsw_thread.cuh), i.e. one SIMD lane of the reference's
// smithWaterman128_16 / _8 with its batch wrapper (reference src/bandedSWA.cpp:3448-4121,
// 4172-4837).  What differs is only how a row is executed:
//   * the row is cut into [lo, head) | [head, tail) | [tail, hi): the middle part is the lane's
//     live range, where neither the head/tail masks nor the "past the tail" rule apply, and runs a
//     short straight-line cell (DPX three-way max / add-max-relu);
//   * the substitution score is one PRMT: the target base selects a per-row byte table
//     (match/mismatch/ambiguous), the query base selects the byte, sign-extended by the permute;
//   * the running (row maximum, last column attaining it) is one packed max: h<<16 | (j+1).
// Column word j of thread t lives at col[j * STRIDE] (bank = t: conflict-free).
#pragma once
#include <stdint.h>
#include "bsw_thread.cuh"

// PTX prmt in its default mode: a selector nibble with bit 3 set replicates the sign bit of the
// selected byte (the __byte_perm intrinsic ignores that bit, hence the inline form)
__device__ __forceinline__ uint32_t bsw_prmt(uint32_t a, uint32_t b, uint32_t sel)
{
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}

// Row arg-max bookkeeping: the low half of the packed (row maximum, column) key is a column code.  An exact code names
// the column (1-based) itself; a word code says "in the 8 columns that end here".  Codes order like the columns they
// stand for, so the packed maximum still ends up on the LAST column attaining the row maximum (or on its word).
#ifndef BSW_V2
#define BSW_V2 1
#endif
#define BSW_KEY_EXACT(c1) (((c1) << 1) | 1)
#define BSW_KEY_WORD(c1) ((c1) << 1)

// BSW_FMA_PIPE: 1 = upper half-word, gap-extension decrements and (left, F) packing as integer multiply-adds; 2 = the same
// with a run-time 1 as multiplier so that the compiler keeps both decrements on the FMA pipe (ALU : FMA instructions per
// 8 cells 73 : 16 -> 61 : 35 -> 52 : 43).  MEASURED (scripts/gpu_r2_aa.sh, ms per 5 M reads at config-3 scale): 265.2 off,
// 272.8 with 1, 263.8 with 2 — no gain, so it stays OFF: the ALU-pipe figure of the ncu capture is one tier's, the sum
// over the tiers is bound by issue and by the rows' serial bookkeeping.
#ifndef BSW_FMA_PIPE
#define BSW_FMA_PIPE 0
#endif
// a * b + c as an integer multiply-add that stays one (the compiler would turn a * 1 + c back into an add)
__device__ __forceinline__ int bsw_mad(int a, int b, int c)
{
    int d;
    asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

#ifndef BSW_UNROLL
#define BSW_UNROLL 4
#endif
constexpr int kBswUnroll = BSW_UNROLL;

template <int BITS, int STRIDE>
__device__ BswOut bsw_lane_fast(const bwagpu_bsw_params &p, const uint8_t *__restrict__ query,
                                int len2, const uint8_t *__restrict__ target, int len1, int w,
                                int h0, uint32_t *__restrict__ col, uint32_t *__restrict__ qsm, int dir = 1)
{
#define COL(j) col[(j) * STRIDE]
    // (asking for every 128-byte line of query and target up front with prefetch.global.L1 was measured: the staging loop
    // below shows 8 % of the stall samples, profiles/r02_ncu_bsw_fast_v3_lines.txt, but the kernel time did not move —
    // 264.6 vs 260.3 ms per 5 M reads, scripts/gpu_r2_z.sh — other warps already cover it)
    // stage the query once: 4-bit codes, 8 per word, word k of this thread at qsm[k * STRIDE]
    // (a byte load per cell from global memory was the kernel's bottleneck: L1 request rate)
    for (int k = 0; k * 8 < len2; ++k) {
        uint32_t v = 0;
        for (int b = 0; b < 8 && k * 8 + b < len2; ++b) v |= (uint32_t)(query[(k * 8 + b) * dir] & 7u) << (4 * b);
        qsm[k * STRIDE] = v;
    }
    const int wrap = BITS == 8 ? 0xFF : 0xFFFF;
    [[maybe_unused]] const int one_rt = (int)(blockDim.x / STRIDE);     // 1, but not to the compiler (BSW_FMA_PIPE 2)
    const int gapo_d = p.o_del + p.e_del, gapo_i = p.o_ins + p.e_ins;
    const int e_del = p.e_del, e_ins = p.e_ins;
    const int sc_match = p.w_match, sc_mis = -p.w_mismatch, sc_amb = -1;
    const int nrow = len1, ncol = len2;
    // byte tables for the PRMT score look-up: bytes 0..3 = query A,C,G,T; byte 4 = query N
    const uint32_t mis4 = (uint32_t)(sc_mis & 0xFF) * 0x01010101u;
    const uint32_t amb4 = (uint32_t)(sc_amb & 0xFF) * 0x01010101u;

    {   // first row: seed score decaying along an insertion run; columns ncol..ncol+2 start at 0
        int v = h0;
        COL(0) = bsw_pack(v, 0);
        v = h0 > gapo_i ? h0 - gapo_i : 0;
        COL(1) = bsw_pack(v, 0);
        for (int k = 2; k < ncol; ++k) { v = max(v - e_ins, 0); COL(k) = bsw_pack(v, 0); }
        for (int k = max(ncol, 2); k < ncol + 3; ++k) COL(k) = 0;
    }
    int band;
    {
        int best_sub = max(max(sc_match, sc_mis), max(sc_amb, 0));
        int qmax = (len2 * best_sub) & wrap;
        int t = (qmax + p.end_bonus - p.o_ins) & wrap;
        band = min(w, max(t / e_ins + 1, 1));
        t = (qmax + p.end_bonus - p.o_del) & wrap;
        band = min(band, max(t / e_del + 1, 1)) & wrap;
    }
    int head = 0, tail = len2;
    const int mlen = min((len2 + band) & wrap, len1);
    int best = h0, bx = 0, by = 0, gsc = -1, gi = 0, off = 0;
    bool live = true;
    int nlo = 0, nhi = ncol;
    int cells = 0;

    int tc_next = nrow > 0 ? target[0] : 0;       // one row ahead: keeps the byte load off the row's critical path
    for (int i = 0; i < nrow; ++i) {
        const int tc = tc_next;
        if (i + 1 < nrow) tc_next = target[(i + 1) * dir];
        const int t = bsw_sym((uint8_t)tc, wrap);
        uint32_t tlo, thi = amb4;
        if (tc > 3) tlo = amb4;
        else tlo = (mis4 & ~(0xFFu << (8 * tc))) | ((uint32_t)(sc_match & 0xFF) << (8 * tc));

        const int lo = max(nlo, i - w), hi = min(min(nhi, i + w + 1), ncol);
        cells += max(hi - lo, 0);
        int left = lo == 0 ? max(h0 - p.o_del - (i + 1) * e_del, 0) : 0;
        int er = 0;
        int key = 0;                              // rowmax << 16 | rowarg

        const int phead = head, ptail = tail;
        head = max(head, i - band);
        tail = min(min(tail, i + 1 + band), len2);
        if (head != phead || tail != ptail)
            for (int l = lo; l < hi && head > l; ++l) COL(l) = 0;
        if (i + 1 > mlen || head >= tail) live = false;

        // general cell: exactly the body of bsw_lane_pair
#define GENERAL_CELL(j)                                                               \
        {                                                                             \
            const int q = bsw_sym((uint8_t)((qsm[((j) >> 3) * STRIDE] >> (4 * ((j) & 7))) & 7u), wrap); \
            const uint32_t cv = COL(j);                                               \
            const int diag = bsw_hd(cv), fc = bsw_fc(cv);                             \
            int s = t == q ? sc_match : sc_mis;                                       \
            if (t == wrap || q == wrap) s = sc_amb;                                   \
            int M = diag ? diag + s : 0;                                              \
            if (BITS == 8 && M < 0) M = 0;                                            \
            const int h = max(max(M, er), fc);                                        \
            er = max(max(M - gapo_i, 0), er - e_ins);                                 \
            int fnew = max(max(M - gapo_d, 0), fc - e_del);                           \
            const bool masked = head > (j) || (j) > tail;                             \
            if (masked) { left = 0; fnew = 0; }                                       \
            const bool skip = BITS == 8 ? masked : ((j) + 1 > tail);                  \
            if (!skip) key = max(key, (h << 16) | BSW_KEY_EXACT((j) + 1));            \
            COL(j) = bsw_pack(left, fnew);                                            \
            left = h;                                                                 \
            if ((j) + 1 == len2 && live && !((j) + 1 > tail)) {                       \
                if (!(gsc > h)) gi = i + 1;                                           \
                gsc = max(gsc, h);                                                    \
            }                                                                         \
        }

        int j = lo;
        const int ja = min(hi, max(lo, head));    // [lo, ja): before the live range
        for (; j < ja; ++j) GENERAL_CELL(j)
        const int jb = min(hi, tail);             // [ja, jb): live range, j >= head and j + 1 <= tail
        if (j < jb) {
            uint32_t *cp = &COL(j);
            // one live cell: NIB = query code of column J, KEY receives (h << 16) + the exact-column code of J + 1
#define LIVE_CELL(NIB, J, KEY)                                                        \
            {                                                                         \
                const uint32_t cv = *cp;                                              \
                const int diag = (int)(cv & 0xFFFFu), fc = (int)(cv >> 16);           \
                /* NIB: a word whose low nibble is the query code (bit 3 of every nibble clear): byte 0 of the permute */ \
                /* is this cell's score, added to the diagonal as a signed byte */   \
                int M = BSW_V2 ? __dp4a((int)bsw_prmt(tlo, thi, (NIB)), 1, diag)      \
                               : diag + (int)bsw_prmt(tlo, thi, ((NIB) & 7u) * 0x1111u + 0x8880u); \
                M = diag ? M : 0;                                                     \
                if (BITS == 8) M = max(M, 0);                                         \
                const int h = __vimax3_s32(M, er, fc);                                \
                er = __viaddmax_s32_relu(M, -gapo_i, er - e_ins);                     \
                const int fnew = __viaddmax_s32_relu(M, -gapo_d, fc - e_del);         \
                KEY = (h << 16) + BSW_KEY_EXACT((J) + 1);                             \
                *cp = __byte_perm((uint32_t)left, (uint32_t)fnew, 0x5410);            \
                left = h;                                                             \
                cp += STRIDE;                                                         \
            }
#if BSW_V2
            // The staged query holds 8 codes per word.  Four cells share ONE permute for their scores: the codes are 4-bit
            // with bit 3 clear, so 16 bits of the word are themselves a selector that picks the four score bytes, and a
            // signed byte dot product with a one-hot word adds byte c to the diagonal (M = diag + s in one operation).
            // The live range runs as: single cells up to a multiple of 4, a half word up to a multiple of 8, whole
            // words, a trailing half word, single cells.  Inside half and whole words the row's arg-max is kept per GROUP —
            // key = (max h of the group) << 16 | code of "somewhere up to the group's last column" — and the exact column
            // is looked up after the row, only when the row's outcome depends on it.
            // The cell is bound by the ALU pipe (logic, shifts, selects, permutes and the DPX min/max operations all issue
            // there at 16 lanes per clock and scheduler: 81 % busy against 21 % for the FMA pipe, profiles/
            // r02_ncu_bsw_fast_v3_lines.txt), so whatever can be phrased as an integer multiply-add goes to the FMA pipe:
            // the upper half-word as a high multiply, the two gap-extension decrements and the packing of (left, F).
#if BSW_FMA_PIPE
#define BSW_HI16(x) ((int)__umulhi((x), 0x10000u))
#define BSW_SUB(x, c) bsw_mad((x), BSW_FMA_PIPE > 1 ? one_rt : 1, -(c))
#define BSW_PACK(lo, hi) ((uint32_t)bsw_mad((hi), 0x10000, (lo)))
#else
#define BSW_HI16(x) ((int)((x) >> 16))
#define BSW_SUB(x, c) ((x) - (c))
#define BSW_PACK(lo, hi) __byte_perm((uint32_t)(lo), (uint32_t)(hi), 0x5410)
#endif
#define LIVE_CELL4(S, C, H)                                                           \
            {                                                                         \
                const uint32_t cv = *cp;                                              \
                const int diag = (int)(cv & 0xFFFFu), fc = BSW_HI16(cv);              \
                int M = __dp4a((int)(S), 1 << (8 * (C)), diag);                       \
                M = diag ? M : 0;            /* (a predicated dot product comes back from ptxas as IDP + SEL) */ \
                if (BITS == 8) M = max(M, 0);                                         \
                H = __vimax3_s32(M, er, fc);                                          \
                er = __viaddmax_s32_relu(M, -gapo_i, BSW_SUB(er, e_ins));             \
                const int fnew = __viaddmax_s32_relu(M, -gapo_d, BSW_SUB(fc, e_del)); \
                *cp = BSW_PACK(left, fnew);                                           \
                left = H;                                                             \
                cp += STRIDE;                                                         \
            }
#define LIVE_HALF(WQ16)                                                               \
            {                                                                         \
                const uint32_t s03 = bsw_prmt(tlo, thi, (WQ16));                      \
                int h0_, h1_, h2_, h3_;                                               \
                LIVE_CELL4(s03, 0, h0_) LIVE_CELL4(s03, 1, h1_) LIVE_CELL4(s03, 2, h2_) LIVE_CELL4(s03, 3, h3_) \
                const int bm = max(__vimax3_s32(h0_, h1_, h2_), h3_);                 \
                key = max(key, (bm << 16) + BSW_KEY_WORD(j + 4));                     \
                j += 4;                                                               \
            }
            const int jpro = min(jb, (j + 3) & ~3);
            if (j < jpro) {
                uint32_t wq = qsm[(j >> 3) * STRIDE] >> (4 * (j & 7));
                for (; j < jpro; ++j) { int k1; LIVE_CELL(wq, j, k1) key = max(key, k1); wq >>= 4; }
            }
            if ((j & 4) && j + 4 <= jb) LIVE_HALF(qsm[(j >> 3) * STRIDE] >> 16)
            const uint32_t *qp = qsm + (j >> 3) * STRIDE;
            for (; j + 8 <= jb; j += 8, qp += STRIDE) {
                const uint32_t wq = *qp;
                const uint32_t s03 = bsw_prmt(tlo, thi, wq), s47 = bsw_prmt(tlo, thi, wq >> 16);
                int h0_, h1_, h2_, h3_, h4_, h5_, h6_, h7_;
                LIVE_CELL4(s03, 0, h0_) LIVE_CELL4(s03, 1, h1_) LIVE_CELL4(s03, 2, h2_) LIVE_CELL4(s03, 3, h3_)
                LIVE_CELL4(s47, 0, h4_) LIVE_CELL4(s47, 1, h5_) LIVE_CELL4(s47, 2, h6_) LIVE_CELL4(s47, 3, h7_)
                int bm = __vimax3_s32(h0_, h1_, h2_);
                bm = __vimax3_s32(bm, h3_, h4_);
                bm = __vimax3_s32(bm, h5_, h6_);
                bm = max(bm, h7_);
                key = max(key, (bm << 16) + BSW_KEY_WORD(j + 8));
            }
            if (j < jb) {
                uint32_t wq = qsm[(j >> 3) * STRIDE] >> (4 * (j & 7));
                if (j + 4 <= jb) { LIVE_HALF(wq) wq >>= 16; }
                for (; j < jb; ++j) { int k1; LIVE_CELL(wq, j, k1) key = max(key, k1); wq >>= 4; }
            }
#undef LIVE_HALF
#undef LIVE_CELL4
#undef BSW_HI16
#undef BSW_SUB
#undef BSW_PACK
#else
            // columns up to the next multiple of 8 (the query is staged 8 codes per word)
            const int jpro = min(jb, (j + 7) & ~7);
            if (j < jpro) {
                uint32_t wq = qsm[(j >> 3) * STRIDE] >> (4 * (j & 7));
                for (; j < jpro; ++j) { int k1; LIVE_CELL(wq, j, k1) key = max(key, k1); wq >>= 4; }
            }
            // whole words: 8 cells straight-line, shifts known at compile time
            const uint32_t *qp = qsm + (j >> 3) * STRIDE;
            for (; j + 8 <= jb; j += 8, qp += STRIDE) {
                const uint32_t wq = *qp;
                int k0, k1, k2, k3, k4, k5, k6, k7;
                LIVE_CELL(wq & 7u, j, k0)
                LIVE_CELL((wq >> 4) & 7u, j + 1, k1)
                LIVE_CELL((wq >> 8) & 7u, j + 2, k2)
                LIVE_CELL((wq >> 12) & 7u, j + 3, k3)
                LIVE_CELL((wq >> 16) & 7u, j + 4, k4)
                LIVE_CELL((wq >> 20) & 7u, j + 5, k5)
                LIVE_CELL((wq >> 24) & 7u, j + 6, k6)
                LIVE_CELL((wq >> 28) & 7u, j + 7, k7)
                key = __vimax3_s32(key, k0, k1);
                key = __vimax3_s32(key, k2, k3);
                key = __vimax3_s32(key, k4, k5);
                key = __vimax3_s32(key, k6, k7);
            }
            if (j < jb) {
                uint32_t wq = *qp;
                for (; j < jb; ++j) { int k1; LIVE_CELL(wq, j, k1) key = max(key, k1); wq >>= 4; }
            }
#endif
#undef LIVE_CELL
            if (jb == len2 && live) {             // the last query column was in the live range
                if (!(gsc > left)) gi = i + 1;
                gsc = max(gsc, left);
            }
        }
        for (; j < hi; ++j) GENERAL_CELL(j)
#undef GENERAL_CELL
        COL(j) = bsw_pack(left, 0);

        const int rowmax = key >> 16;
        if (rowmax == 0) break;
        int rowarg = (key & 0xFFFF) >> 1;
        if (BSW_V2 && !(key & 1) && ((live && rowmax > best) || best - rowmax > p.zdrop)) {
            // the maximum sits in a whole-word group ending at column `rowarg` and this row's outcome depends on where: the
            // last column that attains it (H of column c is the low half of word c + 1 once the row is written)
            int c = rowarg;
            while ((int)(COL(c) & 0xFFFFu) != rowmax) --c;
            rowarg = c;
        }
        const int prev = best;
        if (live) best = max(best, rowmax);
        if (best > prev) { by = rowarg; bx = i + 1; off = max(off, abs(rowarg - (i + 1))); }
        {
            const int di = (i + 1) - bx, dj = rowarg - by;
            const int gap = di > dj ? di - dj : dj - di;
            if ((best - rowmax) - gap > p.zdrop) live = false;
        }
        // the row's leading and trailing runs of zero words bound the next row (nlo, nhi) and, while the lane is live,
        // its head and tail: one scan from each end serves both
        int lead = lo, trail = hi;
        while (lead < hi && COL(lead) == 0) ++lead;          // first non-zero word of [lo, hi), or hi
        if (left == 0) {                                     // word hi was just written: (left, 0)
            --trail;
            while (trail >= lo && COL(trail) == 0) --trail;  // last non-zero word of [lo, hi], or lo - 1
        }
        if (lead > lo) nlo = lead - 1;
        nhi = min(trail + 2, ncol);
        if (!live) break;
        if (lead > lo) head = lead;
        tail = min((trail < hi ? trail : tail) + 2, len2);
    }
#undef COL
    BswOut o;
    o.score = best; o.qle = by; o.tle = bx; o.gtle = gi; o.gscore = gsc; o.max_off = off; o.cells = cells;
    return o;
}
