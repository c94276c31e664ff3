// The part of the ERT k-mer table that follows from occurrence counts alone, computed from the FM-index:
// for the k-mer with index idx (base i at bits 2i..2i+1) the reference's builder extends the k-mer base by base to
// the right and sets LEP bit i-1 whenever the number of occurrences changes at base i, stopping at the first base
// that no longer occurs (reference src/ertindex.cpp buildIndex, the loop over `i < kmerSize`).  An absent k-mer's
// entry is (LEP << 2) | INVALID; its pointer bits are never read by the seeding code.
// Per-thread code shared with tests/hostsim (test scaffolding, not a product path).
#pragma once
#include "fmi_core.cuh"

struct ErtKmerWalk {
    Biv cur;            // bi-interval of the prefix walked so far
    uint32_t lep;
    int64_t hits;       // occurrences of the prefix; 0 = dead
    BWAGPU_HD void start(const FmView &f, int a)
    {
        cur.k = fm_count(f, a); cur.l = fm_count(f, 3 - a); cur.s = fm_count(f, a + 1) - cur.k;
        lep = 0; hits = cur.s;
    }
    // base number i (1..14) of the k-mer; no-op once dead... the caller must not step a dead walk
    BWAGPU_HD void step(const FmView &f, int i, int a)
    {
        Biv t; t.k = cur.l; t.l = cur.k; t.s = cur.s;
        const Biv r = fm_bext(f, t, 3 - a);                 // forward extension = backward on the other strand
        if (r.s != hits) lep |= 1u << (i - 1);
        if (r.s >= 1) { cur.k = r.l; cur.l = r.k; cur.s = r.s; }
        hits = r.s;
    }
};

// whole k-mer at once (tests, and the reference point for the tiled kernel)
BWAGPU_HD uint32_t ert_kmer_lep(const FmView &f, uint32_t idx, int64_t *hits)
{
    ErtKmerWalk w;
    w.start(f, (int)(idx & 3));
    for (int i = 1; i < 15 && w.hits > 0; ++i) w.step(f, i, (int)((idx >> (2 * i)) & 3));
    *hits = w.hits;
    return w.lep & 0x3FFF;
}
