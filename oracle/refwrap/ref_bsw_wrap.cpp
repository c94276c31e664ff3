// TEST INFRASTRUCTURE ONLY — never linked into the product library.
//
// extern "C" shim around the UNMODIFIED reference class BandedPairWiseSW
// (reference src/bandedSWA.h:113-336, src/bandedSWA.cpp), compiled from the sources where they
// lie under /root/reference by oracle/Makefile into oracle/_ref/libref_bsw_<isa>.so.
// It lets tests/ and bench.py's cpu_baseline leg call the reference's own getScores8 /
// getScores16 / scalarBandedSWAWrapper on caller-provided SeqPair arrays.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include "bandedSWA.h"

uint64_t prof[10][112];   // bandedSWA.cpp:44 declares it extern (lives in main.cpp upstream)

extern "C" {

struct ref_bsw_params {
    int32_t o_del, e_del, o_ins, e_ins, zdrop, end_bonus, w_match, w_mismatch;
};

static void fill_mat(int8_t mat[25], int a, int b) {
    // same matrix bwa_fill_scmat builds (reference src/bwa.cpp:368-378)
    int k = 0;
    for (int i = 0; i < 4; ++i) {
        for (int j = 0; j < 4; ++j) mat[k++] = i == j ? a : -b;
        mat[k++] = -1;
    }
    for (int j = 0; j < 5; ++j) mat[k++] = -1;
}

// kind: 0 = scalarBandedSWAWrapper, 8 = getScores8, 16 = getScores16.
// pairs must have room for SIMD padding entries after n (the reference writes them,
// bandedSWA.cpp:3491-3496) plus PFD read-ahead; we copy into a padded scratch array.
int ref_bsw_run(const ref_bsw_params *p, int kind, SeqPair *pairs, int32_t n,
                const uint8_t *ref_buf, int64_t ref_len, const uint8_t *qer_buf, int64_t qer_len,
                int32_t w)
{
    int8_t mat[25];
    fill_mat(mat, p->w_match, p->w_mismatch);
    BandedPairWiseSW bsw(p->o_del, p->e_del, p->o_ins, p->e_ins, p->zdrop, p->end_bonus, mat,
                         (int8_t)p->w_match, (int8_t)p->w_mismatch, 1);
    SeqPair *tmp = (SeqPair *)_mm_malloc(sizeof(SeqPair) * (size_t)(n + 256), 64);
    memset(tmp, 0, sizeof(SeqPair) * (size_t)(n + 256));
    memcpy(tmp, pairs, sizeof(SeqPair) * (size_t)n);
    // slack after the sequence buffers, like the reference's arenas (fastmap.cpp:172-174)
    uint8_t *rb = (uint8_t *)_mm_malloc((size_t)ref_len + 512, 64);
    uint8_t *qb = (uint8_t *)_mm_malloc((size_t)qer_len + 512, 64);
    memcpy(rb, ref_buf, (size_t)ref_len); memset(rb + ref_len, 0, 512);
    memcpy(qb, qer_buf, (size_t)qer_len); memset(qb + qer_len, 0, 512);
    if (kind == 16)      bsw.getScores16(tmp, rb, qb, n, 1, w);
    else if (kind == 8)  bsw.getScores8(tmp, rb, qb, n, 1, w);
    else                 bsw.scalarBandedSWAWrapper(tmp, rb, qb, n, 1, w);
    memcpy(pairs, tmp, sizeof(SeqPair) * (size_t)n);
    _mm_free(tmp); _mm_free(rb); _mm_free(qb);
    return 0;
}

int ref_bsw_simd_width16(void) { return SIMD_WIDTH16; }
int ref_bsw_simd_width8(void)  { return SIMD_WIDTH8; }
int ref_bsw_sizeof_seqpair(void) { return (int)sizeof(SeqPair); }

}  // extern "C"
