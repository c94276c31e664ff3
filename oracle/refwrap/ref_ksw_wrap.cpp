// TEST INFRASTRUCTURE ONLY — never linked into the product library.
// extern "C" shim around the UNMODIFIED reference ksw_align2 (reference src/ksw.cpp:347-381),
// compiled together with src/ksw.cpp where it lies by oracle/Makefile -> oracle/_ref/libref_ksw.so.
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include "ksw.h"

extern "C" {
struct ref_seqpair { int32_t idr, idq, id, len1, len2, h0, seqid, regid, score, tle, gtle, qle, gscore, max_off; };

// job i: query = qer_buf + idq (len2), target = ref_buf + idr (len1), xtra = h0  -> aln[i]
// (the sequential body of mem_sam_pe_batch, reference src/bwamem_pair.cpp:891-902)
void ref_ksw_batch(const ref_seqpair *pairs, int32_t n, const uint8_t *ref_buf, const uint8_t *qer_buf,
                   const int8_t *mat, int o_del, int e_del, int o_ins, int e_ins, kswr_t *aln)
{
    for (int32_t i = 0; i < n; ++i) {
        // ksw_align2 reverses its inputs in place and restores them: give it private copies
        uint8_t *q = (uint8_t *)malloc(pairs[i].len2 + 64), *t = (uint8_t *)malloc(pairs[i].len1 + 64);
        memcpy(q, qer_buf + pairs[i].idq, pairs[i].len2);
        memcpy(t, ref_buf + pairs[i].idr, pairs[i].len1);
        aln[i] = ksw_align2(pairs[i].len2, q, pairs[i].len1, t, 5, mat, o_del, e_del, o_ins, e_ins, pairs[i].h0, 0);
        free(q); free(t);
    }
}
}
