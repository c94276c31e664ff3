// TEST INFRASTRUCTURE ONLY.  Harvests the mate-rescue jobs of a real paired-end run: oracle/Makefile compiles the
// reference's src/bwamem_pair.cpp with -Dksw_align2=tap_ksw_align2, so every ksw_align2 call mem_matesw makes
// (src/bwamem_pair.cpp:217, :331 — the jobs mem_sam_pe_batch_pre would emit as SeqPairs, :821-870) lands here: the
// unmodified ksw_align2 (src/ksw.cpp:347) computes the answer, and job + answer are appended to the file named by
// $BWAGPU_KSW_TAP at exit.  Record: int32 qlen, tlen, xtra, score, te, qe, score2, te2, tb, qb; then qlen query bytes and
// tlen target bytes.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <vector>
#include "ksw.h"

static std::mutex g_mu;
static std::vector<uint8_t> g_buf;
static bool g_hooked = false;

static void tap_flush()
{
    const char *path = getenv("BWAGPU_KSW_TAP");
    if (!path) return;
    FILE *f = fopen(path, "wb");
    if (!f) return;
    fwrite(g_buf.data(), 1, g_buf.size(), f);
    fclose(f);
}

extern "C" kswr_t tap_ksw_align2(int qlen, uint8_t *query, int tlen, uint8_t *target, int m, const int8_t *mat, int o_del, int e_del,
                      int o_ins, int e_ins, int xtra, kswq_t **qry)
{
    // copies first: ksw_align2 reverses its inputs in place for the start-point pass (and restores them)
    std::vector<uint8_t> q(query, query + qlen), t(target, target + tlen);
    const kswr_t r = ksw_align2(qlen, query, tlen, target, m, mat, o_del, e_del, o_ins, e_ins, xtra, qry);
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_hooked) { atexit(tap_flush); g_hooked = true; }
    const int32_t rec[10] = {qlen, tlen, xtra, r.score, r.te, r.qe, r.score2, r.te2, r.tb, r.qb};
    const uint8_t *p = (const uint8_t *)rec;
    g_buf.insert(g_buf.end(), p, p + sizeof rec);
    g_buf.insert(g_buf.end(), q.begin(), q.end());
    g_buf.insert(g_buf.end(), t.begin(), t.end());
    return r;
}
