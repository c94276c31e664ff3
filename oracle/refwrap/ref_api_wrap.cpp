// TEST INFRASTRUCTURE ONLY — never linked into the product library.
//
// extern "C" taps into the UNMODIFIED reference (all of /root/reference/src except main.cpp and
// runsimd.cpp, compiled where they lie by oracle/Makefile) so tests/ and bench.py's cpu_baseline
// leg can run the reference's own seeding code on caller-provided reads:
//   ref_fmi_open / ref_fmi_close  : FMI_search(prefix) + load_index()   (src/FMI_search.cpp:1251)
//   ref_collect_smem              : mem_collect_smem                    (src/bwamem.cpp:648-786)
//   ref_sa_lookup                 : FMI_search::get_sa_entries_prefetch (src/FMI_search.cpp:2261)
// Built with the scale=0 flag set (plain bwa-mem2 + ERT code, no shm / EMF / FMA).
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include "bwamem.h"
#include "FMI_search.h"

// globals main.cpp owns upstream (src/main.cpp:48)
uint64_t proc_freq = 1, tprof[LIM_R][LIM_C], prof[LIM_R];

SMEM *mem_collect_smem(FMI_search *fmi, const mem_opt_t *opt, const bseq1_t *seq_, int nseq,
                       SMEM *matchArray, int32_t *min_intv_ar, int16_t *query_pos_ar,
                       uint8_t *enc_qdb, int32_t *rid, int64_t &tot_smem);

extern "C" {

int ref_sizeof_smem(void) { return (int)sizeof(SMEM); }

void *ref_fmi_open(const char *prefix)
{
    FMI_search *f = new FMI_search(prefix);
    f->load_index();
    return f;
}

void ref_fmi_close(void *h) { delete (FMI_search *)h; }

// reads: concatenated 2-bit codes (0-3, >=4 = N); off[i] = start of read i; len[i] its length.
// opt: min_seed_len, split_width, max_mem_intv; split_factor as float.  Returns tot_smem, writes
// the sorted SMEM array exactly as mem_collect_smem leaves it in matchArray.
int64_t ref_collect_smem(void *h, int min_seed_len, float split_factor, int split_width,
                         int64_t max_mem_intv, const uint8_t *reads, const int64_t *off,
                         const int32_t *len, int nseq, SMEM *out, int64_t out_cap)
{
    FMI_search *fmi = (FMI_search *)h;
    mem_opt_t *opt = mem_opt_init();
    opt->min_seed_len = min_seed_len;
    opt->split_factor = split_factor;
    opt->split_width = split_width;
    opt->max_mem_intv = max_mem_intv;
    bseq1_t *seqs = (bseq1_t *)calloc(nseq, sizeof(bseq1_t));
    int64_t tot_len = 0;
    int maxlen = 1;
    for (int i = 0; i < nseq; ++i) {
        seqs[i].l_seq = len[i];
        seqs[i].id = i;
        seqs[i].seq = (char *)(reads + off[i]);
        tot_len += len[i];
        if (len[i] > maxlen) maxlen = len[i];
    }
    // arenas sized like memoryAlloc does per thread (src/fastmap.cpp:280-300) but for this batch
    int64_t cap = (int64_t)20 * nseq * maxlen + 1024;
    SMEM *matchArray = (SMEM *)_mm_malloc(cap * sizeof(SMEM), 64);
    int32_t *min_intv_ar = (int32_t *)_mm_malloc(cap * sizeof(int32_t), 64);
    int16_t *query_pos_ar = (int16_t *)_mm_malloc(cap * sizeof(int16_t), 64);
    int32_t *rid = (int32_t *)_mm_malloc(cap * sizeof(int32_t), 64);
    uint8_t *enc_qdb = (uint8_t *)_mm_malloc(tot_len + 1024, 64);
    int64_t tot = 0;
    mem_collect_smem(fmi, opt, seqs, nseq, matchArray, min_intv_ar, query_pos_ar, enc_qdb, rid, tot);
    int64_t ncopy = tot < out_cap ? tot : out_cap;
    memcpy(out, matchArray, ncopy * sizeof(SMEM));
    _mm_free(matchArray); _mm_free(min_intv_ar); _mm_free(query_pos_ar); _mm_free(rid);
    _mm_free(enc_qdb);
    free(seqs); free(opt);
    return tot;
}

// coords[] receives the SA coordinates in SMEM order then row order (k, k+step, ...), as
// mem_chain_seeds consumes them (src/bwamem.cpp:866-893).  Returns the number written.
int64_t ref_sa_lookup(void *h, SMEM *smems, int64_t n, int32_t max_occ, int64_t *coords)
{
    FMI_search *fmi = (FMI_search *)h;
    int64_t cnt = 0;
    // one call per read, as mem_chain_seeds does (its scratch is sized by an int32 sum of s)
    for (int64_t b = 0; b < n;) {
        int64_t e = b + 1;
        while (e < n && smems[e].rid == smems[b].rid) ++e;
        int64_t c = 0, id = 0;
        fmi->get_sa_entries_prefetch(smems + b, coords + cnt, &c, e - b, max_occ, 0, id);
        cnt += c;
        b = e;
    }
    return cnt;
}

}  // extern "C"
