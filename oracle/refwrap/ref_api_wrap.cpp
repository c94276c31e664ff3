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
#include <string>
#include <vector>
#include <fstream>
#include <iostream>
#include <map>
#include <algorithm>
#include <immintrin.h>
// ref_mem_open_arrays below hands an index that is already in memory to FMI_search, whose array members are private
#define private public
#include "FMI_search.h"
#undef private
#include "bwamem.h"
#include "bwa.h"
#include "ertseeding.h"
#include <cstdio>

// globals main.cpp owns upstream (src/main.cpp:48)
uint64_t proc_freq = 1, tprof[LIM_R][LIM_C], prof[LIM_R];

// the definition (src/bwamem.cpp:1209-1217) takes seedBuf/seedBufSize; the header's declaration is stale
int mem_kernel1_core(FMI_search *fmi, const mem_opt_t *opt, bseq1_t *seq_, int nseq, mem_chain_v *chain_ar,
                     mem_seed_t *seedBuf, int64_t seedBufSize, mem_cache *mmc, int tid);

SMEM *mem_collect_smem(FMI_search *fmi, const mem_opt_t *opt, const bseq1_t *seq_, int nseq,
                       SMEM *matchArray, int32_t *min_intv_ar, int16_t *query_pos_ar,
                       uint8_t *enc_qdb, int32_t *rid, int64_t &tot_smem);

#ifdef PERFECT_MATCH
int find_perfect_match_entry(perfect_table_t *pt, bseq1_t *seq, int len);   // src/perfect_map.cpp:638
#endif

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
static int64_t collect_smem_impl(void *h, int min_seed_len, float split_factor, int split_width,
                                 int64_t max_mem_intv, const uint8_t *reads, const int64_t *off,
                                 const int32_t *len, int nseq, SMEM *out, int64_t out_cap, const uint8_t *skip)
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
#ifdef PERFECT_MATCH
        if (skip && skip[i]) { seqs[i].perfect.flags = FLAG_VALID; seqs[i].perfect.location = 0; }   // an EMF answer
#endif
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

int64_t ref_collect_smem(void *h, int min_seed_len, float split_factor, int split_width,
                         int64_t max_mem_intv, const uint8_t *reads, const int64_t *off,
                         const int32_t *len, int nseq, SMEM *out, int64_t out_cap)
{
    return collect_smem_impl(h, min_seed_len, split_factor, split_width, max_mem_intv, reads, off, len, nseq, out, out_cap, nullptr);
}

// same with reads marked as answered by the EMF (seq_[i].perfect.exist != 0, src/bwamem.cpp:672-689);
// only meaningful in the PERFECT_MATCH flavours
int64_t ref_collect_smem_skip(void *h, int min_seed_len, float split_factor, int split_width,
                              int64_t max_mem_intv, const uint8_t *reads, const int64_t *off,
                              const int32_t *len, int nseq, SMEM *out, int64_t out_cap, const uint8_t *skip)
{
    return collect_smem_impl(h, min_seed_len, split_factor, split_width, max_mem_intv, reads, off, len, nseq, out, out_cap, skip);
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

// ---- chaining + extension taps --------------------------------------------------------------
// Runs the reference's own mem_kernel1_core (seeding + mem_chain_seeds + mem_chain_flt, src/bwamem.cpp:
// 1209-1378) and mem_chain2aln_across_reads_V2 (:2810-3745) on caller-provided reads, 512 at a time
// like worker_bwt / worker_aln (src/macro.h:63), and returns flattened copies of
//   chains : per chain  {seqid, rid, n, w, kept, first, is_alt} + frac_rep + pos, seeds {rbeg, qbeg, len, score}
//   regs   : per read the mem_alnreg_t records V2 leaves in av_v (including purged ones, qb = qe = -1)
struct ref_mem_ctx {
    FMI_search *fmi;
    uint8_t *ref_string;
    mem_cache mmc;
    mem_seed_t *seedBuf;
    int use_ert;                 // seeding + chaining through mem_kernel1_core_ert (src/bwamem.cpp:1052-1207)
};

void *ref_mem_open(const char *prefix)
{
    ref_mem_ctx *c = (ref_mem_ctx *)calloc(1, sizeof(ref_mem_ctx));
    c->fmi = new FMI_search(prefix);
    c->fmi->load_index();
    char fn[4096];
    snprintf(fn, sizeof fn, "%s.0123", prefix);
    FILE *fp = fopen(fn, "rb");
    fseek(fp, 0, SEEK_END); long rlen = ftell(fp); fseek(fp, 0, SEEK_SET);
    c->ref_string = (uint8_t *)_mm_malloc(rlen + 256, 64);
    if (fread(c->ref_string, 1, rlen, fp) != (size_t)rlen) return nullptr;
    fclose(fp);
    // arenas of one worker thread, sized as memoryAlloc does (src/fastmap.cpp:222-300)
    const int64_t wsize = BATCH_SIZE * SEEDS_PER_READ;
    const int readLen = 1024;
    mem_cache &m = c->mmc;
    m.seqBufLeftRef[0] = (uint8_t *)_mm_malloc(wsize * MAX_SEQ_LEN_REF + MAX_LINE_LEN, 64);
    m.seqBufLeftQer[0] = (uint8_t *)_mm_malloc(wsize * MAX_SEQ_LEN_QER + MAX_LINE_LEN, 64);
    m.seqBufRightRef[0] = (uint8_t *)_mm_malloc(wsize * MAX_SEQ_LEN_REF + MAX_LINE_LEN, 64);
    m.seqBufRightQer[0] = (uint8_t *)_mm_malloc(wsize * MAX_SEQ_LEN_QER + MAX_LINE_LEN, 64);
    m.wsize_buf_ref[0] = wsize * MAX_SEQ_LEN_REF;
    m.wsize_buf_qer[0] = wsize * MAX_SEQ_LEN_QER;
    m.seqPairArrayAux[0] = (SeqPair *)malloc((wsize + MAX_LINE_LEN) * sizeof(SeqPair));
    m.seqPairArrayLeft128[0] = (SeqPair *)malloc((wsize + MAX_LINE_LEN) * sizeof(SeqPair));
    m.seqPairArrayRight128[0] = (SeqPair *)malloc((wsize + MAX_LINE_LEN) * sizeof(SeqPair));
    m.wsize[0] = wsize;
    m.wsize_mem[0] = (int64_t)BATCH_MUL * BATCH_SIZE * readLen;
    m.matchArray[0] = (SMEM *)_mm_malloc(m.wsize_mem[0] * sizeof(SMEM), 64);
    m.min_intv_ar[0] = (int32_t *)malloc(m.wsize_mem[0] * sizeof(int32_t));
    m.query_pos_ar[0] = (int16_t *)malloc(m.wsize_mem[0] * sizeof(int16_t));
    m.enc_qdb[0] = (uint8_t *)malloc(m.wsize_mem[0] * sizeof(uint8_t));
    m.rid[0] = (int32_t *)malloc(m.wsize_mem[0] * sizeof(int32_t));
    m.lim[0] = (int32_t *)_mm_malloc((BATCH_SIZE + 32) * sizeof(int32_t), 64);
    c->seedBuf = (mem_seed_t *)calloc(sizeof(mem_seed_t), (size_t)BATCH_SIZE * AVG_SEEDS_PER_READ);
    return c;
}

void *ref_mem_clone(void *h0);

// Same worker over an index that is ALREADY IN MEMORY (bench.py at human scale: the index is built on the GPU and
// copied back; writing 16 GB of index files only for load_index to read them again would cost minutes).  The arrays
// are what FMI_search::load_index leaves in its members (src/FMI_search.cpp:1251-1351; count[] in the loaded form);
// the caller keeps them alive.  bns carries the contig table only (what mem_chain_seeds / V2 read).
void *ref_mem_open_arrays(int64_t reference_seq_len, const int64_t *count_loaded, void *cp_occ, int8_t *sa_ms_byte,
                          uint32_t *sa_ls_word, int64_t sentinel_index, uint8_t *ref_string, int64_t l_pac, int n_seqs,
                          const int64_t *contig_off, const int32_t *contig_len)
{
    ref_mem_ctx *p = (ref_mem_ctx *)calloc(1, sizeof(ref_mem_ctx));
    FMI_search *f = new FMI_search("in-memory");
    f->one_hot_mask_array = (uint64_t *)_mm_malloc(64 * sizeof(uint64_t), 64);      // as load_index sets it up (:1253-1261)
    f->one_hot_mask_array[0] = 0;
    f->one_hot_mask_array[1] = 0x8000000000000000ULL;
    for (int i = 2; i < 64; ++i) f->one_hot_mask_array[i] = (f->one_hot_mask_array[i - 1] >> 1) | 0x8000000000000000ULL;
    f->reference_seq_len = reference_seq_len;
    for (int i = 0; i < 5; ++i) f->count[i] = count_loaded[i];
    f->cp_occ = (CP_OCC *)cp_occ; f->sa_ms_byte = sa_ms_byte; f->sa_ls_word = sa_ls_word;
    f->sentinel_index = sentinel_index;
    bntseq_t *bns = (bntseq_t *)calloc(1, sizeof(bntseq_t));
    bns->l_pac = l_pac; bns->n_seqs = n_seqs; bns->seed = 11;
    bns->anns = (bntann1_t *)calloc(n_seqs, sizeof(bntann1_t));
    for (int i = 0; i < n_seqs; ++i) {
        bns->anns[i].offset = contig_off[i]; bns->anns[i].len = contig_len[i];
        bns->anns[i].name = strdup("chr"); bns->anns[i].anno = strdup("");
    }
    f->idx->bns = bns;
    p->fmi = f; p->ref_string = ref_string;
    ref_mem_ctx *c = (ref_mem_ctx *)ref_mem_clone(p);           // arenas of one worker
    free(p);
    return c;
}

#ifdef SMEM_ACCEL
// FMA tables already in memory (what FMI_search::load_smem_table leaves, src/FMI_search.cpp:279-356); NULL = off
void ref_mem_set_fma(void *h, void *all_smem_table, void *last_smem_table)
{
    ref_mem_ctx *c = (ref_mem_ctx *)h;
    c->fmi->all_smem_table = (all_smem_t *)all_smem_table;
    c->fmi->last_smem_table = (last_smem_t *)last_smem_table;
}
#endif

#ifdef PERFECT_MATCH
// EMF table already in memory (what load_perfect_table leaves in fmi->perfect_table, src/perfect_map.cpp:204-291);
// shared by every clone of this worker.  loc_table / seed_table stay the caller's.
void ref_mem_set_perfect(void *h, int seed_len, uint32_t num_loc_entry, uint32_t num_seed_entry, uint32_t seq_len,
                         uint32_t *loc_table, void *seed_table)
{
    ref_mem_ctx *c = (ref_mem_ctx *)h;
    if (!seed_table) { c->fmi->perfect_table = NULL; perfect_table_seed_len = PT_SEED_LEN_NO_TABLE; return; }
    perfect_table_t *pt = (perfect_table_t *)calloc(1, sizeof(perfect_table_t));
    pt->seed_len = seed_len; pt->num_loc_entry = num_loc_entry; pt->num_seed_entry = num_seed_entry;
#ifdef MEMSCALE
    pt->num_seed_load = num_seed_entry;
#endif
    pt->ref_string = c->ref_string; pt->loc_table = loc_table; pt->seed_table = (seed_entry_t *)seed_table;
    pt->seq_len = seq_len;
    c->fmi->perfect_table = pt;
    perfect_table_seed_len = seed_len;
}
#endif

// a second worker on the same (read-only) index: own arenas, shared FMI_search / reference text,
// like the reference's threads sharing w.fmi (src/bwamem.h:228-253)
void *ref_mem_clone(void *h0)
{
    ref_mem_ctx *p = (ref_mem_ctx *)h0;
    ref_mem_ctx *c = (ref_mem_ctx *)calloc(1, sizeof(ref_mem_ctx));
    c->fmi = p->fmi; c->ref_string = p->ref_string;
    const int64_t wsize = BATCH_SIZE * SEEDS_PER_READ;
    const int readLen = 1024;
    mem_cache &m = c->mmc;
    m.seqBufLeftRef[0] = (uint8_t *)_mm_malloc(wsize * MAX_SEQ_LEN_REF + MAX_LINE_LEN, 64);
    m.seqBufLeftQer[0] = (uint8_t *)_mm_malloc(wsize * MAX_SEQ_LEN_QER + MAX_LINE_LEN, 64);
    m.seqBufRightRef[0] = (uint8_t *)_mm_malloc(wsize * MAX_SEQ_LEN_REF + MAX_LINE_LEN, 64);
    m.seqBufRightQer[0] = (uint8_t *)_mm_malloc(wsize * MAX_SEQ_LEN_QER + MAX_LINE_LEN, 64);
    m.wsize_buf_ref[0] = wsize * MAX_SEQ_LEN_REF;
    m.wsize_buf_qer[0] = wsize * MAX_SEQ_LEN_QER;
    m.seqPairArrayAux[0] = (SeqPair *)malloc((wsize + MAX_LINE_LEN) * sizeof(SeqPair));
    m.seqPairArrayLeft128[0] = (SeqPair *)malloc((wsize + MAX_LINE_LEN) * sizeof(SeqPair));
    m.seqPairArrayRight128[0] = (SeqPair *)malloc((wsize + MAX_LINE_LEN) * sizeof(SeqPair));
    m.wsize[0] = wsize;
    m.wsize_mem[0] = (int64_t)BATCH_MUL * BATCH_SIZE * readLen;
    m.matchArray[0] = (SMEM *)_mm_malloc(m.wsize_mem[0] * sizeof(SMEM), 64);
    m.min_intv_ar[0] = (int32_t *)malloc(m.wsize_mem[0] * sizeof(int32_t));
    m.query_pos_ar[0] = (int16_t *)malloc(m.wsize_mem[0] * sizeof(int16_t));
    m.enc_qdb[0] = (uint8_t *)malloc(m.wsize_mem[0] * sizeof(uint8_t));
    m.rid[0] = (int32_t *)malloc(m.wsize_mem[0] * sizeof(int32_t));
    m.lim[0] = (int32_t *)_mm_malloc((BATCH_SIZE + 32) * sizeof(int32_t), 64);
    c->seedBuf = (mem_seed_t *)calloc(sizeof(mem_seed_t), (size_t)BATCH_SIZE * AVG_SEEDS_PER_READ);
    return c;
}

// switch this worker to the ERT path: the k-mer table / tree blob come from ref_ert_open (ref_ert_wrap.cpp)
void ref_mem_use_ert(void *h, uint64_t *kmer_offsets, uint8_t *mlt_table)
{
    ref_mem_ctx *c = (ref_mem_ctx *)h;
    c->fmi->kmer_offsets = kmer_offsets; c->fmi->mlt_table = mlt_table;
    c->use_ert = kmer_offsets != nullptr;
}

int64_t ref_mem_l_pac(void *h) { return ((ref_mem_ctx *)h)->fmi->idx->bns->l_pac; }
int ref_mem_n_seqs(void *h) { return ((ref_mem_ctx *)h)->fmi->idx->bns->n_seqs; }
void ref_mem_contigs(void *h, int64_t *offset, int32_t *len)
{
    const bntseq_t *bns = ((ref_mem_ctx *)h)->fmi->idx->bns;
    for (int i = 0; i < bns->n_seqs; ++i) { offset[i] = bns->anns[i].offset; len[i] = bns->anns[i].len; }
}

struct ref_chain_rec { int32_t seqid, rid, n, w, kept, first, is_alt; float frac_rep; int64_t pos; int64_t seed_off; };
struct ref_seed_rec { int64_t rbeg; int32_t qbeg, len, score, pad_; };
struct ref_reg_rec { int64_t rb, re; int32_t qb, qe, rid, score, truesc, w, seedcov, seedlen0, chain; float frac_rep; };

// opt_i: a, b, o_del, e_del, o_ins, e_ins, pen_clip5, pen_clip3, w, zdrop, min_seed_len, max_occ, max_chain_gap,
//        min_chain_weight, max_chain_extend, split_width (16 ints); opt_f: split_factor, mask_level, drop_ratio
// returns 0; counts in n_out[0..2] = chains, seeds, regs.  Buffers must be large enough (caps in cap[0..2]).
int ref_mem_core(void *h, const int32_t *opt_i, const float *opt_f, int64_t max_mem_intv,
                 const uint8_t *reads, const int64_t *off, const int32_t *len, int nseq,
                 ref_chain_rec *chains, ref_seed_rec *seeds, ref_reg_rec *regs, int64_t *reg_off,
                 const int64_t *cap, int64_t *n_out)
{
    ref_mem_ctx *c = (ref_mem_ctx *)h;
    mem_opt_t *opt = mem_opt_init();
    opt->a = opt_i[0]; opt->b = opt_i[1]; opt->o_del = opt_i[2]; opt->e_del = opt_i[3]; opt->o_ins = opt_i[4];
    opt->e_ins = opt_i[5]; opt->pen_clip5 = opt_i[6]; opt->pen_clip3 = opt_i[7]; opt->w = opt_i[8];
    opt->zdrop = opt_i[9]; opt->min_seed_len = opt_i[10]; opt->max_occ = opt_i[11]; opt->max_chain_gap = opt_i[12];
    opt->min_chain_weight = opt_i[13]; opt->max_chain_extend = opt_i[14]; opt->split_width = opt_i[15];
    opt->split_factor = opt_f[0]; opt->mask_level = opt_f[1]; opt->drop_ratio = opt_f[2];
    opt->max_mem_intv = max_mem_intv;
    bwa_fill_scmat(opt->a, opt->b, opt->mat);
    int64_t nc = 0, ns = 0, nr = 0;
    for (int b0 = 0; b0 < nseq; b0 += BATCH_SIZE) {
        const int nb = nseq - b0 < BATCH_SIZE ? nseq - b0 : BATCH_SIZE;
        bseq1_t *seqs = (bseq1_t *)calloc(nb, sizeof(bseq1_t));
        char name[8] = "r";
        for (int i = 0; i < nb; ++i) {
            seqs[i].l_seq = len[b0 + i];
            seqs[i].id = b0 + i;
            seqs[i].name = name;
            seqs[i].seq = (char *)malloc(len[b0 + i] + 8);      // kernel 1 rewrites seq in place (already 2-bit here)
            memcpy(seqs[i].seq, reads + off[b0 + i], len[b0 + i]);
        }
        mem_chain_v *chain_ar = (mem_chain_v *)malloc(nb * sizeof(mem_chain_v));
        mem_alnreg_v *av = (mem_alnreg_v *)calloc(nb, sizeof(mem_alnreg_v));
        memset(c->seedBuf, 0, sizeof(mem_seed_t) * (size_t)BATCH_SIZE * AVG_SEEDS_PER_READ);
        if (c->use_ert) {
            mem_v smems; kv_init(smems); kv_resize(mem_t, smems, 4096);
            u64v hits; kv_init(hits); kv_resize(uint64_t, hits, 1 << 20);
            mem_kernel1_core_ert(c->fmi, opt, seqs, nb, chain_ar, c->seedBuf, (int64_t)BATCH_SIZE * AVG_SEEDS_PER_READ,
                                 c->ref_string, &smems, &hits, 0);
            kv_destroy(smems); kv_destroy(hits);
        } else
        mem_kernel1_core(c->fmi, opt, seqs, nb, chain_ar, c->seedBuf, (int64_t)BATCH_SIZE * AVG_SEEDS_PER_READ, &c->mmc, 0);
        const int64_t chain_base = nc;
        for (int i = 0; i < nb; ++i)
            for (size_t j = 0; j < chain_ar[i].n; ++j) {
                const mem_chain_t &ch = chain_ar[i].a[j];
                if (nc < cap[0]) {
                    ref_chain_rec &o = chains[nc];
                    o.seqid = b0 + ch.seqid; o.rid = ch.rid; o.n = ch.n; o.w = ch.w; o.kept = ch.kept; o.first = ch.first;
                    o.is_alt = ch.is_alt; o.frac_rep = ch.frac_rep; o.pos = ch.pos; o.seed_off = ns;
                }
                ++nc;
                for (int k = 0; k < ch.n; ++k, ++ns)
                    if (ns < cap[1]) { seeds[ns].rbeg = ch.seeds[k].rbeg; seeds[ns].qbeg = ch.seeds[k].qbeg;
                                       seeds[ns].len = ch.seeds[k].len; seeds[ns].score = ch.seeds[k].score; seeds[ns].pad_ = 0; }
            }
        for (int i = 0; i < nb; ++i) kv_init(av[i]);
        mem_chain2aln_across_reads_V2(opt, c->fmi->idx->bns, c->fmi->idx->pac, seqs, nb, chain_ar, av, &c->mmc, c->ref_string, 0);
        int64_t cbase = chain_base;
        for (int i = 0; i < nb; ++i) {
            reg_off[b0 + i] = nr;
            for (size_t r = 0; r < av[i].n; ++r, ++nr) {
                const mem_alnreg_t &a = av[i].a[r];
                if (nr < cap[2]) {
                    ref_reg_rec &o = regs[nr];
                    o.rb = a.rb; o.re = a.re; o.qb = a.qb; o.qe = a.qe; o.rid = a.rid; o.score = a.score; o.truesc = a.truesc;
                    o.w = a.w; o.seedcov = a.seedcov; o.seedlen0 = a.seedlen0; o.frac_rep = a.frac_rep;
                    o.chain = (int32_t)(cbase + (a.c - chain_ar[i].a));
                }
            }
            cbase += chain_ar[i].n;
        }
        for (int i = 0; i < nb; ++i) {
            for (size_t j = 0; j < chain_ar[i].n; ++j)
                if (chain_ar[i].a[j].m > SEEDS_PER_CHAIN) free(chain_ar[i].a[j].seeds);
            free(chain_ar[i].a); free(av[i].a); free(seqs[i].seq);
        }
        free(chain_ar); free(av); free(seqs);
    }
    reg_off[nseq] = nr;
    n_out[0] = nc; n_out[1] = ns; n_out[2] = nr;
    free(opt);
    return 0;
}

// ---- CIGAR tap: the reference's own mem_reg2aln (src/bwamem.cpp:2533-2627) over caller-provided regions -------------
// job: rb, re, read, qb, qe, truesc, w (the mem_alnreg_t fields mem_reg2aln's alignment part reads).
// out per job: pos, cigar_off, md_off, is_rev, n_cigar, NM, l_md, score (unused: 0), ok; cigar / md are appended.
struct tap_cigar_job { int64_t rb, re; int32_t read, qb, qe, truesc, w, pad_; };
struct tap_cigar_out { int64_t pos, cigar_off, md_off; int32_t is_rev, n_cigar, nm, l_md, score, ok; };
int ref_reg2aln(void *h, const int32_t *opt_i, const uint8_t *reads, const int64_t *off, const int32_t *len,
                const tap_cigar_job *jobs, int64_t n_jobs, tap_cigar_out *outs, uint32_t *cigar, int64_t cigar_cap,
                char *md, int64_t md_cap, int64_t *n_out)
{
    ref_mem_ctx *c = (ref_mem_ctx *)h;
    mem_opt_t *opt = mem_opt_init();
    opt->a = opt_i[0]; opt->b = opt_i[1]; opt->o_del = opt_i[2]; opt->e_del = opt_i[3]; opt->o_ins = opt_i[4];
    opt->e_ins = opt_i[5]; opt->w = opt_i[8];
    bwa_fill_scmat(opt->a, opt->b, opt->mat);
    const bntseq_t *bns = c->fmi->idx->bns;
    int64_t nc = 0, nm = 0;
    for (int64_t j = 0; j < n_jobs; ++j) {
        const tap_cigar_job &b = jobs[j];
        mem_alnreg_t ar;
        memset(&ar, 0, sizeof ar);
        ar.rb = b.rb; ar.re = b.re; ar.qb = b.qb; ar.qe = b.qe; ar.truesc = b.truesc; ar.score = b.truesc; ar.w = b.w;
        ar.secondary = -1;
        int is_rev;
        ar.rid = bns_pos2rid(bns, bns_depos(bns, b.rb < bns->l_pac ? b.rb : b.re - 1, &is_rev));
        mem_aln_t a = mem_reg2aln(opt, bns, c->fmi->idx->pac, len[b.read], (const char *)(reads + off[b.read]), &ar);
        tap_cigar_out &o = outs[j];
        memset(&o, 0, sizeof o);
        o.pos = a.pos; o.is_rev = a.is_rev; o.n_cigar = a.n_cigar; o.nm = a.NM; o.ok = a.cigar != 0;
        o.cigar_off = nc; o.md_off = nm;
        if (a.cigar) {
            const char *mdp = (const char *)(a.cigar + a.n_cigar);
            o.l_md = (int32_t)strlen(mdp);
            for (int k = 0; k < a.n_cigar; ++k, ++nc) if (nc < cigar_cap) cigar[nc] = a.cigar[k];
            for (int k = 0; k <= o.l_md; ++k, ++nm) if (nm < md_cap) md[nm] = mdp[k];
            free(a.cigar);
        }
    }
    n_out[0] = nc; n_out[1] = nm;
    free(opt);
    return (nc <= cigar_cap && nm <= md_cap) ? 0 : -1;
}

#ifdef PERFECT_MATCH
// ---- EMF tap: the reference's own find_perfect_match_entry (src/perfect_map.cpp:638-659) over a table
// file written by its `perfect-index` sub-command.  The file is read here (header, loc_table,
// seed_table back to back, src/perfect_index.cpp:1418-1424) instead of through load_perfect_table,
// which insists on the shm store in the scale flavour.
struct ref_emf_ctx { perfect_table_t pt; uint8_t *blob; uint8_t *ref; };

void *ref_emf_open(const char *table_path, const char *ref0123_path, uint32_t num_seed_load)
{
    ref_emf_ctx *c = new ref_emf_ctx();
    FILE *fp = fopen(table_path, "rb");
    if (!fp) { delete c; return nullptr; }
    fseek(fp, 0, SEEK_END); long sz = ftell(fp); fseek(fp, 0, SEEK_SET);
    c->blob = (uint8_t *)malloc(sz);
    if (fread(c->blob, 1, sz, fp) != (size_t)sz) { fclose(fp); free(c->blob); delete c; return nullptr; }
    fclose(fp);
    memcpy(&c->pt, c->blob, sizeof(perfect_table_t));
    c->pt.loc_table = (uint32_t *)(c->blob + sizeof(perfect_table_t));
    c->pt.seed_table = (seed_entry_t *)(c->blob + sizeof(perfect_table_t) + sizeof(uint32_t) * (size_t)c->pt.num_loc_entry);
#ifdef MEMSCALE
    c->pt.num_seed_load = (num_seed_load > 0 && num_seed_load < c->pt.num_seed_entry) ? num_seed_load : c->pt.num_seed_entry;
#endif
    fp = fopen(ref0123_path, "rb");
    if (!fp) { free(c->blob); delete c; return nullptr; }
    fseek(fp, 0, SEEK_END); sz = ftell(fp); fseek(fp, 0, SEEK_SET);
    c->ref = (uint8_t *)malloc(sz + 64);
    if (fread(c->ref, 1, sz, fp) != (size_t)sz) { fclose(fp); free(c->blob); free(c->ref); delete c; return nullptr; }
    fclose(fp);
    c->pt.ref_string = c->ref;
    perfect_table_seed_len = c->pt.seed_len;
    return c;
}

void ref_emf_close(void *h) { ref_emf_ctx *c = (ref_emf_ctx *)h; free(c->blob); free(c->ref); delete c; }

void ref_emf_probe(void *h, uint8_t *reads, const int64_t *off, const int32_t *len, int32_t nseq,
                   int32_t *ret, uint32_t *perfect)
{
    ref_emf_ctx *c = (ref_emf_ctx *)h;
    perfect_table_seed_len = c->pt.seed_len;
    for (int32_t i = 0; i < nseq; ++i) {
        bseq1_t s;
        memset(&s, 0, sizeof(s));
        s.seq = (char *)(reads + off[i]);
        s.l_seq = len[i];
        ret[i] = find_perfect_match_entry(&c->pt, &s, len[i]);
        perfect[2 * i] = s.perfect.flags;
        perfect[2 * i + 1] = s.perfect.location;
    }
}
#endif  // PERFECT_MATCH

}  // extern "C"
