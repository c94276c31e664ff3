// bwagpu_shim.cpp — the reference-side binding of libbwagpu.so (C ABI: include/bwagpu.h).
//
// This is the ONE file a maintainer adds to etri/bwa-mem-scale to run the seeding + extension hot path on the
// GPU.  It is compiled IN PLACE OF src/bwamem.cpp: it includes that file unmodified (so that its static
// worker_sam stays reachable), renames the reference's mem_process_seqs out of the way and provides a
// mem_process_seqs whose two kt_for passes over worker_bwt / worker_aln (reference src/bwamem.cpp:1872-1876) are
// one call into the library per chunk.  Everything else — FASTQ reading, kt_pipeline, mem_pestat, worker_sam (mapq,
// CIGAR, SAM text), the index loaders, the shm store — is the reference's own code, untouched.
//
//   build (what oracle/Makefile's `gpu_*` flavours do; nothing of the reference is copied or edited):
//     g++ <reference flags> -I<ref>/src -I<repo>/include -c integration/bwagpu_shim.cpp     # instead of src/bwamem.cpp
//     g++ ... <all other reference objects> bwagpu_shim.o -L<repo>/bwa_mem_scale_b200 -lbwagpu
//
// The reference exit()s on every error; so does the shim (a non-zero bwagpu_status is fatal).  There is no CPU
// fallback: without a CUDA device bwagpu_create fails and the process exits.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <string>
#include <vector>
#include <algorithm>
#include <fstream>
#include <iostream>
#include <map>
#include <limits>
#include <immintrin.h>
// FMI_search keeps cp_occ / sa_* / count / the FMA tables private and has no accessors; the upload below needs them.
// (A maintainer would add five getters to FMI_search.h instead of this define.)
#define private public
#include "FMI_search.h"
#undef private

#define mem_process_seqs mem_process_seqs_reference      /* the reference's own driver stays available under this name */
#include "bwamem.cpp"
#undef mem_process_seqs

#include "bwagpu.h"

static_assert(sizeof(SMEM) == sizeof(bwagpu_smem), "SMEM layout");
static_assert(sizeof(SeqPair) == sizeof(bwagpu_seqpair), "SeqPair layout");
#ifdef PERFECT_MATCH
static_assert(sizeof(seed_entry_t) == sizeof(bwagpu_seed_entry), "seed_entry_t layout");
static_assert(sizeof(bseq1_perfect_t) == sizeof(bwagpu_perfect), "bseq1_perfect_t layout");
#endif

namespace {

struct Shim {
    bwagpu_ctx *ctx = nullptr;
    const FMI_search *fmi = nullptr;          // the index the replica was made from
    // pinned staging buffers, grown on demand
    uint8_t *enc = nullptr; int64_t enc_cap = 0;
    int64_t *off = nullptr; int32_t *len = nullptr; int64_t *reg_off = nullptr; int32_t *emf_ret = nullptr;
    bwagpu_perfect *perfect = nullptr; int64_t seq_cap = 0;
    bwagpu_alnreg *regs = nullptr; int64_t regs_cap = 0;
    int has_emf = 0;
} g;

[[noreturn]] void die(const char *what)
{
    fprintf(stderr, "[bwagpu] %s: %s\n", what, bwagpu_last_error(g.ctx));
    exit(EXIT_FAILURE);
}
#define CK(call) do { if ((call) != 0) die(#call); } while (0)

template <typename T> void grow(T *&p, int64_t n)
{
    if (p) CK(bwagpu_host_free(g.ctx, p));
    void *q = nullptr;
    CK(bwagpu_host_alloc(g.ctx, n * (int64_t)sizeof(T), &q));
    p = (T *)q;
}

// once per process: replicate the index the reference has just loaded (src/fastmap.cpp:1187-1234) in HBM
void init(worker_t &w)
{
    FMI_search *fmi = w.fmi;
    const double t_init = realtime();
    int dev = getenv("BWAGPU_DEVICE") ? atoi(getenv("BWAGPU_DEVICE")) : 0;
    if (bwagpu_create(dev, &g.ctx)) die("bwagpu_create (libbwagpu has no CPU path)");
    const bntseq_t *bns = fmi->idx->bns;
    std::vector<int64_t> coff(bns->n_seqs); std::vector<int32_t> clen(bns->n_seqs); std::vector<uint8_t> calt(bns->n_seqs);
    for (int i = 0; i < bns->n_seqs; ++i) { coff[i] = bns->anns[i].offset; clen[i] = bns->anns[i].len; calt[i] = bns->anns[i].is_alt; }
    CK(bwagpu_index_load_ref(g.ctx, w.ref_string, bns->l_pac, bns->n_seqs, coff.data(), clen.data(), calt.data()));
    if (!w.useErt) {
        bwagpu_fm_index_desc d;
        d.reference_seq_len = fmi->reference_seq_len;
        for (int i = 0; i < 5; ++i) d.count[i] = fmi->count[i];            // loaded form (+1), FMI_search.cpp:1304-1307
        d.cp_occ = fmi->cp_occ; d.sa_ms_byte = fmi->sa_ms_byte; d.sa_ls_word = fmi->sa_ls_word;
        d.sentinel_index = fmi->sentinel_index;
        CK(bwagpu_index_load_fm(g.ctx, &d));
#ifdef SMEM_ACCEL
        if (fmi->all_smem_table || fmi->last_smem_table)                   // the planner may have loaded neither, one or both
            CK(bwagpu_index_load_fma(g.ctx, fmi->all_smem_table, fmi->last_smem_table));
#endif
    } else {
        struct stat st;
        std::string p = std::string(fmi->file_name) + ".mlt_table";
        if (stat(p.c_str(), &st)) { fprintf(stderr, "[bwagpu] cannot stat %s\n", p.c_str()); exit(EXIT_FAILURE); }
        CK(bwagpu_index_load_ert(g.ctx, fmi->kmer_offsets, nullptr, nullptr, 0, fmi->mlt_table, (int64_t)st.st_size));
    }
#ifdef PERFECT_MATCH
    if (fmi->perfect_table) {
        const perfect_table_t *pt = fmi->perfect_table;
        bwagpu_emf_desc ed;
        memset(&ed, 0, sizeof ed);
        ed.seed_len = pt->seed_len; ed.num_loc_entry = pt->num_loc_entry; ed.num_seed_entry = pt->num_seed_entry;
#ifdef MEMSCALE
        ed.num_seed_load = pt->num_seed_load;
#endif
        ed.seq_len = pt->seq_len; ed.num_seed_used = pt->num_seed_used; ed.num_seed_key = pt->num_seed_key;
        CK(bwagpu_emf_load(g.ctx, &ed, pt->loc_table, (const bwagpu_seed_entry *)pt->seed_table));
        g.has_emf = 1;
    }
#endif
    g.fmi = fmi;
    fprintf(stderr, "[bwagpu] index replicated in HBM (context + uploads): %.3f s\n", realtime() - t_init);
}

struct Job { worker_t *w; const bwagpu_mem_result *res; };
Job g_job;

// kt_for bodies (signature of src/kthread.cpp:79): per-read host work before / after the GPU call
void pack_worker(void *, long seq_id, long batch_size, int)
{
    bseq1_t *seqs = g_job.w->seqs;
    for (long l = seq_id; l < seq_id + batch_size; ++l) {
        char *seq = seqs[l].seq;
        const int n = seqs[l].l_seq;
        for (int i = 0; i < n; ++i) seq[i] = seq[i] < 4 ? seq[i] : nst_nt4_table[(int)seq[i]];   // bwamem.cpp:1226-1233
        memcpy(g.enc + g.off[l], seq, (size_t)n);
    }
}

void unpack_worker(void *, long seq_id, long batch_size, int tid)
{
    worker_t &w = *g_job.w;
    const mem_opt_t *opt = w.opt;
    const bntseq_t *bns = w.fmi->idx->bns;
    for (long l = seq_id; l < seq_id + batch_size; ++l) {
        mem_alnreg_v *av = &w.regs[l];
        kv_init(*av);
#ifdef PERFECT_MATCH
        w.seqs[l].perfect.exist = 0;
        if (g.has_emf) {
            pprof[tid][g.emf_ret[l]]++;
            w.seqs[l].perfect.flags = g.perfect[l].flags;
            w.seqs[l].perfect.location = g.perfect[l].location;
        } else {
            pprof[tid][FIND_PERFECT_NO_TABLE]++;
        }
#ifndef DO_NORMAL
        if (w.seqs[l].perfect.exist) continue;
#endif
#endif
        const int64_t n = g.reg_off[l + 1] - g.reg_off[l];
        if (n == 0) continue;
        av->m = av->n = (size_t)n;
        av->a = (mem_alnreg_t *)calloc((size_t)n, sizeof(mem_alnreg_t));
        const bwagpu_alnreg *r = g.regs + g.reg_off[l];
        for (int64_t i = 0; i < n; ++i) {                       // the fields mem_chain2aln_across_reads_V2 sets (:2960-2967 ...)
            mem_alnreg_t *a = &av->a[i];
            a->rb = r[i].rb; a->re = r[i].re; a->qb = r[i].qb; a->qe = r[i].qe; a->rid = r[i].rid;
            a->score = r[i].score; a->truesc = r[i].truesc; a->w = r[i].w; a->seedcov = r[i].seedcov;
            a->seedlen0 = r[i].seedlen0; a->frac_rep = r[i].frac_rep;
        }
        av->n = mem_sort_dedup_patch(opt, bns, w.fmi->idx->pac, (uint8_t *)w.seqs[l].seq, (int)av->n, av->a);   // :1462-1468
        for (size_t i = 0; i < av->n; ++i) {                                                                    // :1470-1483
            mem_alnreg_t *p = &av->a[i];
            if (p->rid >= 0 && bns->anns[p->rid].is_alt) p->is_alt = 1;
        }
    }
}

// worker_bwt + worker_aln of one chunk (reference src/bwamem.cpp:1872-1876) on the GPU
void process_chunk(mem_opt_t *opt, int n, bseq1_t *seqs, worker_t &w)
{
    if (!g.ctx) init(w);
    if (w.fmi != g.fmi) { fprintf(stderr, "[bwagpu] the index changed under the shim\n"); exit(EXIT_FAILURE); }
    if (n > g.seq_cap) {
        g.seq_cap = n + n / 4 + 1024;
        grow(g.off, g.seq_cap + 1); grow(g.len, g.seq_cap + 1); grow(g.reg_off, g.seq_cap + 1);
        grow(g.emf_ret, g.seq_cap + 1); grow(g.perfect, g.seq_cap + 1);
    }
    int64_t pos = 0;
    for (int i = 0; i < n; ++i) { g.off[i] = pos; g.len[i] = seqs[i].l_seq; pos += seqs[i].l_seq; }
    if (pos + 64 > g.enc_cap) { g.enc_cap = pos + pos / 4 + 4096; grow(g.enc, g.enc_cap); }
    g_job.w = &w;
    const double t0 = realtime();
    kt_for(pack_worker, &w, n);
    const double t1 = realtime();

    bwagpu_mem_opt mo;
    memset(&mo, 0, sizeof mo);
    mo.seed.min_seed_len = opt->min_seed_len; mo.seed.split_factor = opt->split_factor; mo.seed.split_width = opt->split_width;
    mo.seed.max_occ = opt->max_occ; mo.seed.max_mem_intv = opt->max_mem_intv;
    mo.chain.w = opt->w; mo.chain.max_chain_gap = opt->max_chain_gap; mo.chain.max_occ = opt->max_occ;
    mo.chain.min_seed_len = opt->min_seed_len; mo.chain.min_chain_weight = opt->min_chain_weight;
    mo.chain.max_chain_extend = opt->max_chain_extend; mo.chain.mask_level = opt->mask_level; mo.chain.drop_ratio = opt->drop_ratio;
    mo.ext.a = opt->a; mo.ext.b = opt->b; mo.ext.o_del = opt->o_del; mo.ext.e_del = opt->e_del; mo.ext.o_ins = opt->o_ins;
    mo.ext.e_ins = opt->e_ins; mo.ext.pen_clip5 = opt->pen_clip5; mo.ext.pen_clip3 = opt->pen_clip3; mo.ext.w = opt->w;
    mo.ext.zdrop = opt->zdrop;
    mo.use_ert = w.useErt; mo.use_emf = g.has_emf; mo.batch_size = BATCH_SIZE;

    bwagpu_mem_result res;
    CK(bwagpu_mem_process_chunk(g.ctx, &mo, g.enc, g.off, g.len, n, &res));
    const double t2 = realtime();
    if (res.n_regs + 1 > g.regs_cap) { g.regs_cap = res.n_regs + res.n_regs / 4 + 4096; grow(g.regs, g.regs_cap); }
    CK(bwagpu_mem_fetch(g.ctx, &res, n, g.regs, g.reg_off, g.emf_ret, g.perfect));
    const double t3 = realtime();
    g_job.res = &res;
    kt_for(unpack_worker, &w, n);
    const double t4 = realtime();
    fprintf(stderr, "[bwagpu] chunk of %d reads: %lld EMF hits, %lld SMEMs, %lld chains, %lld SW jobs (%lld retried), %lld regions\n",
            n, (long long)res.n_emf_hits, (long long)res.n_smem, (long long)res.n_chains,
            (long long)(res.n_sw_left + res.n_sw_right), (long long)res.n_sw_retry, (long long)res.n_regs);
    fprintf(stderr, "[bwagpu] chunk times (s): pack %.3f, reads -> regions on the GPU (copies in, scratch growth included) %.3f, "
                    "results out %.3f, mem_alnreg_v + mem_sort_dedup_patch %.3f\n", t1 - t0, t2 - t1, t3 - t2, t4 - t3);
}

}  // namespace

// Same contract as the reference's mem_process_seqs (src/bwamem.cpp:1850-1903); only the first two passes differ.
void mem_process_seqs(mem_opt_t *opt, int64_t n_processed, int n, bseq1_t *seqs, const mem_pestat_t *pes0, worker_t &w)
{
    mem_pestat_t pes[4];
    double ctime = cputime(), rtime = realtime();
    FMI_search *fmi = w.fmi;
    w.opt = opt;
    w.seqs = seqs; w.n_processed = n_processed;
    w.pes = &pes[0];

    uint64_t tim = __rdtsc();
    fprintf(stderr, "[0000] 1+2. worker_bwt + worker_aln on the GPU (libbwagpu)\n");
    process_chunk(opt, n, seqs, w);
    tprof[WORKER10][0] += __rdtsc() - tim;

    if (opt->flag & MEM_F_PE) {                                 // insert-size inference: host barrier, as upstream (:1881-1891)
        if (pes0) memcpy_bwamem(pes, 4 * sizeof(mem_pestat_t), pes0, 4 * sizeof(mem_pestat_t), __FILE__, __LINE__);
        else mem_pestat(opt, fmi->idx->bns->l_pac, n, w.regs, pes);
    }
    tim = __rdtsc();
    fprintf(stderr, "[0000] 3. Calling kt_for - worker_sam\n");
    kt_for(worker_sam, &w, n);
    tprof[WORKER20][0] += __rdtsc() - tim;
    fprintf(stderr, "\t[0000][ M::%s] Processed %d reads in %.3f CPU sec, %.3f real sec\n", __func__, n, cputime() - ctime,
            realtime() - rtime);
}
