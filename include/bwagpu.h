/* bwagpu.h — C ABI of the B200-native seeding + extension hot path of bwa-mem-scale.
 *
 * Drop-in boundary: the reference has no plugin/FFI layer; what the host pipeline
 * calls on this path are C++ member functions of BandedPairWiseSW and FMI_search from
 * worker_bwt / worker_aln (reference src/bwamem.cpp:1489-1543).  Every entry point below names the
 * reference interface it replaces.  Plain pointers and sizes only; structs mirror the reference's
 * byte-for-byte so the host passes its own arrays.  All functions return 0 on success, a negative
 * bwagpu_status otherwise (the reference itself exit()s on error; the shim in INTEGRATION.md maps
 * non-zero to exit(EXIT_FAILURE)).  There is NO CPU fallback: without a CUDA device every call
 * fails with BWAGPU_ERR_NO_DEVICE.
 */
#ifndef BWAGPU_H
#define BWAGPU_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    BWAGPU_OK = 0,
    BWAGPU_ERR_NO_DEVICE = -1,   /* no CUDA device / driver */
    BWAGPU_ERR_CUDA = -2,        /* a CUDA call failed; see bwagpu_last_error */
    BWAGPU_ERR_ARG = -3,         /* invalid argument */
    BWAGPU_ERR_NOMEM = -4,
    BWAGPU_ERR_RANGE = -5        /* input outside what the kernel kind supports (e.g. 8-bit rule) */
} bwagpu_status;

typedef struct bwagpu_ctx bwagpu_ctx;   /* one per GPU: device id, stream, scratch, index replica */

/* ---- context ------------------------------------------------------------------------------- */
int  bwagpu_create(int device, bwagpu_ctx **out);
void bwagpu_destroy(bwagpu_ctx *ctx);
/* Several worker contexts on ONE GPU over one index replica: `ctx` (freshly created, same device) gets its own stream,
 * scratch and error state but works on the FM-index / FMA tables / reference text / EMF table / ERT held by `owner`.
 * This is the reference's threading model (src/bwamem.cpp:1872-1876: every kt_for thread runs worker_bwt/worker_aln on
 * its own slice of the chunk over the shared, read-only index) moved to the GPU: one host thread per context, so that one
 * thread's copies and host-side waits overlap another's kernels.  Destroy the borrower before the owner; a borrower can
 * not load, build or unload index parts (BWAGPU_ERR_ARG). */
int  bwagpu_ctx_share_index(bwagpu_ctx *ctx, const bwagpu_ctx *owner);
const char *bwagpu_last_error(const bwagpu_ctx *ctx);   /* never NULL; ctx may be NULL */
int  bwagpu_device_count(void);                          /* 0 when no driver/device */
int  bwagpu_synchronize(bwagpu_ctx *ctx);                /* wait for ctx's stream */
/* number of kernel launches this context has issued (bench.py's gpu_launches) */
int64_t bwagpu_launch_count(const bwagpu_ctx *ctx);

/* CUDA-event timing on the context's own stream (bench.py times kernels with these) */
int bwagpu_timer_start(bwagpu_ctx *ctx);
int bwagpu_timer_stop(bwagpu_ctx *ctx, float *elapsed_ms);   /* records, synchronises, reports */

/* Per-kernel device time, for roofline reporting: when profiling is on every launch of the listed
 * kernels is bracketed by CUDA events on the context's stream (this serialises the stream, so
 * leave it off when measuring whole-step throughput).  set_profiling also clears the counters. */
enum { BWAGPU_KERN_SEED_COLLECT = 0, BWAGPU_KERN_SEED_SORT = 1, BWAGPU_KERN_SA_WALK = 2,
       BWAGPU_KERN_BSW = 3, BWAGPU_KERN_BSW_GENERIC = 4, BWAGPU_KERN_FMA_BUILD = 5, BWAGPU_KERN_KSW = 6, BWAGPU_KERN_C2A = 7, BWAGPU_KERN_CHAIN = 8, BWAGPU_KERN_EMF = 9,
       BWAGPU_KERN_ERT_SEED = 10, BWAGPU_KERN_ERT_EMIT = 11, BWAGPU_KERN_CIGAR = 12,
       BWAGPU_KERN_N = 16 };
int bwagpu_set_profiling(bwagpu_ctx *ctx, int on);
int bwagpu_kernel_time(const bwagpu_ctx *ctx, int which, double *ms, int64_t *launches);

/* Measured issue rate (G thread-operations / s over the whole GPU) of one integer instruction, from a register-only
 * loop: the denominator of the banded-SW kernel's roofline.  IADD = IADD3, VIMAX3 = __vimax3_s32,
 * VIADDMAX = __viaddmax_s32_relu, VIADDMAX_S16X2 = __viaddmax_s16x2_relu (two int16 lanes per operation),
 * ADDMAX_PLAIN = max(a + b, c) without DPX. */
enum { BWAGPU_PEAK_IADD = 0, BWAGPU_PEAK_VIMAX3 = 1, BWAGPU_PEAK_VIADDMAX = 2, BWAGPU_PEAK_VIADDMAX_S16X2 = 3,
       BWAGPU_PEAK_ADDMAX_PLAIN = 4 };
int bwagpu_measure_int_peak(bwagpu_ctx *ctx, int which, double *gops);

/* Design probe (not a product path): FMI_search::backwardExt walks along the resident text, one walk per thread (mode 0,
 * what the seeding kernel does) or per 8-lane group with coalesced 8-byte loads and shuffles (mode 1, the design
 * BASELINE.json's north_star describes); CP_OCC records fetched per second and a checksum of the final intervals
 * (equal between modes).  bench.py --workload peaks records both next to the integer issue rates. */
int bwagpu_probe_bext(bwagpu_ctx *ctx, int mode, int64_t n_walks, int32_t steps, double *records_per_s,
                      uint64_t *checksum, float *ms);

/* raw device memory helpers for callers that keep inputs resident in HBM */
int bwagpu_malloc(bwagpu_ctx *ctx, int64_t bytes, void **dptr);
int bwagpu_free(bwagpu_ctx *ctx, void *dptr);
int bwagpu_h2d(bwagpu_ctx *ctx, void *dst_dev, const void *src_host, int64_t bytes);
int bwagpu_d2h(bwagpu_ctx *ctx, void *dst_host, const void *src_dev, int64_t bytes);
int bwagpu_memset(bwagpu_ctx *ctx, void *dst_dev, int value, int64_t bytes);   /* async on ctx's stream */
/* Page-locked host memory for the caller's batch buffers (the reference keeps its per-thread arenas
 * for the whole run, src/fastmap.cpp:280-300; pinning them once makes every copy a DMA at PCIe speed),
 * and copies that only enqueue on the context's stream (pair with bwagpu_synchronize). */
int bwagpu_host_alloc(bwagpu_ctx *ctx, int64_t bytes, void **hptr);
int bwagpu_host_free(bwagpu_ctx *ctx, void *hptr);
int bwagpu_h2d_async(bwagpu_ctx *ctx, void *dst_dev, const void *src_host, int64_t bytes);
int bwagpu_d2h_async(bwagpu_ctx *ctx, void *dst_host, const void *src_dev, int64_t bytes);
/* Copies on the context's SECOND stream, for a double-buffered batch loop (what the reference's
 * kt_pipeline does with its read / compute / write steps, src/fastmap.cpp:430-640):
 *   bwagpu_h2d_overlap : starts copying the NEXT batch into a buffer the running batch does not use
 *   bwagpu_h2d_fence   : later work on the main stream waits for that copy (call before using the buffer)
 *   bwagpu_d2h_overlap : copies a finished result out, ordered after everything enqueued so far; the
 *                        stage that next rewrites the source buffer waits for the copy by itself
 *   bwagpu_copy_synchronize : host waits for the second stream (the copied-out result is then valid) */
int bwagpu_h2d_overlap(bwagpu_ctx *ctx, void *dst_dev, const void *src_host, int64_t bytes);
int bwagpu_h2d_fence(bwagpu_ctx *ctx);
int bwagpu_d2h_overlap(bwagpu_ctx *ctx, void *dst_host, const void *src_dev, int64_t bytes);
int bwagpu_copy_synchronize(bwagpu_ctx *ctx);

/* ---- S2: banded Smith-Waterman extension ------------------------------------------------------
 * bwagpu_seqpair mirrors SeqPair (reference src/bandedSWA.h:90-99), 56 bytes. */
typedef struct {
    int32_t idr, idq, id;
    int32_t len1, len2;
    int32_t h0;
    int32_t seqid, regid;
    int32_t score, tle, gtle, qle;
    int32_t gscore, max_off;
} bwagpu_seqpair;

/* bwagpu_smem mirrors SMEM (reference src/FMI_search.h:85-93, non-DEBUG build), 40 bytes */
typedef struct {
    uint32_t rid, m, n, pad_;
    int64_t k, l, s;
} bwagpu_smem;

/* constructor arguments of BandedPairWiseSW (reference src/bandedSWA.cpp:51-98 as called from
 * src/bwamem.cpp:3210-3216): w_match = opt->a, w_mismatch = opt->b (positive), end_bonus =
 * opt->pen_clip5 (left) or pen_clip3 (right); the ambiguous-base score is fixed at -1. */
typedef struct {
    int32_t o_del, e_del, o_ins, e_ins, zdrop, end_bonus, w_match, w_mismatch;
} bwagpu_bsw_params;

#define BWAGPU_BSW_SCALAR 0   /* scalarBandedSWAWrapper  (src/bandedSWA.cpp:242-262) */
#define BWAGPU_BSW_8BIT   8   /* getScores8              (src/bandedSWA.cpp:4172-4837) */
#define BWAGPU_BSW_16BIT  16  /* getScores16             (src/bandedSWA.cpp:3448-4121) */

/* Replaces BandedPairWiseSW::getScores16 / getScores8 / scalarBandedSWAWrapper
 * (reference src/bandedSWA.h:130-172): same pairArray / seqBufRef / seqBufQer / numPairs / w;
 * results (score, tle, gtle, qle, gscore, max_off) are written into pairArray in place and are
 * bit-identical to the reference's per-lane results.  HOST buffers; copies are inside the call.
 * Unlike the reference it neither writes SIMD pad entries after numPairs nor reads past them. */
int bwagpu_bsw_get_scores(bwagpu_ctx *ctx, const bwagpu_bsw_params *prm, int kind,
                          bwagpu_seqpair *pairArray, const uint8_t *seqBufRef,
                          const uint8_t *seqBufQer, int32_t numPairs, int32_t w);

/* Same computation with every buffer already resident in HBM (device pointers); asynchronous on
 * the context's stream.  ref_bytes / qer_bytes are the extents of the two sequence arenas. */
int bwagpu_bsw_get_scores_device(bwagpu_ctx *ctx, const bwagpu_bsw_params *prm, int kind,
                                 bwagpu_seqpair *d_pairArray, const uint8_t *d_seqBufRef,
                                 int64_t ref_bytes, const uint8_t *d_seqBufQer, int64_t qer_bytes,
                                 int32_t numPairs, int32_t w);

/* DP cells computed by every banded-SW launch of this context since the last reset: the sum over pairs and rows of
 * (end - beg), i.e. one count per execution of the inner-loop body — what the reference counts as SW_cells
 * (src/bandedSWA.cpp:175-200, :199) — accumulated on the device.  bench.py's GCUPS numerator. */
int bwagpu_bsw_cells(bwagpu_ctx *ctx, int reset, int64_t *cells);

/* No counterpart in the reference.
 * The same banded-SW extension for callers that cannot allocate (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc): this call issues no cudaMalloc / cudaFree and no allocating library
 * routine, and does not synchronise.  Device pointers in, results written into d_pairArray in place, asynchronous on
 * the context's stream, bit-identical to bwagpu_bsw_get_scores_device.  The DP column state lives in one
 * file-scope device array of BWAGPU_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.
 * max_len2 = the caller's bound on len2 over the batch (it sizes the per-thread columns; a pair with len2 > max_len2 is
 * not computed and gets -1 in all six result fields); max_len2 > BWAGPU_NOALLOC_MAX_LEN2 is rejected with
 * BWAGPU_ERR_RANGE.  numPairs and len1 are unbounded.  All calls on one GPU share the array: calls that overlap
 * on that GPU (other streams or contexts) race; calls on one context are ordered by its stream.  The array keeps the
 * previous call's contents, which is harmless: every pair initialises the columns it reads. */
#define BWAGPU_NOALLOC_SCRATCH_BYTES (16 << 20)
#define BWAGPU_NOALLOC_MAX_LEN2 1020
int bwagpu_bsw_get_scores_noalloc(bwagpu_ctx *ctx, const bwagpu_bsw_params *prm, int kind,
                                  bwagpu_seqpair *d_pairArray, const uint8_t *d_seqBufRef,
                                  const uint8_t *d_seqBufQer, int32_t numPairs, int32_t max_len2, int32_t w);

/* ---- S2': chains -> alignment regions ---------------------------------------------------------
 * The whole of mem_chain2aln_across_reads_V2 (reference src/bwamem.cpp:2810-3745) on the device:
 * extension jobs are built from the chains, the read and the HBM-resident reference text, run through
 * the banded-SW kernels with the reference's 8/16-bit/scalar routing and two-try band doubling,
 * applied to the regions, and the containment filter purges redundant regions.
 *
 * bwagpu_seed  = the payload of mem_seed_t        (reference src/bwamem.h:127-139)
 * bwagpu_chain = mem_chain_t with seeds flattened (reference src/bwamem.h:141-148): seeds of chain c
 *                are seeds[c.seed_off .. c.seed_off + c.n); chains come in read order (chain_ar[0],
 *                chain_ar[1], ...) with their seeds laid out contiguously in that order
 * bwagpu_alnreg = the mem_alnreg_t fields V2 sets (reference src/bwamem.h:152-173); `chain` is the
 *                index of the owning chain (the reference stores the pointer a->c) */
typedef struct { int64_t rbeg; int32_t qbeg, len, score, pad_; } bwagpu_seed;
typedef struct {
    int32_t seqid, rid, n, w, kept, first, is_alt; float frac_rep; int64_t pos, seed_off;
} bwagpu_chain;
typedef struct {
    int64_t rb, re; int32_t qb, qe, rid, score, truesc, w, seedcov, seedlen0, chain; float frac_rep;
} bwagpu_alnreg;
/* the mem_opt_t fields extension reads (reference src/bwamem.h:89-124) */
typedef struct { int32_t a, b, o_del, e_del, o_ins, e_ins, pen_clip5, pen_clip3, w, zdrop; } bwagpu_ext_opt;

/* Reference text replica: ref_string = the .0123 file (2*l_pac bytes, values 0-3; reference
 * src/fastmap.cpp:773-806) plus bns->anns[i].offset / .len for the n_seqs contigs (src/bntseq.h). */
int bwagpu_index_load_ref(bwagpu_ctx *ctx, const uint8_t *ref_string, int64_t l_pac, int32_t n_seqs,
                          const int64_t *contig_off, const int32_t *contig_len,
                          const uint8_t *contig_is_alt /* bns->anns[i].is_alt; NULL = none */);

/* ---- S1'': seeds -> chains --------------------------------------------------------------------
 * Replaces mem_chain_seeds + mem_chain_flt as worker_bwt runs them (reference src/bwamem.cpp:1341-1363;
 * :789-936, :528-646): the SMEMs (sorted as bwagpu_collect_smem returns them) and their SA
 * coordinates (as bwagpu_get_sa_entries returns them) become, per read, the filtered chain list in
 * the reference's final order — weight-descending as its ks_introsort leaves it — with every field
 * the reference sets (w, kept, first, frac_rep, pos, rid, is_alt).  Output layout = the input layout
 * of bwagpu_chain2aln.  The chain B-tree (kbtree.h) and the unstable sort (ksort.h) are replayed
 * step for step because their tie behaviour is observable. */
typedef struct {
    int32_t w, max_chain_gap, max_occ, min_seed_len, min_chain_weight, max_chain_extend;   /* mem_opt_t */
    float mask_level, drop_ratio;
} bwagpu_chain_opt;
int bwagpu_chain_seeds(bwagpu_ctx *ctx, const bwagpu_chain_opt *opt, const bwagpu_smem *smems, int64_t n_smem,
                       const int64_t *coords, int64_t n_coords, const int32_t *len, int32_t nseq,
                       bwagpu_chain *chains, int64_t chains_cap, bwagpu_seed *seeds, int64_t seeds_cap,
                       int64_t *n_chains, int64_t *n_seeds);                 /* HOST buffers */
typedef struct {
    const bwagpu_chain *d_chains; int64_t n_chains;     /* context-owned, valid until the next chaining call */
    const bwagpu_seed *d_seeds;   int64_t n_seeds;
} bwagpu_chain_result;
/* device-resident inputs: d_smem_off[nseq+1] = SMEM range of each read, d_coord_off[n_smem+1] =
 * coordinate range of each SMEM (both are outputs of bwagpu_seed_fm_device) */
int bwagpu_chain_seeds_device(bwagpu_ctx *ctx, const bwagpu_chain_opt *opt, const bwagpu_smem *d_smems,
                              int64_t n_smem, const int64_t *d_smem_off, const int64_t *d_coords,
                              int64_t n_coords, const int64_t *d_coord_off, const int32_t *d_len,
                              int32_t nseq, int32_t max_len, bwagpu_chain_result *res);

/* The ERT path's chaining: mem_chain_new + mem_chain_flt per read (reference src/bwamem.cpp:960-1050, 1194-1202),
 * over the output of bwagpu_seed_ert_device.  Same B-tree / merge / filter code as bwagpu_chain_seeds_device; the
 * only difference is that mem_chain_new runs read by read, so the batch-wide `pos < num_smem - 1` bound of
 * mem_chain_seeds (:819) does not apply. */
int bwagpu_chain_mems_device(bwagpu_ctx *ctx, const bwagpu_chain_opt *opt, const bwagpu_smem *d_smems,
                             int64_t n_smem, const int64_t *d_smem_off, const int64_t *d_coords,
                             int64_t n_coords, const int64_t *d_coord_off, const int32_t *d_len,
                             int32_t nseq, int32_t max_len, bwagpu_chain_result *res);

/* regs[n_seeds]: the av_v records in creation order — region of the k-th best seed (descending
 * score, ties by descending index) of chain c at regs[c.seed_off + k]; purged regions have
 * qb = qe = -1 exactly as the reference leaves them.  HOST buffers. */
int bwagpu_chain2aln(bwagpu_ctx *ctx, const bwagpu_ext_opt *opt, const uint8_t *enc_reads, const int64_t *off,
                     const int32_t *len, int32_t nseq, const bwagpu_chain *chains, int64_t n_chains,
                     const bwagpu_seed *seeds, int64_t n_seeds, bwagpu_alnreg *regs);

typedef struct {
    const bwagpu_alnreg *d_regs; int64_t n_regs;       /* context-owned, valid until the next call */
    int64_t n_left, n_right;                           /* extension jobs built */
    int64_t n_retry_left, n_retry_right;               /* jobs re-run with the doubled band */
    int64_t n_pairs8, n_pairs16, n_pairs_scalar;       /* jobs per kernel kind (first try) */
    int64_t seq_bytes;                                 /* bytes of job sequences materialised (0: the jobs read the reads and the text in place) */
    int64_t sw_cells;                                  /* DP cells the extension kernels computed (see bwagpu_bsw_cells) */
} bwagpu_c2a_result;
/* same with device-resident inputs */
int bwagpu_chain2aln_device(bwagpu_ctx *ctx, const bwagpu_ext_opt *opt, const uint8_t *d_reads,
                            const int64_t *d_off, const int32_t *d_len, int32_t nseq,
                            const bwagpu_chain *d_chains, int64_t n_chains, const bwagpu_seed *d_seeds,
                            int64_t n_seeds, bwagpu_c2a_result *res);

/* ---- S4: CIGAR / NM / MD of final alignment regions -------------------------
 * Replaces the alignment work of mem_reg2aln (reference src/bwamem.cpp:2533-2600): band inference, up to three
 * bwa_gen_cigar2 tries (src/bwa.cpp:380-467) each running ksw_global2 (src/ksw.cpp:558-668) with backtrack, NM / MD,
 * then position, removal of a leading / trailing deletion and soft clips.  One job per region that reaches
 * mem_reg2aln (worker_sam); mapq, flags and the SAM text stay the reference's host code.
 * job: the mem_alnreg_t fields read (rb, re, qb, qe, truesc, w) + the read it belongs to.
 * out: pos = 0-based forward coordinate in the concatenated reference (bns_depos, shifted past a leading deletion),
 *      is_rev, n_cigar BAM-encoded ops (len << 4 | op, op 0 M / 1 I / 2 D / 3 S) at cigar[cigar_off ..], NM, and the
 *      NUL-terminated MD text at md[md_off ..]; ok = 0 where the reference would return a NULL cigar. */
typedef struct { int64_t rb, re; int32_t read, qb, qe, truesc, w, pad_; } bwagpu_cigar_job;
typedef struct { int64_t pos, cigar_off, md_off; int32_t is_rev, n_cigar, nm, l_md, score, ok; } bwagpu_cigar_out;
typedef struct { int8_t mat[25]; int8_t pad_[3]; int32_t a, o_del, e_del, o_ins, e_ins, w; } bwagpu_cigar_opt;
/* HOST buffers.  *n_cigar_total / *md_bytes always report the sizes needed; BWAGPU_ERR_RANGE (outs[] still filled,
 * cigar / md not written) when cigar_cap / md_cap are too small. */
int bwagpu_reg2aln_batch(bwagpu_ctx *ctx, const bwagpu_cigar_opt *opt, const uint8_t *enc_reads, const int64_t *off,
                         const int32_t *len, int32_t nseq, const bwagpu_cigar_job *jobs, int64_t n_jobs,
                         bwagpu_cigar_out *outs, uint32_t *cigar, int64_t cigar_cap, char *md, int64_t md_cap,
                         int64_t *n_cigar_total, int64_t *md_bytes);

/* Device-resident form: jobs, reads and results stay in HBM (results context-owned, valid until the next CIGAR
 * call); the job order and the scratch layout are planned on the device (plan kernel, radix sort, scans). */
typedef struct {
    const bwagpu_cigar_out *d_outs; int64_t n_jobs;
    const uint32_t *d_cigar;        int64_t n_cigar;
    const char *d_md;               int64_t md_bytes;
    int64_t n_gapfree;              /* regions that took bwa_gen_cigar2's gap-free shortcut */
} bwagpu_cigar_result;
int bwagpu_reg2aln_device(bwagpu_ctx *ctx, const bwagpu_cigar_opt *opt, const uint8_t *d_reads, const int64_t *d_off,
                          const int32_t *d_len, int32_t nseq, const bwagpu_cigar_job *d_jobs, int64_t n_jobs,
                          bwagpu_cigar_result *res);

/* ---- S3: mate-rescue local Smith-Waterman ------------------------------------------------------
 * bwagpu_kswr mirrors kswr_t (reference src/ksw.h), 28 bytes. */
typedef struct {
    int32_t score, te, qe, score2, te2, tb, qb;
} bwagpu_kswr;

/* arguments of ksw_align2 that are constant over a batch (reference src/ksw.cpp:347-350):
 * m = 5, mat = opt->mat (5x5, row = target base, column = query base; bwa_fill_scmat
 * src/bwa.cpp:368-378), gap penalties as mem_matesw passes them (src/bwamem_pair.cpp:217-219). */
typedef struct {
    int8_t mat[25];
    int8_t pad_[3];
    int32_t o_del, e_del, o_ins, e_ins;
} bwagpu_ksw_params;

/* Replaces the per-job ksw_align2 calls of mate rescue — the body of mem_sam_pe_batch (reference
 * src/bwamem_pair.cpp:880-979, whose own sequential form is :891-902) and the call in mem_matesw
 * (:217): for job i, query = seqBufQer + idq (len2 bases), target = seqBufRef + idr (len1 bases),
 * xtra = h0 (KSW_XBYTE 0x10000 selects the 8-bit kernel, KSW_XSUBO 0x40000 | minimum score,
 * KSW_XSTART 0x80000 = also find (tb, qb) with the reversed pass); aln[i] = the kswr_t that
 * ksw_align2(len2, query, len1, target, 5, mat, o_del, e_del, o_ins, e_ins, xtra, 0) returns,
 * bit-identical to the reference's striped SSE2 kernels.  Inputs are not modified (the reference
 * reverses and restores them in place).  HOST buffers; queries up to 4096 bases. */
int bwagpu_ksw_align2_batch(bwagpu_ctx *ctx, const bwagpu_ksw_params *prm, const bwagpu_seqpair *pairArray,
                            const uint8_t *seqBufRef, const uint8_t *seqBufQer, bwagpu_kswr *aln,
                            int32_t numPairs);
/* same with every buffer resident in HBM; asynchronous on the context's stream after one scalar read-back */
int bwagpu_ksw_align2_batch_device(bwagpu_ctx *ctx, const bwagpu_ksw_params *prm,
                                   const bwagpu_seqpair *d_pairArray, const uint8_t *d_seqBufRef,
                                   const uint8_t *d_seqBufQer, bwagpu_kswr *d_aln, int32_t numPairs);

/* ---- S1: FM-index seeding --------------------------------------------------------------------
 * bwagpu_smem mirrors SMEM (reference src/FMI_search.h:85-93, non-DEBUG build), 40 bytes. */

/* The arrays FMI_search::load_index leaves in host memory (reference src/FMI_search.cpp:1251-1351,
 * members at src/FMI_search.h:230-254).  count[] is the LOADED form (file value + 1,
 * cpp:1304-1307).  bwagpu_index_load_fm copies them into this GPU's HBM (one replica per GPU). */
typedef struct {
    int64_t reference_seq_len;          /* 2*l_pac + 1 */
    int64_t count[5];
    const void *cp_occ;                 /* CP_OCC[(reference_seq_len >> 6) + 1], 64 B each */
    const int8_t *sa_ms_byte;           /* [(reference_seq_len >> 3) + 1] */
    const uint32_t *sa_ls_word;         /* [(reference_seq_len >> 3) + 1] */
    int64_t sentinel_index;
} bwagpu_fm_index_desc;

int bwagpu_index_load_fm(bwagpu_ctx *ctx, const bwagpu_fm_index_desc *desc);

/* Builds the FM-index replica in HBM from the resident reference text (bwagpu_index_load_ref) — what the reference's
 * `index` sub-command computes for the search side (FMI_search::build_index / build_fm_index, src/FMI_search.cpp:611-849:
 * suffix array of the forward + reverse-complement text, BWT, CP_OCC, SA sampled every 8th row), same bytes.  Bucketed
 * radix sort of suffix prefixes with segmented refinement of repeats; no whole-text suffix array is held, so the
 * 2 x 3.1 G-symbol text of the human-scale configuration is indexed in HBM in seconds.  Not on the per-batch hot path.
 * bwagpu_index_get_fm copies the resident index out (any pointer may be NULL): count[5] in the LOADED form,
 * cp_occ[(len >> 6) + 1] x 64 B, sa_ms_byte / sa_ls_word[(len >> 3) + 1] — the members of FMI_search after load_index. */
int bwagpu_index_build_fm(bwagpu_ctx *ctx);
int bwagpu_index_get_fm(bwagpu_ctx *ctx, int64_t *reference_seq_len, int64_t *count, void *cp_occ,
                        int8_t *sa_ms_byte, uint32_t *sa_ls_word, int64_t *sentinel_index);

/* FMA ("smem-table") acceleration tables: all_smem_t[4^11] (128 B each) and last_smem_t[4^13]
 * (16 B each), reference src/FMI_search.h:99-135.  They memoise the first forward extensions of a
 * pivot (src/FMI_search.cpp:1414-1463) and of a round-3 walk (:1708-1752).
 *   bwagpu_index_build_fma : builds both tables in HBM from the loaded FM-index (what the
 *                            reference's `smem-table` sub-command, build_smem_tables :229-277, does)
 *   bwagpu_index_load_fma  : uploads tables read from <prefix>.all_smem.11 / .last_smem.13 (HOST
 *                            pointers; either may be NULL = leave that table off)
 *   bwagpu_index_get_fma   : copies the resident tables back (HOST buffers of 512 MiB / 1 GiB)
 *   bwagpu_index_use_fma   : on/off switch for the seeding kernels (the planner of the reference may
 *                            run with or without them, src/bwa_shm.cpp:1703-1717) */
#define BWAGPU_FMA_ALL_BYTES  ((int64_t)128 << 22)
#define BWAGPU_FMA_LAST_BYTES ((int64_t)16 << 26)
int bwagpu_index_build_fma(bwagpu_ctx *ctx);
int bwagpu_index_load_fma(bwagpu_ctx *ctx, const void *all_smem_table, const void *last_smem_table);
int bwagpu_index_get_fma(bwagpu_ctx *ctx, void *all_smem_table, void *last_smem_table);
int bwagpu_index_use_fma(bwagpu_ctx *ctx, int on);

/* the mem_opt_t fields the seeding stage reads (reference src/bwamem.h:89-124) */
typedef struct {
    int32_t min_seed_len;     /* opt->min_seed_len  (19) */
    float   split_factor;     /* opt->split_factor  (1.5) */
    int32_t split_width;      /* opt->split_width   (10) */
    int32_t max_occ;          /* opt->max_occ       (500) */
    int64_t max_mem_intv;     /* opt->max_mem_intv  (20); 0 disables the third round */
} bwagpu_seed_opt;

/* Replaces mem_collect_smem (reference src/bwamem.cpp:648-786), i.e. FMI_search::
 * getSMEMsAllPosOneThread + getSMEMsOnePosOneThread + bwtSeedStrategyAllPosOneThread + sortSMEMs
 * and the per-read introsort: writes the SMEMs of all reads in the reference's final order
 * (rid, m, n ascending, duplicates kept).  enc_reads: 2-bit codes (>= 4 = N), read i at
 * enc_reads[off[i] .. off[i]+len[i]).  skip (may be NULL): non-zero = read answered by the EMF
 * (seq_[i].perfect.exist).  HOST buffers.  *tot_smem is always the full count; if it exceeds
 * out_cap nothing is written and BWAGPU_ERR_RANGE is returned. */
int bwagpu_collect_smem(bwagpu_ctx *ctx, const bwagpu_seed_opt *opt, const uint8_t *enc_reads,
                        const int64_t *off, const int32_t *len, const uint8_t *skip, int32_t nseq,
                        bwagpu_smem *out, int64_t out_cap, int64_t *tot_smem);

/* Replaces FMI_search::get_sa_entries_prefetch (reference src/FMI_search.cpp:2261-2379) over a
 * whole batch: for each SMEM the rows k, k+step, ... (at most max_occ), coordinates in SMEM
 * order then row order, as mem_chain_seeds consumes them (src/bwamem.cpp:866-893).  HOST buffers. */
int bwagpu_get_sa_entries(bwagpu_ctx *ctx, const bwagpu_smem *smems, int64_t n_smem,
                          int32_t max_occ, int64_t *coords, int64_t coords_cap, int64_t *n_coords);

/* Both stages with every buffer resident in HBM (device pointers in, context-owned device
 * buffers out; asynchronous except for two scalar read-backs).  Results stay valid until the
 * next seeding call on this context. */
typedef struct {
    const bwagpu_smem *d_smems; int64_t n_smem;
    const int64_t *d_coords;    int64_t n_coords;
    const int64_t *d_coord_off;          /* [n_smem + 1] first coordinate of each SMEM */
    int64_t n_ext, n_cp_records;         /* backward extensions done, CP_OCC records fetched */
    int64_t n_lf_steps;                  /* LF steps of the SA stage */
    const int64_t *d_smem_off;           /* [nseq + 1] first SMEM of each read */
} bwagpu_seed_result;
int bwagpu_seed_fm_device(bwagpu_ctx *ctx, const bwagpu_seed_opt *opt, const uint8_t *d_reads,
                          const int64_t *d_off, const int32_t *d_len, const uint8_t *d_skip,
                          int32_t nseq, int32_t max_len, bwagpu_seed_result *res);

/* ---- S1': ERT seeding --------------------------------------------------------------------------
 * Replaces the per-read seeding body of mem_kernel1_core_ert (reference src/bwamem.cpp:1124-1193):
 * get_seeds[_prefix] -> reseed[_prefix] -> last (src/ertseeding.cpp:2925/3062, 3200/3315, 3425) and the
 * ks_introsort(mem_smem_sort_lt) that follows.  The index is the one `index -a ert` writes:
 * <prefix>.kmer_table (2^30 x 8 B, fmi->kmer_offsets) and <prefix>.mlt_table (fmi->mlt_table),
 * uploaded verbatim; the text is the replica of bwagpu_index_load_ref.  Reads longer than 151 bases are
 * rejected with BWAGPU_ERR_RANGE (the reference asserts len <= READ_LEN, bwamem.cpp:1135), and
 * min_seed_len must be >= 19 (k-mer + x-mer; the reference asserts below that, ertseeding.cpp:2847).
 *
 * kmer_offsets: the dense table, or NULL with the non-zero entries given as (sparse_idx, sparse_val) —
 * absent k-mers are all-zero entries (type INVALID), so both forms load the same table. */
int bwagpu_index_load_ert(bwagpu_ctx *ctx, const uint64_t *kmer_offsets, const uint32_t *sparse_idx,
                          const uint64_t *sparse_val, int64_t n_sparse, const uint8_t *mlt_table,
                          int64_t mlt_bytes);
int bwagpu_index_unload_ert(bwagpu_ctx *ctx);
/* Entries of the k-mers that do NOT occur in the reference, computed on the device from the resident FM-index
 * (bwagpu_index_load_fm): such an entry is (LEP mask << 2) | INVALID, where LEP bit i-1 says that the number of
 * occurrences changed when base i was appended (reference src/ertindex.cpp, buildIndex); its pointer bits are never
 * read by seeding and are left 0.  Entries of k-mers that occur are not touched, so an index can be shipped as the
 * entries of the occurring k-mers (sparse form of bwagpu_index_load_ert) + mlt_table and completed here. */
int bwagpu_index_complete_ert(bwagpu_ctx *ctx);
/* read back single k-mer table entries (tests) */
int bwagpu_index_get_ert_kmers(bwagpu_ctx *ctx, const uint32_t *idx, int64_t n, uint64_t *out);
/* the whole resident k-mer table (2^30 x 8 B) back to the host, e.g. to hand a completed table to host code */
int bwagpu_index_get_ert_table(bwagpu_ctx *ctx, uint64_t *kmer_offsets);

/* One maximal exact match of a read: mem_t {start, end, hitcount} (src/ertseeding.h:62-77) with hitbeg
 * rebased into the batch-wide hit list.  hits[] holds, for every MEM, the coordinate mem_chain_new
 * derives from each of its hits (src/bwamem.cpp:991-998: forward / re-fetched hits as they are, hits of
 * backward searches mirrored with the end correction), i.e. s.rbeg in the forward-reverse text; a host
 * shim passes them on as mem_t{forward = 1, hitbeg = hit_off} + u64v hits. */
typedef struct { int32_t rid, start, end, hitcount; int64_t hit_off; } bwagpu_ert_mem;

/* HOST buffers.  MEMs come out read by read in the reference's sorted order (start, end ascending;
 * ties as its unstable introsort leaves them).  mem_off[nseq + 1] (may be NULL) = first MEM of each read.
 * The full counts are always reported; BWAGPU_ERR_RANGE (nothing written) when a capacity is too small. */
int bwagpu_seed_ert(bwagpu_ctx *ctx, const bwagpu_seed_opt *opt, const uint8_t *enc_reads, const int64_t *off,
                    const int32_t *len, const uint8_t *skip, int32_t nseq, bwagpu_ert_mem *mems, int64_t mems_cap,
                    int64_t *hits, int64_t hits_cap, int64_t *mem_off, int64_t *n_mems, int64_t *n_hits);

/* Device-resident form, shaped like bwagpu_seed_fm_device so that chaining follows without leaving HBM:
 * d_smems[i] = {rid, m = start, n = end - 1, k = first coordinate index, l = 0, s = hitcount} and the
 * coordinates are the ones mem_chain_new visits (k = 0, step, 2*step ...; at most opt->max_occ per MEM). */
typedef struct {
    const bwagpu_smem *d_smems; int64_t n_smem;
    const int64_t *d_coords;    int64_t n_coords;
    const int64_t *d_coord_off;          /* [n_smem + 1] */
    const int64_t *d_smem_off;           /* [nseq + 1] */
    int64_t n_index_bytes;               /* k-mer entries, tree bytes, hit records and text bytes read */
    int64_t n_hits;                      /* sum of hitcount over the MEMs */
    int64_t n_retry_reads;               /* reads re-run with larger scratch slabs */
} bwagpu_ert_result;
int bwagpu_seed_ert_device(bwagpu_ctx *ctx, const bwagpu_seed_opt *opt, const uint8_t *d_reads,
                           const int64_t *d_off, const int32_t *d_len, const uint8_t *d_skip,
                           int32_t nseq, int32_t max_len, bwagpu_ert_result *res);

/* ---- S0: EMF, the exact-match filter -----------------------------------------------------------
 * Replaces the per-read loop around find_perfect_match_entry in worker_bwt / mem_kernel1_core
 * (reference src/bwamem.cpp:1083-1119, 1245-1272; src/perfect_map.cpp:638-659 with
 * __find_perfect_match_entry :583-629 and seedmatch_further :528-581).  The table is the one the
 * reference's `perfect-index` sub-command writes to <prefix>.perfect.<seed_len>: a 64-byte header,
 * loc_table[num_loc_entry], seed_table[num_seed_entry] (src/perfect_index.cpp:1418-1424).
 * bwagpu_seed_entry mirrors seed_entry_t (src/perfect.h:81-97); bwagpu_perfect mirrors
 * bseq1_perfect_t (:131-138); bwagpu_emf_desc carries the header fields of perfect_table_t
 * (:175-199) the probe reads. */
typedef struct { uint32_t flags, location, left, right; } bwagpu_seed_entry;
typedef struct { uint32_t flags, location; } bwagpu_perfect;
typedef struct {
    int32_t  seed_len;          /* pt->seed_len */
    uint32_t num_loc_entry;     /* pt->num_loc_entry */
    uint32_t num_seed_entry;    /* pt->num_seed_entry: the hash modulus */
    uint32_t num_seed_load;     /* MEMSCALE partial load (perfect.h:752-758): entries >= this are absent; 0 = all */
    uint32_t seq_len;           /* pt->seq_len = l_pac */
    uint32_t num_seed_used, num_seed_key;   /* statistics kept in the header */
} bwagpu_emf_desc;
enum { BWAGPU_EMF_NO_TABLE = 0, BWAGPU_EMF_WITH_N = 1, BWAGPU_EMF_NOT_MATCHED = 2, BWAGPU_EMF_FW_MATCHED = 3,
       BWAGPU_EMF_RC_MATCHED = 4, BWAGPU_EMF_SEED_ONLY_MATCHED = 5 };   /* FIND_PERFECT_*, perfect.h:902-907 */

/* Copies the table into this GPU's HBM (load_perfect_table, reference src/perfect_map.cpp:204-291).
 * Needs the reference text (bwagpu_index_load_ref) resident first.  HOST pointers. */
int bwagpu_emf_load(bwagpu_ctx *ctx, const bwagpu_emf_desc *desc, const uint32_t *loc_table,
                    const bwagpu_seed_entry *seed_table);
int bwagpu_emf_unload(bwagpu_ctx *ctx);
/* One probe per read.  enc_reads as in bwagpu_collect_smem.  ret[i] = FIND_PERFECT_* code,
 * perfect[i] = seqs[i].perfect ({0,0} unless ret is FW/RC_MATCHED), skip[i] (may be NULL) = 1 when
 * the read is answered here and must be left out of seeding.  With no table loaded every ret is
 * NO_TABLE (bwamem.cpp:1094-1097).  HOST buffers. */
int bwagpu_emf_probe(bwagpu_ctx *ctx, const uint8_t *enc_reads, const int64_t *off, const int32_t *len,
                     int32_t nseq, int32_t *ret, bwagpu_perfect *perfect, uint8_t *skip);
/* same with every buffer resident in HBM; asynchronous on the context's stream */
int bwagpu_emf_probe_device(bwagpu_ctx *ctx, const uint8_t *d_reads, const int64_t *d_off, const int32_t *d_len,
                            int32_t nseq, int32_t *d_ret, bwagpu_perfect *d_perfect, uint8_t *d_skip);
/* Offline table builder on the HOST (no context needed): what `bwa-mem2 perfect-index -l seed_len
 * -s slack` does (reference src/perfect_index.cpp:614-757, 926-990, 1317-1435), same table bytes.
 * ref_string: forward .0123 text, l_pac bases; contigs as in .ann, holes as in .amb.  The two
 * tables are malloc'ed here; release them with bwagpu_emf_free_tables. */
int bwagpu_emf_build(const uint8_t *ref_string, int64_t l_pac, int32_t n_seqs, const int64_t *contig_off,
                     const int32_t *contig_len, int32_t n_holes, const int64_t *hole_off, const int32_t *hole_len,
                     int32_t seed_len, double slack, int32_t n_threads, bwagpu_emf_desc *desc,
                     uint32_t **loc_table, bwagpu_seed_entry **seed_table);
void bwagpu_emf_free_tables(uint32_t *loc_table, bwagpu_seed_entry *seed_table);
/* The same table built in HBM from the resident reference text and contig table (bwagpu_index_load_ref), in seconds at
 * human scale where the sequential insertion above — like the reference's perfect-index — needs minutes to hours.  Same
 * format and same CONTENT (every seed's primary location, strand flag and multi-location lists, src/perfect.h:81-129);
 * collision entries and lists are placed by a parallel rule, so the bytes differ from the reference-built file while
 * find_perfect_match_entry answers every read with the same code and location (the list id in perfect.flags >> 2 indexes
 * THIS table's loc_table).  No holes (.amb) support: use bwagpu_emf_build for references with N runs.
 * bwagpu_emf_get copies the resident table out (pointers may be NULL to read only the descriptor). */
int bwagpu_emf_build_device(bwagpu_ctx *ctx, int32_t seed_len, double slack);
int bwagpu_emf_get(bwagpu_ctx *ctx, bwagpu_emf_desc *desc, uint32_t *loc_table, bwagpu_seed_entry *seed_table);

/* ---- W: one chunk of reads -> per-read alignment regions (the worker-level entry) -------------------
 * Replaces the two kt_for passes of mem_process_seqs (reference src/bwamem.cpp:1872-1876), i.e. every
 * worker_bwt (:1507-1543 -> mem_kernel1_core :1209-1363 / mem_kernel1_core_ert :1056-1207) and worker_aln
 * (:1489-1505 -> mem_kernel2_core :1365-1487) call of a chunk, up to and including the compaction loop that drops
 * regions with qe <= qb (:1449-1460).  Left to the host, on the returned lists: mem_sort_dedup_patch (:1462-1468)
 * and is_alt (:1470-1483), then worker_sam as before.
 *   use_emf    : probe the exact-match filter first (PERFECT_MATCH builds with a table loaded); reads it answers
 *                get perfect[i] / emf_ret[i] and no regions (bwamem.cpp:1245-1272, 1402-1410)
 *   use_ert    : w.useErt — ERT seeding + mem_chain_new instead of SMEM seeding + mem_chain_seeds
 *   batch_size : BATCH_SIZE of kt_for (src/macro.h:63, 512): mem_chain_seeds runs once per batch upstream and its
 *                `pos < num_smem - 1` loop bound (:819) makes a batch whose SMEMs total <= 1 yield no chains;
 *                0 = treat the whole chunk as one batch
 * enc_reads: one byte per base, 0-3 / >= 4 = N (what worker_bwt leaves in seq, :1226-1233); read i at
 * enc_reads[off[i] .. off[i] + len[i]).  Results stay in HBM (context-owned, valid until the next call on this
 * context) and are described by bwagpu_mem_result; bwagpu_mem_fetch copies them into host buffers:
 * regs[n_regs] read by read in creation order (the order of av_v after the compaction), reg_off[nseq + 1],
 * emf_ret[nseq] (FIND_PERFECT_* codes, all NO_TABLE without use_emf), perfect[nseq] (seqs[i].perfect). */
typedef struct {
    bwagpu_seed_opt seed;
    bwagpu_chain_opt chain;
    bwagpu_ext_opt ext;
    int32_t use_ert, use_emf, batch_size, pad_;
} bwagpu_mem_opt;
typedef struct {
    const bwagpu_alnreg *d_regs; int64_t n_regs;
    const int64_t *d_reg_off;
    const int32_t *d_emf_ret; const bwagpu_perfect *d_perfect;     /* NULL without use_emf */
    /* work done, for throughput / roofline reporting */
    int64_t n_emf_hits, n_smem, n_coords, n_chains, n_seeds, n_regs_raw;
    int64_t n_sw_left, n_sw_right, n_sw_retry, n_pairs8, n_pairs16, n_pairs_scalar, sw_cells;
    int64_t n_ext, n_cp_records, n_lf_steps, n_index_bytes;
} bwagpu_mem_result;
int bwagpu_mem_process_chunk(bwagpu_ctx *ctx, const bwagpu_mem_opt *opt, const uint8_t *enc_reads, const int64_t *off,
                             const int32_t *len, int32_t nseq, bwagpu_mem_result *res);          /* HOST reads */
int bwagpu_mem_process_chunk_device(bwagpu_ctx *ctx, const bwagpu_mem_opt *opt, const uint8_t *d_reads,
                                    const int64_t *d_off, const int32_t *d_len, int32_t nseq, int32_t max_len,
                                    bwagpu_mem_result *res);                                     /* reads in HBM */
/* any of the four output pointers may be NULL */
int bwagpu_mem_fetch(bwagpu_ctx *ctx, const bwagpu_mem_result *res, int32_t nseq, bwagpu_alnreg *regs,
                     int64_t *reg_off, int32_t *emf_ret, bwagpu_perfect *perfect);

#ifdef __cplusplus
}
#endif
#endif /* BWAGPU_H */
