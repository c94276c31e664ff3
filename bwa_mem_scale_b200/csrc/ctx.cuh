// Internal: per-GPU context shared by every kernel family of libbwagpu.so.
#pragma once
#include <mutex>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/bwagpu.h"

enum { SCR_BSW_PAIRS = 0, SCR_BSW_REF, SCR_BSW_QER, SCR_BSW_WORK, SCR_MISC0, SCR_MISC1, SCR_MISC2,
       SCR_MISC3, SCR_SEED_READS, SCR_SEED_OFF, SCR_SEED_LEN, SCR_SEED_SKIP, SCR_SEED_TMP,
       SCR_SEED_CNT, SCR_SEED_SCAN, SCR_SEED_PREV, SCR_SEED_OUT, SCR_SA_CNT, SCR_SA_OFF,
       SCR_SA_COORD, SCR_CUB, SCR_STATS, SCR_SA_IN, SCR_BSW_KEYS, SCR_BSW_ORDER,
       SCR_C2A_BK0, SCR_C2A_BK1, SCR_C2A_BK2, SCR_C2A_RETRY, SCR_C2A_CNT, SCR_C2A_RMAX, SCR_C2A_FIRST,
       SCR_C2A_KEYS, SCR_C2A_SRT, SCR_C2A_CHAINOF, SCR_C2A_SIZES, SCR_C2A_OFFS, SCR_C2A_REGS, SCR_C2A_PURGED,
       SCR_C2A_L, SCR_C2A_R, SCR_C2A_LQ, SCR_C2A_LR, SCR_C2A_RQ, SCR_C2A_RR, SCR_C2A_CHAINS, SCR_C2A_SEEDS,
       SCR_CH_SD, SCR_CH_CH, SCR_CH_NODES, SCR_CH_ORD, SCR_CH_CNT, SCR_CH_OFFS, SCR_CH_FRAC,
       SCR_ERT_MEMS, SCR_ERT_HITS, SCR_ERT_KEYS, SCR_ERT_MEMS2, SCR_ERT_HITS2, SCR_ERT_KEYS2, SCR_ERT_CNT,
       SCR_ERT_SLAB, SCR_ERT_LIST, SCR_ERT_COFF,
       SCR_CIG_JOBS, SCR_CIG_OUT, SCR_CIG_SLABOFF, SCR_CIG_EH, SCR_CIG_Z, SCR_CIG_CIGAR, SCR_CIG_MD, SCR_CIG_CNT, SCR_CIG_SCAN,
       SCR_CIG_OUTC, SCR_CIG_OUTM, SCR_CIG_ORDER, SCR_CIG_REDO,
       SCR_MEM_READS, SCR_MEM_OFF, SCR_MEM_LEN, SCR_MEM_EMF_RET, SCR_MEM_PERFECT, SCR_MEM_SKIP, SCR_MEM_CNT, SCR_MEM_KEEP,
       SCR_MEM_POS, SCR_MEM_REGS, SCR_MEM_REGOFF, SCR_BSW_CELLS, SCR_CH_HEAVY, SCR_SEED_OVF, SCR_SEED_TMP2,
       SCR_CH_PERM,
       BWAGPU_NSCRATCH };

struct bwagpu_ctx {
    int device;
    int sm_count;
    cudaStream_t stream;
    cudaEvent_t ev_start, ev_stop;
    // second stream for copies that overlap the kernels of the next batch (bwagpu_*_overlap)
    cudaStream_t copy_stream;
    cudaEvent_t ev_main_done, ev_h2d_done, ev_d2h_done;
    int d2h_pending_slot;        // scratch slot an overlapped D2H is still reading (-1: none)
    int h2d_pending;
    int64_t launches;
    char err[512];
    // grow-only device scratch owned by the context (one chunk of reads at a time is in flight,
    // SURVEY.md §3.1, so no locking is needed)
    void *scratch[BWAGPU_NSCRATCH];
    int64_t scratch_bytes[BWAGPU_NSCRATCH];
    // 1: every index pointer below belongs to another context of the same GPU (bwagpu_ctx_share_index); never freed or replaced here
    int index_borrowed;
    // FM-index replica in this GPU's HBM (bwagpu_index_load_fm)
    bool has_fm;
    void *d_cp_occ, *d_sa_ms, *d_sa_ls;
    int64_t fm_count[5], fm_sentinel, fm_ref_seq_len;
    // FMA tables (all_smem_t[4^11], last_smem_t[4^13]); built on device or uploaded
    void *d_fma_all, *d_fma_last;
    int fma_enabled;            // bwagpu_index_use_fma: 0 = walk CP_OCC only
    // reference text replica (.0123: forward + reverse complement, one base per byte) and contig table
    void *d_ref_string, *d_contig_off, *d_contig_len, *d_contig_alt;
    int64_t l_pac; int n_contigs;
    // EMF table replica (bwagpu_emf_load)
    void *d_emf_loc, *d_emf_seed;
    bwagpu_emf_desc emf;
    // ERT replica (bwagpu_index_load_ert): k-mer table (2^30 x 8 B) and the tree blob, reference format
    void *d_ert_kmer, *d_ert_mlt;
    int64_t ert_mlt_bytes;
    int ert_mem_cap, ert_hit_cap;   // per-read slab capacities of the first ERT pass
    int blocking_sync;          // 1: host waits sleep on a blocking event instead of spinning (BWAGPU_BLOCKING_SYNC=1, bwagpu_stream_sync)
    cudaEvent_t ev_block;
    int bsw_direct;             // set around bsw_launch by bwagpu_bsw_scores_direct: pairs address the reads / the text directly
    int chain_batch_reads;      // unit of mem_chain_seeds' batch-wide SMEM bound (0 = the whole call); set by bwagpu_mem_process_chunk
    int seed_read_cap;          // per-read SMEM capacity of the seeding scratch (grows on overflow)
    // optional per-kernel CUDA-event timing (bwagpu_set_profiling): bench.py's roofline numbers
    int profiling;
    cudaEvent_t kev0, kev1;
    double kern_ms[BWAGPU_KERN_N];
    int64_t kern_n[BWAGPU_KERN_N];
};

// Brackets one kernel launch with events on the context's stream when profiling is on.
#define BWAGPU_KERNEL_BEGIN(ctx) \
    do { if ((ctx)->profiling) cudaEventRecord((ctx)->kev0, (ctx)->stream); } while (0)
#define BWAGPU_KERNEL_END(ctx, which)                                        \
    do {                                                                     \
        (ctx)->launches++;                                                   \
        if ((ctx)->profiling) {                                              \
            float ms__ = 0.f;                                                \
            cudaEventRecord((ctx)->kev1, (ctx)->stream);                     \
            cudaEventSynchronize((ctx)->kev1);                               \
            cudaEventElapsedTime(&ms__, (ctx)->kev0, (ctx)->kev1);           \
            (ctx)->kern_ms[which] += ms__;                                   \
            (ctx)->kern_n[which]++;                                          \
        }                                                                    \
    } while (0)



#define BWAGPU_CUDA(ctx, call)                                                              \
    do {                                                                                    \
        cudaError_t e__ = (call);                                                           \
        if (e__ != cudaSuccess) {                                                           \
            snprintf((ctx)->err, sizeof((ctx)->err), "%s:%d: %s -> %s", __FILE__, __LINE__, \
                     #call, cudaGetErrorString(e__));                                       \
            return BWAGPU_ERR_CUDA;                                                         \
        }                                                                                   \
    } while (0)

// cudaFuncAttributeMaxDynamicSharedMemorySize belongs to the kernel (per device), not to a stream or context: two contexts
// driven by two host threads must never lower each other's bound between setting it and launching.  The bound therefore
// only ever GROWS, under a lock; a launch that needs less stays valid.  One BwagpuSmemBound per kernel (instantiation).
struct BwagpuSmemBound { std::mutex mu; int cur[64]; BwagpuSmemBound() { for (int i = 0; i < 64; ++i) cur[i] = 0; } };
template <class F>
static inline cudaError_t bwagpu_raise_dyn_smem(BwagpuSmemBound &b, int device, F func, int bytes)
{
    std::lock_guard<std::mutex> lk(b.mu);
    if (device < 0 || device >= 64) return cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (bytes <= b.cur[device]) return cudaSuccess;
    const cudaError_t e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) b.cur[device] = bytes;
    return e;
}

// Every host-side wait for the context's stream goes through here.  cudaStreamSynchronize spins; with several ranks and
// worker contexts per host (8 GPUs x 2 contexts on 32 cores) the spinning threads take the cores the other ranks' launches
// need, so the wait can instead sleep on an event created with cudaEventBlockingSync (BWAGPU_BLOCKING_SYNC=1).
static inline cudaError_t bwagpu_stream_sync(bwagpu_ctx *ctx)
{
    if (!ctx->blocking_sync) return cudaStreamSynchronize(ctx->stream);
    const cudaError_t e = cudaEventRecord(ctx->ev_block, ctx->stream);
    return e != cudaSuccess ? e : cudaEventSynchronize(ctx->ev_block);
}

// first statement of every entry point that loads, builds or unloads a part of the index
#define BWAGPU_OWNS_INDEX(ctx)                                                                                      \
    do {                                                                                                            \
        if ((ctx) && (ctx)->index_borrowed) {                                                                       \
            snprintf((ctx)->err, sizeof((ctx)->err), "%s: this context borrows its index (bwagpu_ctx_share_index)", __func__); \
            return BWAGPU_ERR_ARG;                                                                                  \
        }                                                                                                           \
    } while (0)

#define BWAGPU_FAIL(ctx, code, ...)                                  \
    do {                                                             \
        snprintf((ctx)->err, sizeof((ctx)->err), __VA_ARGS__);       \
        return (code);                                               \
    } while (0)

// ensure scratch slot `slot` holds at least `bytes`; contents are NOT preserved on growth
int bwagpu_scratch(bwagpu_ctx *ctx, int slot, int64_t bytes, void **out);
int bwagpu_bsw_scores_direct(bwagpu_ctx *ctx, const bwagpu_bsw_params *prm, int kind, bwagpu_seqpair *d_pairs,
                             const uint8_t *d_ref_string, const uint8_t *d_reads, int32_t numPairs, int32_t w);   // bsw.cu
// CP_OCC records between the reference's word order and the replica's (fmi_core.cuh FM_PAIRED); src == dst: in place (seed.cu)
int bwagpu_fm_convert_layout(bwagpu_ctx *ctx, const void *src, void *dst, int64_t n_cp, int to_paired);
