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
const char *bwagpu_last_error(const bwagpu_ctx *ctx);   /* never NULL; ctx may be NULL */
int  bwagpu_device_count(void);                          /* 0 when no driver/device */
int  bwagpu_synchronize(bwagpu_ctx *ctx);                /* wait for ctx's stream */
/* number of kernel launches this context has issued (bench.py's gpu_launches) */
int64_t bwagpu_launch_count(const bwagpu_ctx *ctx);

/* CUDA-event timing on the context's own stream (bench.py times kernels with these) */
int bwagpu_timer_start(bwagpu_ctx *ctx);
int bwagpu_timer_stop(bwagpu_ctx *ctx, float *elapsed_ms);   /* records, synchronises, reports */

/* raw device memory helpers for callers that keep inputs resident in HBM */
int bwagpu_malloc(bwagpu_ctx *ctx, int64_t bytes, void **dptr);
int bwagpu_free(bwagpu_ctx *ctx, void *dptr);
int bwagpu_h2d(bwagpu_ctx *ctx, void *dst_dev, const void *src_host, int64_t bytes);
int bwagpu_d2h(bwagpu_ctx *ctx, void *dst_host, const void *src_dev, int64_t bytes);

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

#ifdef __cplusplus
}
#endif
#endif /* BWAGPU_H */
