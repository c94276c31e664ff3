// Internal: per-GPU context shared by every kernel family of libbwagpu.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/bwagpu.h"

struct bwagpu_ctx {
    int device;
    int sm_count;
    cudaStream_t stream;
    cudaEvent_t ev_start, ev_stop;
    int64_t launches;
    char err[512];
    // grow-only device scratch owned by the context (one chunk of reads at a time is in flight,
    // SURVEY.md §3.1, so no locking is needed)
    void *scratch[8];
    int64_t scratch_bytes[8];
};

enum { SCR_BSW_PAIRS = 0, SCR_BSW_REF = 1, SCR_BSW_QER = 2, SCR_BSW_WORK = 3,
       SCR_MISC0 = 4, SCR_MISC1 = 5, SCR_MISC2 = 6, SCR_MISC3 = 7 };

#define BWAGPU_CUDA(ctx, call)                                                              \
    do {                                                                                    \
        cudaError_t e__ = (call);                                                           \
        if (e__ != cudaSuccess) {                                                           \
            snprintf((ctx)->err, sizeof((ctx)->err), "%s:%d: %s -> %s", __FILE__, __LINE__, \
                     #call, cudaGetErrorString(e__));                                       \
            return BWAGPU_ERR_CUDA;                                                         \
        }                                                                                   \
    } while (0)

#define BWAGPU_FAIL(ctx, code, ...)                                  \
    do {                                                             \
        snprintf((ctx)->err, sizeof((ctx)->err), __VA_ARGS__);       \
        return (code);                                               \
    } while (0)

// ensure scratch slot `slot` holds at least `bytes`; contents are NOT preserved on growth
int bwagpu_scratch(bwagpu_ctx *ctx, int slot, int64_t bytes, void **out);
