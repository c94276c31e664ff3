// bwagpu_measure_int_peak (include/bwagpu.h): register-only issue-rate micro-benchmarks of the integer / DPX
// instructions the banded-SW cell is made of, the measured denominator of the SW kernel's roofline (SURVEY.md 8d asks
// for a measured __viaddmax issue rate recorded next to MEASURED_PEAKS.json).  Each thread runs ILP independent
// dependency chains of one instruction for `iters` iterations; rate = threads x ILP x iters / time.
#include "ctx.cuh"

#define PEAK_ILP 8
template <int WHICH>
__global__ void __launch_bounds__(256) peak_kernel(int iters, int seed, int *out)
{
    int a[PEAK_ILP], b = seed + threadIdx.x, c = seed * 3 + 1;
#pragma unroll
    for (int k = 0; k < PEAK_ILP; ++k) a[k] = threadIdx.x + k * seed;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < PEAK_ILP; ++k) {
            if (WHICH == BWAGPU_PEAK_IADD) a[k] = a[k] + b + c;                                           // IADD3
            else if (WHICH == BWAGPU_PEAK_VIMAX3) a[k] = __vimax3_s32(a[k], b, c);                        // VIMNMX3
            else if (WHICH == BWAGPU_PEAK_VIADDMAX) a[k] = __viaddmax_s32_relu(a[k], b, c);               // VIADDMNMX .RELU
            else if (WHICH == BWAGPU_PEAK_VIADDMAX_S16X2) a[k] = (int)__viaddmax_s16x2_relu((unsigned)a[k], (unsigned)b, (unsigned)c);
            else a[k] = max(a[k] + b, c);                                                                  // plain add + max
        }
        b ^= i;                                  // keeps the compiler from folding the loop
    }
    int s = 0;
#pragma unroll
    for (int k = 0; k < PEAK_ILP; ++k) s ^= a[k];
    if (s == 0x7fffffff) *out = s;               // never true in practice; keeps the chains alive
}

extern "C" int bwagpu_measure_int_peak(bwagpu_ctx *ctx, int which, double *gops)
{
    if (!ctx || !gops) return BWAGPU_ERR_ARG;
    if (which < 0 || which > BWAGPU_PEAK_ADDMAX_PLAIN) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "measure_int_peak: unknown instruction %d", which);
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    int *d_out;
    int rc = bwagpu_scratch(ctx, SCR_MISC1, 64, (void **)&d_out);
    if (rc) return rc;
    const int blocks = ctx->sm_count * 8, iters = 1 << 15;
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {          // first repetition warms up
        BWAGPU_CUDA(ctx, cudaEventRecord(ctx->kev0, ctx->stream));
        switch (which) {
        case BWAGPU_PEAK_IADD: peak_kernel<BWAGPU_PEAK_IADD><<<blocks, 256, 0, ctx->stream>>>(iters, 3, d_out); break;
        case BWAGPU_PEAK_VIMAX3: peak_kernel<BWAGPU_PEAK_VIMAX3><<<blocks, 256, 0, ctx->stream>>>(iters, 3, d_out); break;
        case BWAGPU_PEAK_VIADDMAX: peak_kernel<BWAGPU_PEAK_VIADDMAX><<<blocks, 256, 0, ctx->stream>>>(iters, 3, d_out); break;
        case BWAGPU_PEAK_VIADDMAX_S16X2: peak_kernel<BWAGPU_PEAK_VIADDMAX_S16X2><<<blocks, 256, 0, ctx->stream>>>(iters, 3, d_out); break;
        default: peak_kernel<BWAGPU_PEAK_ADDMAX_PLAIN><<<blocks, 256, 0, ctx->stream>>>(iters, 3, d_out);
        }
        BWAGPU_CUDA(ctx, cudaEventRecord(ctx->kev1, ctx->stream));
        BWAGPU_CUDA(ctx, cudaEventSynchronize(ctx->kev1));
        float ms = 0;
        BWAGPU_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->kev0, ctx->kev1));
        if (rep > 0 && ms < best) best = ms;
    }
    BWAGPU_CUDA(ctx, cudaGetLastError());
    ctx->launches += 4;
    *gops = (double)blocks * 256 * PEAK_ILP * iters / (best * 1e-3) / 1e9;
    return BWAGPU_OK;
}
