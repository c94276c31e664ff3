// S3 mate-rescue seam of the C ABI: bwagpu_ksw_align2_batch{,_device}  (include/bwagpu.h).
// Replaces ksw_align2 (reference src/ksw.cpp:347-381; kernels ksw_u8 :111-232, ksw_i16 :234-338,
// profile ksw_qinit :62-109) as called per job by the sequential body of mem_sam_pe_batch
// (reference src/bwamem_pair.cpp:891-902) and by mem_matesw (:217).
//
// The reference is Farrar's striped local SW on 128-bit vectors in every build: 16 unsigned
// saturating byte lanes (KSW_XBYTE) or 8 signed 16-bit lanes.  Its output is NOT plain affine SW:
// E(i+1,j) comes from H(i,j) before the lazy-F correction (ksw.cpp:168), and which cells that
// correction touches depends on the striping.  To be bit-exact the kernel keeps the striping: one
// job per group of P = 16 / 8 threads, thread l = SIMD lane l, owning query positions
// l*slen .. l*slen + slen-1 (slen = ceil(qlen / P)); the vector byte-shift is a sub-warp shuffle,
// the lazy-F exit test a sub-warp vote, the row maximum a sub-warp reduction.  Column state
// (H0, H1, E, Hmax, query codes) lives in shared memory in the vectors' own memory order
// (segment j, lane l at j*P + l), so the lanes of a group touch consecutive 16-bit elements.
#include <stdlib.h>
#include "ctx.cuh"

#define KSW_XBYTE 0x10000
#define KSW_XSTOP 0x20000
#define KSW_XSUBO 0x40000
#define KSW_XSTART 0x80000
#define KSW_MAX_QLEN 4096

struct KswParams {
    uint32_t row_lo[5], row_hi[5];    // score-matrix row t as bytes: query code 0..3 | code 4, pad -> 0
    int o_del, e_del, o_ins, e_ins;
    int shift, qmax;                  // ksw_qinit: 256 - min(mat), max(mat)
};

struct KswRes { int score, te, qe, score2, te2; };

__device__ __forceinline__ int ksw_prmt_sx(uint32_t a, uint32_t b, uint32_t code)
{
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(code * 0x1111u + 0x8880u));
    return (int)d;
}

// One pass of ksw_u8 (P = 16) / ksw_i16 (P = 8).  Query position k reads query[qrev > 0 ? qrev-1-k : k];
// target row i reads target[i < trev ? trev-1-i : i]  (the in-place reversals of ksw_align2 :367).
template <int P>
__device__ KswRes ksw_pass(const KswParams &p, const uint8_t *__restrict__ query, int qlen, int qrev,
                           const uint8_t *__restrict__ target, int tlen, int trev, int xtra,
                           uint16_t *__restrict__ sm, int qpad, uint64_t *__restrict__ b, unsigned gmask, int l)
{
    constexpr bool BYTE = P == 16;
    const int slen = (qlen + P - 1) / P, n = slen * P;
    uint16_t *H0 = sm, *H1 = sm + qpad, *E = sm + 2 * qpad, *Hmax = sm + 3 * qpad;
    uint8_t *qs = reinterpret_cast<uint8_t *>(sm + 4 * qpad);
    const int minsc = (xtra & KSW_XSUBO) ? (xtra & 0xffff) : 0x10000;
    const int endsc = (xtra & KSW_XSTOP) ? (xtra & 0xffff) : 0x10000;
    const int oe_del = p.o_del + p.e_del, oe_ins = p.o_ins + p.e_ins, ed = p.e_del, ei = p.e_ins;
    const int shift = p.shift;
    for (int j = 0; j < slen; ++j) {
        const int k = j + l * slen, idx = j * P + l;
        int c = 5;
        if (k < qlen) { c = query[qrev > 0 ? qrev - 1 - k : k]; c = c > 4 ? 4 : c; }
        qs[idx] = (uint8_t)c;
        H0[idx] = 0; E[idx] = 0; Hmax[idx] = 0;
    }
    int gmax = 0, te = -1, n_b = 0;
    int last_val = 0, last_row = 0;              // copy of b[n_b - 1]
    for (int i = 0; i < tlen; ++i) {
        int t = target[i < trev ? trev - 1 - i : i];
        t = t > 4 ? 4 : t;
        const uint32_t rlo = p.row_lo[t], rhi = p.row_hi[t];
        int h = slen > 0 ? (int)H0[(slen - 1) * P + l] : 0;
        h = __shfl_up_sync(gmask, h, 1, P);
        if (l == 0) h = 0;
        int f = 0, mx = 0;
        for (int j = 0; j < slen; ++j) {
            const int idx = j * P + l;
            int e = E[idx];
            const int s = ksw_prmt_sx(rlo, rhi, qs[idx]);
            if (BYTE) { h = min(h + ((s + shift) & 0xFF), 255); h = max(h - shift, 0); }
            else h = max(min(h + s, 32767), -32768);
            h = max(h, e); h = max(h, f); mx = max(mx, h);
            H1[idx] = (uint16_t)h;
            E[idx] = (uint16_t)max(max(e - ed, 0), max(h - oe_del, 0));
            f = max(max(f - ei, 0), max(h - oe_ins, 0));
            h = H0[idx];
        }
        // lazy-F: shift f by one lane, sweep the segments until no lane can still raise an H
        for (int kk = 0; kk < 16; ++kk) {
            f = __shfl_up_sync(gmask, f, 1, P);
            if (l == 0) f = 0;
            bool done = false;
            for (int j = 0; j < slen; ++j) {
                const int idx = j * P + l;
                int hh = H1[idx];
                hh = max(hh, f);
                H1[idx] = (uint16_t)hh;
                hh = max(hh - oe_ins, 0);
                f = max(f - ei, 0);
                if (!__any_sync(gmask, f > hh)) { done = true; break; }
            }
            if (done) break;
        }
        const int imax = (int)__reduce_max_sync(gmask, (unsigned)mx);
        if (imax >= minsc) {
            if (n_b == 0 || last_row + 1 != i) {
                last_val = imax; last_row = i;
                if (l == 0) b[n_b] = ((uint64_t)(uint32_t)imax << 32) | (uint32_t)i;
                ++n_b;
            } else if (last_val < imax) {
                last_val = imax; last_row = i;
                if (l == 0) b[n_b - 1] = ((uint64_t)(uint32_t)imax << 32) | (uint32_t)i;
            }
        }
        if (imax > gmax) {
            gmax = imax; te = i;
            for (int j = 0; j < slen; ++j) Hmax[j * P + l] = H1[j * P + l];
            if (BYTE ? (gmax + shift >= 255 || gmax >= endsc) : (gmax >= endsc)) break;
        }
        uint16_t *S = H1; H1 = H0; H0 = S;
    }
    KswRes r;
    r.score = BYTE ? (gmax + shift < 255 ? gmax : 255) : gmax;
    r.te = te; r.qe = -1; r.score2 = -1; r.te2 = -1;
    if (!BYTE || r.score != 255) {
        // qe: smallest query position holding the maximum of the kept H row
        unsigned key = 0; bool any = false;
        for (int j = 0; j < slen; ++j) {
            const unsigned v = Hmax[j * P + l], pos = (unsigned)(j + l * slen);
            const unsigned kk = (v << 16) | (0xFFFFu - pos);
            if (!any || kk > key) key = kk;
            any = true;
        }
        if (n > 0) {
            key = __reduce_max_sync(gmask, key);
            r.qe = (int)(0xFFFFu - (key & 0xFFFFu));
        }
        if (n_b > 0) {
            __syncwarp(gmask);                  // lane 0's b[] stores -> visible to the group
            const int w = (r.score + p.qmax - 1) / p.qmax, low = te - w, high = te + w;
            unsigned long long best = 0;        // value << 32 | ~index : first entry holding the largest value
            for (int x = l; x < n_b; x += P) {
                const uint64_t ent = b[x];
                const int e = (int)(uint32_t)ent, val = (int)(ent >> 32);
                if ((e < low || e > high) && val > -1) {
                    const unsigned long long kk = ((unsigned long long)(uint32_t)val << 32) | (0xFFFFFFFFu - (unsigned)x);
                    if (kk > best) best = kk;
                }
            }
            unsigned hi = __reduce_max_sync(gmask, (unsigned)(best >> 32));
            unsigned lo = __reduce_max_sync(gmask, (unsigned)(best >> 32) == hi ? (unsigned)best : 0u);
            if (lo != 0) {                      // some entry qualified (index complement is never 0 here)
                const unsigned x = 0xFFFFFFFFu - lo;
                r.score2 = (int)hi; r.te2 = (int)(uint32_t)b[x];
            }
            __syncwarp(gmask);
        }
    }
    return r;
}

template <int P>
__global__ void __launch_bounds__(128)
ksw_align2_kernel(KswParams p, const bwagpu_seqpair *__restrict__ pairs, const uint8_t *__restrict__ ref,
                  const uint8_t *__restrict__ qer, bwagpu_kswr *__restrict__ aln, int n, int qpad,
                  uint64_t *__restrict__ bscratch, int max_tlen)
{
    extern __shared__ uint16_t ksw_sm[];
    const int gpb = blockDim.x / P, gi = threadIdx.x / P, l = threadIdx.x % P;
    const int gg = blockIdx.x * gpb + gi, tg = gridDim.x * gpb;
    const unsigned gmask = (P == 32 ? 0xffffffffu : ((1u << P) - 1u)) << ((threadIdx.x & 31) / P * P);
    uint16_t *sm = ksw_sm + (size_t)gi * (qpad * 4 + qpad / 2);
    uint64_t *b = bscratch + (size_t)gg * max_tlen;
    for (int job = gg; job < n; job += tg) {
        const bwagpu_seqpair sp = pairs[job];
        const int xtra = sp.h0;
        if (((xtra & KSW_XBYTE) != 0) != (P == 16)) continue;
        const uint8_t *q = qer + sp.idq, *t = ref + sp.idr;
        KswRes r = ksw_pass<P>(p, q, sp.len2, 0, t, sp.len1, 0, xtra, sm, qpad, b, gmask, l);
        int tb = -1, qb = -1;
        if ((xtra & KSW_XSTART) != 0 && !((xtra & KSW_XSUBO) && r.score < (xtra & 0xffff))) {
            const KswRes rr = ksw_pass<P>(p, q, r.qe + 1, r.qe + 1, t, sp.len1, r.te + 1, KSW_XSTOP | r.score, sm, qpad, b, gmask, l);
            if (r.score == rr.score) { tb = r.te - rr.te; qb = r.qe - rr.qe; }
        }
        if (l == 0) {
            bwagpu_kswr o;
            o.score = r.score; o.te = r.te; o.qe = r.qe; o.score2 = r.score2; o.te2 = r.te2; o.tb = tb; o.qb = qb;
            aln[job] = o;
        }
    }
}

__global__ void ksw_scan_kernel(const bwagpu_seqpair *__restrict__ pairs, int n, int *out)
{
    int m1 = 0, m2 = 0, n8 = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        m1 = max(m1, pairs[i].len1); m2 = max(m2, pairs[i].len2);
        n8 += (pairs[i].h0 & KSW_XBYTE) ? 1 : 0;
    }
    m1 = __reduce_max_sync(0xffffffffu, m1); m2 = __reduce_max_sync(0xffffffffu, m2);
    n8 = __reduce_add_sync(0xffffffffu, n8);
    if ((threadIdx.x & 31) == 0) { atomicMax(&out[0], m1); atomicMax(&out[1], m2); atomicAdd(&out[2], n8); }
}

static int ksw_params(bwagpu_ctx *ctx, const bwagpu_ksw_params *prm, KswParams *kp)
{
    if (!prm) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "ksw: params is NULL");
    if (prm->o_del < 0 || prm->e_del < 0 || prm->o_ins < 0 || prm->e_ins < 0 || prm->o_del + prm->e_del > 255 ||
        prm->o_ins + prm->e_ins > 255)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "ksw: gap penalties must be in [0,255]");
    int mn = 127, mx = 0;                         // ksw_qinit :79-85 (with its int8 casts)
    for (int a = 0; a < 25; ++a) {
        if (prm->mat[a] < (int8_t)mn) mn = prm->mat[a];
        if (prm->mat[a] > (int8_t)mx) mx = prm->mat[a];
    }
    if (mx <= 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "ksw: the score matrix has no positive entry");
    kp->shift = (uint8_t)(256 - mn); kp->qmax = (uint8_t)mx;
    for (int t = 0; t < 5; ++t) {
        uint32_t lo = 0;
        for (int c = 0; c < 4; ++c) lo |= (uint32_t)(uint8_t)prm->mat[t * 5 + c] << (8 * c);
        kp->row_lo[t] = lo; kp->row_hi[t] = (uint32_t)(uint8_t)prm->mat[t * 5 + 4];
    }
    kp->o_del = prm->o_del; kp->e_del = prm->e_del; kp->o_ins = prm->o_ins; kp->e_ins = prm->e_ins;
    return BWAGPU_OK;
}

template <int P>
static int ksw_launch(bwagpu_ctx *ctx, const KswParams &kp, const bwagpu_seqpair *d_pairs, const uint8_t *d_ref,
                      const uint8_t *d_qer, bwagpu_kswr *d_aln, int n, int max_qlen, int max_tlen)
{
    const int qpad = ((max_qlen + 15) / 16) * 16 + 16;
    const size_t per_group = (size_t)(qpad * 4 + qpad / 2) * 2;
    int tpb = 128;
    while (tpb > 32 && per_group * (tpb / P) > 160 * 1024) tpb >>= 1;
    const size_t smem = per_group * (tpb / P);
    if (smem > 200 * 1024) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "ksw: query of %d bases does not fit shared memory", max_qlen);
    static BwagpuSmemBound smem_bound;
    BWAGPU_CUDA(ctx, bwagpu_raise_dyn_smem(smem_bound, ctx->device, ksw_align2_kernel<P>, (int)smem));
    int per_sm = 1;
    BWAGPU_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ksw_align2_kernel<P>, tpb, smem));
    if (per_sm < 1) per_sm = 1;
    const int gpb = tpb / P;
    int blocks = ctx->sm_count * per_sm;
    const int need = (n + gpb - 1) / gpb;
    if (blocks > need) blocks = need;
    uint64_t *bs;
    int rc = bwagpu_scratch(ctx, P == 16 ? SCR_MISC1 : SCR_MISC2, (int64_t)blocks * gpb * (max_tlen + 1) * 8, (void **)&bs);
    if (rc) return rc;
    BWAGPU_KERNEL_BEGIN(ctx);
    ksw_align2_kernel<P><<<blocks, tpb, smem, ctx->stream>>>(kp, d_pairs, d_ref, d_qer, d_aln, n, qpad, bs, max_tlen + 1);
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_KSW);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    return BWAGPU_OK;
}

static int ksw_run(bwagpu_ctx *ctx, const KswParams &kp, const bwagpu_seqpair *d_pairs, const uint8_t *d_ref,
                   const uint8_t *d_qer, bwagpu_kswr *d_aln, int n, int max_tlen, int max_qlen, int n8)
{
    if (max_qlen > KSW_MAX_QLEN)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "ksw: queries longer than %d are not supported (got %d)", KSW_MAX_QLEN, max_qlen);
    int rc;
    if (n8 > 0 && (rc = ksw_launch<16>(ctx, kp, d_pairs, d_ref, d_qer, d_aln, n, max_qlen, max_tlen))) return rc;
    if (n8 < n && (rc = ksw_launch<8>(ctx, kp, d_pairs, d_ref, d_qer, d_aln, n, max_qlen, max_tlen))) return rc;
    return BWAGPU_OK;
}

extern "C" int bwagpu_ksw_align2_batch_device(bwagpu_ctx *ctx, const bwagpu_ksw_params *prm,
                                              const bwagpu_seqpair *d_pairArray, const uint8_t *d_seqBufRef,
                                              const uint8_t *d_seqBufQer, bwagpu_kswr *d_aln, int32_t numPairs)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    KswParams kp;
    int rc = ksw_params(ctx, prm, &kp);
    if (rc) return rc;
    if (numPairs < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "ksw: negative numPairs");
    if (numPairs == 0) return BWAGPU_OK;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    int *d_m, h_m[4];
    if ((rc = bwagpu_scratch(ctx, SCR_MISC0, 64, (void **)&d_m))) return rc;
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_m, 0, 16, ctx->stream));
    ksw_scan_kernel<<<ctx->sm_count * 2, 256, 0, ctx->stream>>>(d_pairArray, numPairs, d_m);
    ctx->launches++;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(h_m, d_m, 16, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return ksw_run(ctx, kp, d_pairArray, d_seqBufRef, d_seqBufQer, d_aln, numPairs, h_m[0], h_m[1], h_m[2]);
}

extern "C" int bwagpu_ksw_align2_batch(bwagpu_ctx *ctx, const bwagpu_ksw_params *prm, const bwagpu_seqpair *pairArray,
                                       const uint8_t *seqBufRef, const uint8_t *seqBufQer, bwagpu_kswr *aln,
                                       int32_t numPairs)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    KswParams kp;
    int rc = ksw_params(ctx, prm, &kp);
    if (rc) return rc;
    if (numPairs < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "ksw: negative numPairs");
    if (numPairs == 0) return BWAGPU_OK;
    if (!pairArray || !seqBufRef || !seqBufQer || !aln) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "ksw: NULL buffer");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    int64_t ref_bytes = 0, qer_bytes = 0;
    int max_t = 0, max_q = 0, n8 = 0;
    for (int32_t i = 0; i < numPairs; ++i) {
        const bwagpu_seqpair &sp = pairArray[i];
        if (sp.len1 < 0 || sp.len2 < 0 || sp.idr < 0 || sp.idq < 0)
            BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "ksw: pair %d has a negative length/offset", i);
        const int64_t r = (int64_t)sp.idr + sp.len1, q = (int64_t)sp.idq + sp.len2;
        if (r > ref_bytes) ref_bytes = r;
        if (q > qer_bytes) qer_bytes = q;
        if (sp.len1 > max_t) max_t = sp.len1;
        if (sp.len2 > max_q) max_q = sp.len2;
        n8 += (sp.h0 & KSW_XBYTE) ? 1 : 0;
    }
    bwagpu_seqpair *d_pairs; uint8_t *d_ref, *d_qer; bwagpu_kswr *d_aln;
    if ((rc = bwagpu_scratch(ctx, SCR_BSW_PAIRS, (int64_t)numPairs * sizeof(bwagpu_seqpair), (void **)&d_pairs))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_BSW_REF, ref_bytes + 16, (void **)&d_ref))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_BSW_QER, qer_bytes + 16, (void **)&d_qer))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_MISC3, (int64_t)numPairs * sizeof(bwagpu_kswr), (void **)&d_aln))) return rc;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_pairs, pairArray, (size_t)numPairs * sizeof(bwagpu_seqpair), cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_ref, seqBufRef, (size_t)ref_bytes, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_qer, seqBufQer, (size_t)qer_bytes, cudaMemcpyHostToDevice, ctx->stream));
    if ((rc = ksw_run(ctx, kp, d_pairs, d_ref, d_qer, d_aln, numPairs, max_t, max_q, n8))) return rc;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(aln, d_aln, (size_t)numPairs * sizeof(bwagpu_kswr), cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}
