// CIGAR stage of the C ABI: bwagpu_reg2aln_batch (include/bwagpu.h, seam S4).  Replaces the alignment work of
// mem_reg2aln (reference src/bwamem.cpp:2533-2600); the per-region algorithm lives in cigar_core.cuh.
// One region per thread: most regions take the gap-free shortcut (a 151-step sum); the rest run a banded global
// alignment whose backtrack matrix (one byte per cell) lives in a per-job slab of global memory.  Jobs are
// processed in chunks so that the slabs stay below a fixed budget.  Results are compacted by a scan + copy kernel.
#include <cub/device/device_scan.cuh>
#include <cub/device/device_radix_sort.cuh>
#include <stdlib.h>
#include <algorithm>
#include "ctx.cuh"
#include "cigar_core.cuh"

struct CigSlabOff { int64_t eh, z, cigar, md; };

// sort key of a job: 0 = gap-free; else (tier + 1) in the top bits, then the size of its backtrack matrix
__host__ __device__ inline uint32_t cig_key(bool gapfree, int tier, int64_t zb)
{
    return gapfree ? 0u : ((uint32_t)(tier + 1) << 28) | (uint32_t)(zb < (1 << 28) - 1 ? zb : (1 << 28) - 1);
}
__host__ __device__ inline int cig_key_tier(uint32_t key) { return (int)(key >> 28) - 1; }      // -1 = gap-free

#define CIG_TPB 64
#define CIG_SMEM_QLEN 151                      // H/E rows of reads up to this length live in shared memory

// thread t of the launch handles job order[t] (jobs sorted by the host: gap-free ones first, then by DP size, so that
// the lanes of a warp run similar trip counts); results land at the job's own index
__global__ void __launch_bounds__(CIG_TPB)
cigar_kernel(CigarOpt o, int64_t l_pac, const uint8_t *__restrict__ ref_string, const uint8_t *__restrict__ reads,
             const int64_t *__restrict__ off, const int32_t *__restrict__ len, const bwagpu_cigar_job *__restrict__ jobs,
             const int32_t *__restrict__ order, int64_t first, int64_t n, int tier, const CigSlabOff *__restrict__ so, int32_t *eh, uint8_t *z, uint32_t *cigar,
             char *md, bwagpu_cigar_out *outs, int64_t *cnt_cigar, int64_t *cnt_md, int32_t *redo, unsigned int *n_redo)
{
    extern __shared__ int32_t cig_smem[];
    const int64_t t = first + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int64_t j = order[t];
    const bwagpu_cigar_job jb = jobs[j];
    CigarScratch sc;
    // tier -1: gap-free group (no rows); 0-3: circular window of 16 << tier columns; 4: full rows
    if (jb.qe - jb.qb <= CIG_SMEM_QLEN) { sc.eh = cig_smem + threadIdx.x; sc.eh_stride = CIG_TPB; sc.eh_mask = tier >= 0 && tier < 4 ? (16 << tier) - 1 : 0x7fffffff; }
    else { sc.eh = eh + so[j].eh; sc.eh_stride = 1; sc.eh_mask = 0x7fffffff; }
    sc.z = z + so[j].z; sc.cigar = cigar + so[j].cigar; sc.md = md + so[j].md;
    CigarResult r;
    r.score = 0;
    bwagpu_cigar_out out;
    out.pos = -1; out.cigar_off = out.md_off = 0; out.is_rev = 0; out.n_cigar = 0; out.nm = -1; out.l_md = 0; out.score = 0;
    const int rcode = cig_reg2aln(o, l_pac, ref_string, reads + off[jb.read], jb.qb, jb.qe, jb.rb, jb.re, jb.truesc, jb.w, sc, &r);
    if (rcode == 2) {                       // a retry's band outgrew this tier's window: second pass with full rows
        if (redo) redo[atomicAdd(n_redo, 1u)] = (int32_t)j;
        return;
    }
    out.ok = rcode == 1 ? 1 : 0;
    if (out.ok) {
        int is_rev;
        out.n_cigar = cig_finish(l_pac, jb.rb, jb.re, jb.qb, jb.qe, len[jb.read], sc.cigar, r.n_cigar, &out.pos, &is_rev);
        out.is_rev = is_rev; out.nm = r.nm; out.l_md = r.l_md; out.score = r.score;
    }
    outs[j] = out;
    cnt_cigar[j] = out.n_cigar;
    cnt_md[j] = out.ok ? out.l_md + 1 : 0;
}

__global__ void cigar_emit_kernel(int64_t n, const CigSlabOff *__restrict__ so, const uint32_t *__restrict__ cigar, const char *__restrict__ md,
                                  const int64_t *__restrict__ coff, const int64_t *__restrict__ moff, int64_t cbase, int64_t mbase,
                                  bwagpu_cigar_out *outs, uint32_t *out_cigar, char *out_md)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    bwagpu_cigar_out o = outs[j];
    o.cigar_off = cbase + coff[j]; o.md_off = mbase + moff[j];
    outs[j] = o;
    for (int k = 0; k < o.n_cigar; ++k) out_cigar[coff[j] + k] = cigar[so[j].cigar + k];
    if (o.ok) {
        for (int k = 0; k < o.l_md; ++k) out_md[moff[j] + k] = md[so[j].md + k];
        out_md[moff[j] + o.l_md] = 0;
    }
}

static int cig_scan(bwagpu_ctx *ctx, const int64_t *in, int64_t *out, int64_t n)
{
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum((void *)nullptr, bytes, in, out, (int)n, ctx->stream);
    void *tmp;
    int rc = bwagpu_scratch(ctx, SCR_CUB, (int64_t)bytes + 256, &tmp);
    if (rc) return rc;
    BWAGPU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, bytes, in, out, (int)n, ctx->stream));
    ctx->launches += 2;
    return BWAGPU_OK;
}


// launches the groups of a sorted job list: bounds[g] = first position of group g (0 = gap-free, 1..5 = tiers 0..4), bounds[6] = n
static int cig_launch_groups(bwagpu_ctx *ctx, const CigarOpt &o, const uint8_t *d_reads, const int64_t *d_off, const int32_t *d_len,
                             const bwagpu_cigar_job *d_jobs, const int32_t *d_order, const int64_t *bounds, const CigSlabOff *d_so,
                             int32_t *d_eh, uint8_t *d_z, uint32_t *d_cig, char *d_md, bwagpu_cigar_out *d_out, int64_t *d_cnt, int64_t n)
{
    BWAGPU_CUDA(ctx, cudaFuncSetAttribute(cigar_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)((size_t)2 * (CIG_SMEM_QLEN + 1) * CIG_TPB * 4)));
    int32_t *d_redo; unsigned int *d_nredo; void *stats;
    int rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CIG_REDO, n * 4 + 16, (void **)&d_redo))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_STATS, 256, &stats))) return rc;
    d_nredo = (unsigned int *)((char *)stats + 128);
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_nredo, 0, 4, ctx->stream));
    int launched = 0;
    for (int g = 0; g < 6; ++g) {
        const int64_t b0 = bounds[g], b1 = bounds[g + 1];
        if (b1 <= b0) continue;
        const int tier = g - 1;
        const size_t smem = tier < 0 ? 0 : tier < 4 ? (size_t)2 * (16 << tier) * CIG_TPB * 4 : (size_t)2 * (CIG_SMEM_QLEN + 1) * CIG_TPB * 4;
        cigar_kernel<<<(unsigned)((b1 - b0 + CIG_TPB - 1) / CIG_TPB), CIG_TPB, smem, ctx->stream>>>(
            o, ctx->l_pac, (const uint8_t *)ctx->d_ref_string, d_reads, d_off, d_len, d_jobs, d_order, b0, b1, tier, d_so, d_eh, d_z, d_cig,
            d_md, d_out, d_cnt, d_cnt + n + 1, d_redo, d_nredo);
        ++launched;
    }
    unsigned int n_redo = 0;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&n_redo, d_nredo, 4, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    if (n_redo > 0) {
        cigar_kernel<<<(unsigned)((n_redo + CIG_TPB - 1) / CIG_TPB), CIG_TPB, (size_t)2 * (CIG_SMEM_QLEN + 1) * CIG_TPB * 4, ctx->stream>>>(
            o, ctx->l_pac, (const uint8_t *)ctx->d_ref_string, d_reads, d_off, d_len, d_jobs, d_redo, 0, (int64_t)n_redo, 4, d_so, d_eh, d_z,
            d_cig, d_md, d_out, d_cnt, d_cnt + n + 1, nullptr, nullptr);
        ++launched;
    }
    ctx->launches += launched > 1 ? launched - 1 : 0;
    BWAGPU_CUDA(ctx, cudaGetLastError());
    return BWAGPU_OK;
}

// ---- device-resident form ----------------------------------------------------------------------------------
// per job: slab sizes (eh words, z bytes, cigar words, md bytes) and the sort key (0 = gap-free, else 1 + z bytes)
__global__ void cigar_plan_kernel(CigarOpt o, const bwagpu_cigar_job *__restrict__ jobs, int64_t n, const int32_t *__restrict__ len,
                                  int32_t nseq, int64_t *sizes /* 4 x (n + 1) */, uint32_t *keys, int32_t *idx, int32_t *bad)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const bwagpu_cigar_job b = jobs[j];
    if (b.read < 0 || b.read >= nseq || b.qb < 0 || b.qb > b.qe || b.qe > len[b.read] || b.re - b.rb > (1 << 20)) { atomicExch(bad, 1); }
    const int64_t lq = b.qe > b.qb ? b.qe - b.qb : 0, lt = b.re > b.rb ? b.re - b.rb : 0;
    const bool gapfree = cig_is_gapfree(o, b.qb, b.qe, b.rb, b.re, b.truesc, b.w);
    const int64_t zb = gapfree ? 0 : cig_z_bytes(o, b.qb, b.qe, b.rb, b.re, b.truesc, b.w);
    sizes[j] = lq > CIG_SMEM_QLEN ? 2 * (lq + 1) : 0;
    sizes[(n + 1) + j] = zb;
    sizes[2 * (n + 1) + j] = lq + lt + 4;
    sizes[3 * (n + 1) + j] = 3 * lt + 16;
    keys[j] = cig_key(gapfree, lq > CIG_SMEM_QLEN ? 4 : cig_tier(o, b.qb, b.qe, b.rb, b.re, b.truesc, b.w), zb);
    idx[j] = (int32_t)j;
}

__global__ void cigar_slab_kernel(int64_t n, const int64_t *__restrict__ scans, const uint32_t *__restrict__ sorted_keys, CigSlabOff *so,
                                  int64_t *first /* [6]: first sorted position of each group, preset to n */)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    CigSlabOff s = { scans[j], scans[(n + 1) + j], scans[2 * (n + 1) + j], scans[3 * (n + 1) + j] };
    so[j] = s;
    const int g = cig_key_tier(sorted_keys[j]) + 1;
    if (j == 0 || cig_key_tier(sorted_keys[j - 1]) + 1 != g) first[g] = j;
}

// group bounds from the first positions of the groups present (absent groups are empty: they start where the next one does)
static void cig_bounds(const int64_t *first, int64_t n, int64_t *bounds)
{
    bounds[6] = n;
    for (int g = 5; g >= 0; --g) bounds[g] = first[g] < bounds[g + 1] ? first[g] : bounds[g + 1];
}

extern "C" int bwagpu_reg2aln_device(bwagpu_ctx *ctx, const bwagpu_cigar_opt *opt, const uint8_t *d_reads, const int64_t *d_off,
                                     const int32_t *d_len, int32_t nseq, const bwagpu_cigar_job *d_jobs, int64_t n_jobs,
                                     bwagpu_cigar_result *res)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!opt || !res || nseq < 0 || n_jobs < 0 || n_jobs > 0x7fffffff) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "reg2aln_device: bad argument");
    if (!ctx->d_ref_string) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "reg2aln_device: no reference loaded (bwagpu_index_load_ref)");
    if (opt->e_del <= 0 || opt->e_ins <= 0 || opt->w < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "reg2aln_device: gap extension penalties must be positive");
    memset(res, 0, sizeof *res);
    if (n_jobs == 0) return BWAGPU_OK;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    CigarOpt o;
    memcpy(o.mat, opt->mat, 25);
    o.a = opt->a; o.o_del = opt->o_del; o.e_del = opt->e_del; o.o_ins = opt->o_ins; o.e_ins = opt->e_ins; o.w = opt->w;
    const int64_t n = n_jobs;
    int rc;
    int64_t *d_sizes, *d_scans, *d_cnt, *d_scan; uint32_t *d_keys, *d_keys2; int32_t *d_idx, *d_order, *d_flag; CigSlabOff *d_so;
    int64_t h_ngf = 0; (void)h_ngf;
    bwagpu_cigar_out *d_out;
    if ((rc = bwagpu_scratch(ctx, SCR_CIG_JOBS, 4 * (n + 1) * 8 + 64, (void **)&d_sizes))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CIG_EH, 4 * (n + 1) * 8 + 64, (void **)&d_scans))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_MISC0, n * 8 + 64, (void **)&d_keys))) return rc;
    d_keys2 = d_keys + n;
    if ((rc = bwagpu_scratch(ctx, SCR_MISC1, n * 4 + 64, (void **)&d_idx))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CIG_ORDER, n * 4 + 64, (void **)&d_order))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CIG_SLABOFF, n * (int64_t)sizeof(CigSlabOff), (void **)&d_so))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CIG_OUT, n * (int64_t)sizeof(bwagpu_cigar_out), (void **)&d_out))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CIG_CNT, (n + 1) * 16, (void **)&d_cnt))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CIG_SCAN, (n + 1) * 16 + 64, (void **)&d_scan))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_STATS, 256, (void **)&d_flag))) return rc;
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_flag, 0, 16, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_sizes, 0, (size_t)4 * (n + 1) * 8, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_cnt, 0, (size_t)(n + 1) * 16, ctx->stream));
    cigar_plan_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(o, d_jobs, n, d_len, nseq, d_sizes, d_keys, d_idx, d_flag);
    ctx->launches++;
    for (int a = 0; a < 4; ++a)
        if ((rc = cig_scan(ctx, d_sizes + a * (n + 1), d_scans + a * (n + 1), n + 1))) return rc;
    size_t sb = 0;
    cub::DeviceRadixSort::SortPairs((void *)nullptr, sb, d_keys, d_keys2, d_idx, d_order, (int)n, 0, 32, ctx->stream);
    void *tmp;
    if ((rc = bwagpu_scratch(ctx, SCR_CUB, (int64_t)sb + 256, &tmp))) return rc;
    BWAGPU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, sb, d_keys, d_keys2, d_idx, d_order, (int)n, 0, 32, ctx->stream));
    ctx->launches += 4;
    int64_t *d_first = (int64_t *)(d_flag + 2);
    int64_t h_first[6] = {n, n, n, n, n, n};
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_first, h_first, sizeof h_first, cudaMemcpyHostToDevice, ctx->stream));
    cigar_slab_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(n, d_scans, d_keys2, d_so, d_first);
    ctx->launches++;
    int64_t tot[4]; int32_t h_bad = 0;
    for (int a = 0; a < 4; ++a) BWAGPU_CUDA(ctx, cudaMemcpyAsync(&tot[a], d_scans + a * (n + 1) + n, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(h_first, d_first, sizeof h_first, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&h_bad, d_flag, 4, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    if (h_bad) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "reg2aln_device: a job is out of range (read index, query span or window)");
    if (tot[1] + (tot[0] + tot[2]) * 4 + tot[3] > (32LL << 30))
        BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "reg2aln_device: %lld jobs need %.1f GB of scratch; pass smaller batches", (long long)n, (tot[1] + (tot[0] + tot[2]) * 4 + tot[3]) / 1e9);
    int32_t *d_eh; uint8_t *d_z; uint32_t *d_cig; char *d_md;
    if ((rc = bwagpu_scratch(ctx, SCR_MISC2, tot[0] * 4 + 16, (void **)&d_eh))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CIG_Z, tot[1] + 16, (void **)&d_z))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CIG_CIGAR, tot[2] * 4 + 16, (void **)&d_cig))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CIG_MD, tot[3] + 16, (void **)&d_md))) return rc;
    int64_t bounds[7];
    cig_bounds(h_first, n, bounds);
    const int64_t n_gf = bounds[1];
    BWAGPU_KERNEL_BEGIN(ctx);
    if ((rc = cig_launch_groups(ctx, o, d_reads, d_off, d_len, d_jobs, d_order, bounds, d_so, d_eh, d_z, d_cig, d_md, d_out, d_cnt, n))) return rc;
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_CIGAR);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    if ((rc = cig_scan(ctx, d_cnt, d_scan, n + 1))) return rc;
    if ((rc = cig_scan(ctx, d_cnt + n + 1, d_scan + n + 1, n + 1))) return rc;
    int64_t t2[2] = {0, 0};
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&t2[0], d_scan + n, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&t2[1], d_scan + 2 * n + 1, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    uint32_t *d_oc; char *d_om;
    if ((rc = bwagpu_scratch(ctx, SCR_CIG_OUTC, t2[0] * 4 + 16, (void **)&d_oc))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_CIG_OUTM, t2[1] + 16, (void **)&d_om))) return rc;
    BWAGPU_KERNEL_BEGIN(ctx);
    cigar_emit_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(n, d_so, d_cig, d_md, d_scan, d_scan + n + 1, 0, 0, d_out, d_oc, d_om);
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_CIGAR);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    res->d_outs = d_out; res->n_jobs = n; res->d_cigar = d_oc; res->n_cigar = t2[0]; res->d_md = d_om; res->md_bytes = t2[1];
    res->n_gapfree = n_gf;
    return BWAGPU_OK;
}

extern "C" int bwagpu_reg2aln_batch(bwagpu_ctx *ctx, const bwagpu_cigar_opt *opt, const uint8_t *enc_reads, const int64_t *off,
                                    const int32_t *len, int32_t nseq, const bwagpu_cigar_job *jobs, int64_t n_jobs,
                                    bwagpu_cigar_out *outs, uint32_t *cigar, int64_t cigar_cap, char *md, int64_t md_cap,
                                    int64_t *n_cigar_total, int64_t *md_bytes)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!opt || !n_cigar_total || !md_bytes || nseq < 0 || n_jobs < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "reg2aln_batch: bad argument");
    if (!ctx->d_ref_string) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "reg2aln_batch: no reference loaded (bwagpu_index_load_ref)");
    if (opt->e_del <= 0 || opt->e_ins <= 0 || opt->w < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "reg2aln_batch: gap extension penalties must be positive");
    *n_cigar_total = 0; *md_bytes = 0;
    if (n_jobs == 0) return BWAGPU_OK;
    if (!enc_reads || !off || !len || !jobs || !outs) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "reg2aln_batch: NULL buffer");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    int64_t bytes = 0;
    for (int i = 0; i < nseq; ++i) {
        if (len[i] < 0 || off[i] < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "reg2aln_batch: read %d has a negative length/offset", i);
        if (off[i] + len[i] > bytes) bytes = off[i] + len[i];
    }
    for (int64_t j = 0; j < n_jobs; ++j) {
        const bwagpu_cigar_job &b = jobs[j];
        if (b.read < 0 || b.read >= nseq || b.qb < 0 || b.qe > len[b.read] || b.qb > b.qe || b.re - b.rb > (1 << 20))
            BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "reg2aln_batch: job %lld is out of range", (long long)j);
    }
    CigarOpt o;
    memcpy(o.mat, opt->mat, 25);
    o.a = opt->a; o.o_del = opt->o_del; o.e_del = opt->e_del; o.o_ins = opt->o_ins; o.e_ins = opt->e_ins; o.w = opt->w;
    uint8_t *d_reads; int64_t *d_off; int32_t *d_len;
    int rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_READS, bytes + 16, (void **)&d_reads))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_OFF, (int64_t)nseq * 8, (void **)&d_off))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_LEN, (int64_t)nseq * 4, (void **)&d_len))) return rc;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_reads, enc_reads, (size_t)bytes, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_off, off, (size_t)nseq * 8, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_len, len, (size_t)nseq * 4, cudaMemcpyHostToDevice, ctx->stream));
    const int64_t budget = 4LL << 30;                         // slab bytes per chunk
    CigSlabOff *h_so = (CigSlabOff *)malloc(sizeof(CigSlabOff) * (size_t)n_jobs);
    uint64_t *h_key = (uint64_t *)malloc(8 * (size_t)n_jobs);
    int32_t *h_order = (int32_t *)malloc(4 * (size_t)n_jobs);
    if (!h_so || !h_key || !h_order) { free(h_so); free(h_key); free(h_order); BWAGPU_FAIL(ctx, BWAGPU_ERR_NOMEM, "reg2aln_batch: host allocation failed"); }
    int64_t cbase = 0, mbase = 0;
    bool fits = true;
    for (int64_t j0 = 0; j0 < n_jobs;) {
        CigSlabOff cur = {0, 0, 0, 0};
        int64_t j1 = j0;
        while (j1 < n_jobs) {
            const bwagpu_cigar_job &b = jobs[j1];
            const int64_t lq = b.qe - b.qb, lt = b.re > b.rb ? b.re - b.rb : 0;
            const bool gapfree = cig_is_gapfree(o, b.qb, b.qe, b.rb, b.re, b.truesc, b.w);
            const int64_t zb = gapfree ? 0 : cig_z_bytes(o, b.qb, b.qe, b.rb, b.re, b.truesc, b.w);
            if (j1 > j0 && cur.z + zb + (cur.eh + cur.cigar) * 4 + cur.md > budget) break;
            h_so[j1] = cur;
            if (lq > CIG_SMEM_QLEN) cur.eh += 2 * (lq + 1);
            cur.z += zb; cur.cigar += lq + lt + 4; cur.md += 3 * lt + 16;
            // sort key: gap-free jobs first, the others by the size of their backtrack matrix
            h_key[j1 - j0] = ((uint64_t)cig_key(gapfree, lq > CIG_SMEM_QLEN ? 4 : cig_tier(o, b.qb, b.qe, b.rb, b.re, b.truesc, b.w), zb) << 32) | (uint32_t)(j1 - j0);
            ++j1;
        }
        const int64_t n = j1 - j0;
        bwagpu_cigar_job *d_jobs; bwagpu_cigar_out *d_out; CigSlabOff *d_so; int32_t *d_eh; uint8_t *d_z; uint32_t *d_cig; char *d_md;
        int64_t *d_cnt, *d_scan; uint32_t *d_oc; char *d_om;
        #define CIG_TRY(x) do { if ((rc = (x))) { free(h_so); free(h_key); free(h_order); return rc; } } while (0)
        CIG_TRY(bwagpu_scratch(ctx, SCR_CIG_JOBS, n * (int64_t)sizeof(bwagpu_cigar_job), (void **)&d_jobs));
        CIG_TRY(bwagpu_scratch(ctx, SCR_CIG_OUT, n * (int64_t)sizeof(bwagpu_cigar_out), (void **)&d_out));
        CIG_TRY(bwagpu_scratch(ctx, SCR_CIG_SLABOFF, n * (int64_t)sizeof(CigSlabOff), (void **)&d_so));
        CIG_TRY(bwagpu_scratch(ctx, SCR_CIG_EH, cur.eh * 4 + 16, (void **)&d_eh));
        CIG_TRY(bwagpu_scratch(ctx, SCR_CIG_Z, cur.z + 16, (void **)&d_z));
        CIG_TRY(bwagpu_scratch(ctx, SCR_CIG_CIGAR, cur.cigar * 4 + 16, (void **)&d_cig));
        CIG_TRY(bwagpu_scratch(ctx, SCR_CIG_MD, cur.md + 16, (void **)&d_md));
        CIG_TRY(bwagpu_scratch(ctx, SCR_CIG_CNT, (n + 1) * 16, (void **)&d_cnt));
        CIG_TRY(bwagpu_scratch(ctx, SCR_CIG_SCAN, (n + 1) * 16, (void **)&d_scan));
        std::sort(h_key, h_key + n);
        for (int64_t k = 0; k < n; ++k) h_order[k] = (int32_t)(h_key[k] & 0xFFFFFFFFu);
        int32_t *d_order;
        CIG_TRY(bwagpu_scratch(ctx, SCR_CIG_ORDER, n * 4, (void **)&d_order));
        cudaError_t e = cudaMemcpyAsync(d_order, h_order, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(d_jobs, jobs + j0, (size_t)n * sizeof(bwagpu_cigar_job), cudaMemcpyHostToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(d_so, h_so + j0, (size_t)n * sizeof(CigSlabOff), cudaMemcpyHostToDevice, ctx->stream);
        if (e == cudaSuccess) e = cudaMemsetAsync(d_cnt, 0, (size_t)(n + 1) * 16, ctx->stream);
        if (e != cudaSuccess) { free(h_so); free(h_key); free(h_order); BWAGPU_FAIL(ctx, BWAGPU_ERR_CUDA, "reg2aln_batch: upload -> %s", cudaGetErrorString(e)); }
        // groups of the sorted list: gap-free jobs (no DP rows, full occupancy), then the window tiers
        int64_t first[6] = {n, n, n, n, n, n}, bounds[7];
        for (int64_t k = n - 1; k >= 0; --k) first[cig_key_tier((uint32_t)(h_key[k] >> 32)) + 1] = k;
        cig_bounds(first, n, bounds);
        BWAGPU_KERNEL_BEGIN(ctx);
        CIG_TRY(cig_launch_groups(ctx, o, d_reads, d_off, d_len, d_jobs, d_order, bounds, d_so, d_eh, d_z, d_cig, d_md, d_out, d_cnt, n));
        BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_CIGAR);
        CIG_TRY(cig_scan(ctx, d_cnt, d_scan, n + 1));
        CIG_TRY(cig_scan(ctx, d_cnt + n + 1, d_scan + n + 1, n + 1));
        int64_t tot[2] = {0, 0};
        e = cudaMemcpyAsync(&tot[0], d_scan + n, 8, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(&tot[1], d_scan + 2 * n + 1, 8, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = bwagpu_stream_sync(ctx);
        if (e != cudaSuccess) { free(h_so); free(h_key); free(h_order); BWAGPU_FAIL(ctx, BWAGPU_ERR_CUDA, "reg2aln_batch: kernel -> %s", cudaGetErrorString(e)); }
        CIG_TRY(bwagpu_scratch(ctx, SCR_CIG_OUTC, tot[0] * 4 + 16, (void **)&d_oc));
        CIG_TRY(bwagpu_scratch(ctx, SCR_CIG_OUTM, tot[1] + 16, (void **)&d_om));
        BWAGPU_KERNEL_BEGIN(ctx);
        cigar_emit_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(n, d_so, d_cig, d_md, d_scan, d_scan + n + 1, cbase, mbase,
                                                                               d_out, d_oc, d_om);
        BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_CIGAR);
        e = cudaMemcpyAsync(outs + j0, d_out, (size_t)n * sizeof(bwagpu_cigar_out), cudaMemcpyDeviceToHost, ctx->stream);
        if (cbase + tot[0] > cigar_cap || mbase + tot[1] > md_cap) fits = false;
        if (fits && e == cudaSuccess && tot[0] > 0) e = cudaMemcpyAsync(cigar + cbase, d_oc, (size_t)tot[0] * 4, cudaMemcpyDeviceToHost, ctx->stream);
        if (fits && e == cudaSuccess && tot[1] > 0) e = cudaMemcpyAsync(md + mbase, d_om, (size_t)tot[1], cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = bwagpu_stream_sync(ctx);
        if (e != cudaSuccess) { free(h_so); free(h_key); free(h_order); BWAGPU_FAIL(ctx, BWAGPU_ERR_CUDA, "reg2aln_batch: copy back -> %s", cudaGetErrorString(e)); }
        cbase += tot[0]; mbase += tot[1];
        j0 = j1;
    }
    free(h_so); free(h_key); free(h_order);
    *n_cigar_total = cbase; *md_bytes = mbase;
    if (!fits) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "reg2aln_batch: %lld cigar ops / %lld MD bytes do not fit the output buffers", (long long)cbase, (long long)mbase);
    return BWAGPU_OK;
}
