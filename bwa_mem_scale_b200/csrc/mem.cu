// Worker-level entry of the C ABI: bwagpu_mem_process_chunk{,_device}, bwagpu_mem_fetch  (include/bwagpu.h).
// One call does, for a whole chunk of reads, what the reference's two kt_for passes of mem_process_seqs do
// (reference src/bwamem.cpp:1872-1876): worker_bwt -> mem_kernel1_core[_ert] (:1209-1363, :1056-1207) and
// worker_aln -> mem_kernel2_core (:1365-1487) up to the per-read compaction of the alignment regions
// (:1449-1460).  The stages are the existing seams run back to back on the context's stream without leaving HBM:
//   EMF probe -> SMEM / ERT seeding (+ SA) -> chaining (mem_chain_seeds / mem_chain_new + mem_chain_flt)
//   -> mem_chain2aln_across_reads_V2 -> drop regions with qe <= qb, per-read offsets.
// mem_sort_dedup_patch / is_alt (:1462-1483) stay the host's (they need bns and run on the compacted lists).
#include <cub/device/device_scan.cuh>
#include "ctx.cuh"

// keep[i] = region i survives the reference's compaction loop (a[i].qe > a[i].qb, bwamem.cpp:1455)
__global__ void mem_keep_kernel(const bwagpu_alnreg *__restrict__ regs, int64_t n, int64_t *__restrict__ keep)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    keep[i] = i < n && regs[i].qe > regs[i].qb ? 1 : 0;
}

__global__ void mem_compact_kernel(const bwagpu_alnreg *__restrict__ regs, int64_t n, const int64_t *__restrict__ pos,
                                   bwagpu_alnreg *__restrict__ out)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (pos[i + 1] != pos[i]) out[pos[i]] = regs[i];
}

// reg_off[r] = first compacted region of read r.  Raw regions are laid out chain by chain in read order
// (regs[c.seed_off + k]); first_chain[r] is the first chain of read r or -1 (chain2aln's table).
__global__ void mem_regoff_kernel(const bwagpu_chain *__restrict__ chains, int64_t n_chains, int64_t n_raw,
                                  const int64_t *__restrict__ first_chain, int32_t nseq, const int64_t *__restrict__ pos,
                                  int64_t *__restrict__ reg_off)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r > nseq) return;
    if (r == nseq) { reg_off[r] = pos[n_raw]; return; }
    int64_t c = first_chain[r];
    if (c < 0) {                        // no chains: the read's (empty) range starts where the next read with chains starts
        int64_t lo = 0, hi = n_chains;  // first chain with seqid > r
        while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (chains[mid].seqid > r) hi = mid; else lo = mid + 1; }
        c = lo;
    }
    reg_off[r] = pos[c < n_chains ? chains[c].seed_off : n_raw];
}

__global__ void mem_zero_off_kernel(int64_t *reg_off, int32_t nseq)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r <= nseq) reg_off[r] = 0;
}

__global__ void mem_count_skip_kernel(const uint8_t *__restrict__ skip, int32_t nseq, unsigned long long *out)
{
    unsigned long long c = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nseq; i += (int64_t)gridDim.x * blockDim.x) c += skip[i] != 0;
    for (int o = 16; o; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

static int mem_check(bwagpu_ctx *ctx, const bwagpu_mem_opt *opt, bwagpu_mem_result *res, int32_t nseq)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!opt || !res) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "mem_process_chunk: options / result are NULL");
    if (nseq < 0 || opt->batch_size < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "mem_process_chunk: negative nseq / batch_size");
    if (!ctx->d_ref_string) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "mem_process_chunk: no reference loaded (bwagpu_index_load_ref)");
    if (opt->use_ert ? !ctx->d_ert_kmer : !ctx->has_fm)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "mem_process_chunk: the %s index is not loaded", opt->use_ert ? "ERT" : "FM");
    if (opt->use_emf && !ctx->d_emf_seed) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "mem_process_chunk: use_emf without a table (bwagpu_emf_load)");
    return BWAGPU_OK;
}

extern "C" int bwagpu_mem_process_chunk_device(bwagpu_ctx *ctx, const bwagpu_mem_opt *opt, const uint8_t *d_reads,
                                               const int64_t *d_off, const int32_t *d_len, int32_t nseq, int32_t max_len,
                                               bwagpu_mem_result *res)
{
    int rc = mem_check(ctx, opt, res, nseq);
    if (rc) return rc;
    memset(res, 0, sizeof *res);
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    int64_t *reg_off;
    if ((rc = bwagpu_scratch(ctx, SCR_MEM_REGOFF, (int64_t)(nseq + 1) * 8, (void **)&reg_off))) return rc;
    res->d_reg_off = reg_off;
    if (nseq == 0) { BWAGPU_CUDA(ctx, cudaMemsetAsync(reg_off, 0, 8, ctx->stream)); return BWAGPU_OK; }
    // S0: exact-match filter
    uint8_t *d_skip = nullptr;
    if (opt->use_emf) {
        int32_t *d_ret; bwagpu_perfect *d_perfect; unsigned long long *d_n;
        if ((rc = bwagpu_scratch(ctx, SCR_MEM_EMF_RET, (int64_t)nseq * 4, (void **)&d_ret))) return rc;
        if ((rc = bwagpu_scratch(ctx, SCR_MEM_PERFECT, (int64_t)nseq * 8, (void **)&d_perfect))) return rc;
        if ((rc = bwagpu_scratch(ctx, SCR_MEM_SKIP, (int64_t)nseq + 8, (void **)&d_skip))) return rc;
        if ((rc = bwagpu_scratch(ctx, SCR_MEM_CNT, 64, (void **)&d_n))) return rc;
        if ((rc = bwagpu_emf_probe_device(ctx, d_reads, d_off, d_len, nseq, d_ret, d_perfect, d_skip))) return rc;
        BWAGPU_CUDA(ctx, cudaMemsetAsync(d_n, 0, 8, ctx->stream));
        mem_count_skip_kernel<<<ctx->sm_count * 2, 256, 0, ctx->stream>>>(d_skip, nseq, d_n);
        ctx->launches++;
        unsigned long long h = 0;
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(&h, d_n, 8, cudaMemcpyDeviceToHost, ctx->stream));
        BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
        res->d_emf_ret = d_ret; res->d_perfect = d_perfect; res->n_emf_hits = (int64_t)h;
    }
    // S1 / S1': seeding (+ SA), then chaining
    bwagpu_chain_result cres; memset(&cres, 0, sizeof cres);
    if (opt->use_ert) {
        bwagpu_ert_result er;
        if ((rc = bwagpu_seed_ert_device(ctx, &opt->seed, d_reads, d_off, d_len, d_skip, nseq, max_len, &er))) return rc;
        res->n_smem = er.n_smem; res->n_coords = er.n_coords; res->n_index_bytes = er.n_index_bytes;
        if ((rc = bwagpu_chain_mems_device(ctx, &opt->chain, er.d_smems, er.n_smem, er.d_smem_off, er.d_coords, er.n_coords,
                                           er.d_coord_off, d_len, nseq, max_len, &cres))) return rc;
    } else {
        bwagpu_seed_result sr;
        if ((rc = bwagpu_seed_fm_device(ctx, &opt->seed, d_reads, d_off, d_len, d_skip, nseq, max_len, &sr))) return rc;
        res->n_smem = sr.n_smem; res->n_coords = sr.n_coords;
        res->n_ext = sr.n_ext; res->n_cp_records = sr.n_cp_records; res->n_lf_steps = sr.n_lf_steps;
        // mem_chain_seeds runs once per BATCH_SIZE reads upstream and its `pos < num_smem - 1` loop bound (:819) makes a
        // batch with a single SMEM yield nothing: the unit of that rule is opt->batch_size reads
        ctx->chain_batch_reads = opt->batch_size;
        rc = bwagpu_chain_seeds_device(ctx, &opt->chain, sr.d_smems, sr.n_smem, sr.d_smem_off, sr.d_coords, sr.n_coords,
                                       sr.d_coord_off, d_len, nseq, max_len, &cres);
        ctx->chain_batch_reads = 0;
        if (rc) return rc;
    }
    res->n_chains = cres.n_chains; res->n_seeds = cres.n_seeds;
    // S2: extension
    bwagpu_c2a_result ar;
    if ((rc = bwagpu_chain2aln_device(ctx, &opt->ext, d_reads, d_off, d_len, nseq, cres.d_chains, cres.n_chains,
                                      cres.d_seeds, cres.n_seeds, &ar))) return rc;
    res->n_regs_raw = ar.n_regs; res->n_sw_left = ar.n_left; res->n_sw_right = ar.n_right;
    res->n_sw_retry = ar.n_retry_left + ar.n_retry_right;
    res->n_pairs8 = ar.n_pairs8; res->n_pairs16 = ar.n_pairs16; res->n_pairs_scalar = ar.n_pairs_scalar;
    res->sw_cells = ar.sw_cells;
    // compaction (bwamem.cpp:1449-1460) and per-read offsets
    if (ar.n_regs == 0) {
        mem_zero_off_kernel<<<(nseq + 256) / 256, 256, 0, ctx->stream>>>(reg_off, nseq);
        ctx->launches++;
        BWAGPU_CUDA(ctx, cudaGetLastError());
        return BWAGPU_OK;
    }
    const int64_t n = ar.n_regs;
    int64_t *keep, *pos;
    if ((rc = bwagpu_scratch(ctx, SCR_MEM_KEEP, (n + 1) * 8, (void **)&keep))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_MEM_POS, (n + 1) * 8, (void **)&pos))) return rc;
    mem_keep_kernel<<<(unsigned)((n + 256) / 256), 256, 0, ctx->stream>>>(ar.d_regs, n, keep);
    ctx->launches++;
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum((void *)nullptr, bytes, keep, pos, (int)(n + 1), ctx->stream);
    void *tmp;
    if ((rc = bwagpu_scratch(ctx, SCR_CUB, (int64_t)bytes + 256, &tmp))) return rc;
    BWAGPU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, bytes, keep, pos, (int)(n + 1), ctx->stream));
    ctx->launches += 2;
    int64_t total = 0;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&total, pos + n, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    bwagpu_alnreg *out;
    if ((rc = bwagpu_scratch(ctx, SCR_MEM_REGS, (total + 1) * (int64_t)sizeof(bwagpu_alnreg), (void **)&out))) return rc;
    mem_compact_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ar.d_regs, n, pos, out);
    mem_regoff_kernel<<<(nseq + 256) / 256, 256, 0, ctx->stream>>>(cres.d_chains, cres.n_chains, n,
                                                                    (const int64_t *)ctx->scratch[SCR_C2A_FIRST], nseq, pos, reg_off);
    ctx->launches += 2;
    BWAGPU_CUDA(ctx, cudaGetLastError());
    res->d_regs = out; res->n_regs = total;
    return BWAGPU_OK;
}

extern "C" int bwagpu_mem_process_chunk(bwagpu_ctx *ctx, const bwagpu_mem_opt *opt, const uint8_t *enc_reads,
                                        const int64_t *off, const int32_t *len, int32_t nseq, bwagpu_mem_result *res)
{
    int rc = mem_check(ctx, opt, res, nseq);
    if (rc) return rc;
    if (nseq > 0 && (!enc_reads || !off || !len)) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "mem_process_chunk: NULL buffer");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    int64_t bytes = 0; int max_len = 0;
    for (int i = 0; i < nseq; ++i) {
        if (len[i] < 0 || off[i] < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "mem_process_chunk: read %d has a negative length/offset", i);
        if (len[i] > 32767) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "mem_process_chunk: read %d longer than 32767 (int16 query positions upstream)", i);
        if (off[i] + len[i] > bytes) bytes = off[i] + len[i];
        if (len[i] > max_len) max_len = len[i];
    }
    uint8_t *d_reads; int64_t *d_off; int32_t *d_len;
    if ((rc = bwagpu_scratch(ctx, SCR_MEM_READS, bytes + 16, (void **)&d_reads))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_MEM_OFF, (int64_t)(nseq + 1) * 8, (void **)&d_off))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_MEM_LEN, (int64_t)(nseq + 1) * 4, (void **)&d_len))) return rc;
    if (nseq > 0) {
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_reads, enc_reads, (size_t)bytes, cudaMemcpyHostToDevice, ctx->stream));
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_off, off, (size_t)nseq * 8, cudaMemcpyHostToDevice, ctx->stream));
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_len, len, (size_t)nseq * 4, cudaMemcpyHostToDevice, ctx->stream));
    }
    return bwagpu_mem_process_chunk_device(ctx, opt, d_reads, d_off, d_len, nseq, max_len, res);
}

extern "C" int bwagpu_mem_fetch(bwagpu_ctx *ctx, const bwagpu_mem_result *res, int32_t nseq, bwagpu_alnreg *regs,
                                int64_t *reg_off, int32_t *emf_ret, bwagpu_perfect *perfect)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!res || nseq < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "mem_fetch: bad argument");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    if (regs && res->n_regs > 0)
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(regs, res->d_regs, (size_t)res->n_regs * sizeof(bwagpu_alnreg), cudaMemcpyDeviceToHost, ctx->stream));
    if (reg_off && res->d_reg_off)
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(reg_off, res->d_reg_off, (size_t)(nseq + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (emf_ret) {
        if (res->d_emf_ret) BWAGPU_CUDA(ctx, cudaMemcpyAsync(emf_ret, res->d_emf_ret, (size_t)nseq * 4, cudaMemcpyDeviceToHost, ctx->stream));
        else memset(emf_ret, 0, (size_t)nseq * 4);                       // FIND_PERFECT_NO_TABLE
    }
    if (perfect) {
        if (res->d_perfect) BWAGPU_CUDA(ctx, cudaMemcpyAsync(perfect, res->d_perfect, (size_t)nseq * 8, cudaMemcpyDeviceToHost, ctx->stream));
        else memset(perfect, 0, (size_t)nseq * 8);
    }
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}
