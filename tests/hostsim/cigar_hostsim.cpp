// TEST SCAFFOLDING — runs the per-thread CIGAR code of the CUDA kernel (bwa_mem_scale_b200/csrc/cigar_core.cuh) on the
// host, one job at a time, so that it can be checked against the reference tap without a GPU.  Never part of
// libbwagpu.so.  Same contract as ref_reg2aln (oracle/refwrap/ref_api_wrap.cpp).
#include <cstdint>
#include <cstring>
#include <vector>
#include "../../bwa_mem_scale_b200/csrc/cigar_core.cuh"

extern "C" int hostsim_reg2aln(const bwagpu_cigar_opt *opt, int64_t l_pac, const uint8_t *ref_string, const uint8_t *reads,
                               const int64_t *off, const int32_t *len, const bwagpu_cigar_job *jobs, int64_t n_jobs,
                               bwagpu_cigar_out *outs, uint32_t *cigar, int64_t cigar_cap, char *md, int64_t md_cap, int64_t *n_out)
{
    CigarOpt o;
    memcpy(o.mat, opt->mat, 25);
    o.a = opt->a; o.o_del = opt->o_del; o.e_del = opt->e_del; o.o_ins = opt->o_ins; o.e_ins = opt->e_ins; o.w = opt->w;
    int64_t nc = 0, nm = 0;
    n_out[2] = 0;
    for (int64_t j = 0; j < n_jobs; ++j) {
        const bwagpu_cigar_job &b = jobs[j];
        const int64_t lq = b.qe - b.qb, lt = b.re > b.rb ? b.re - b.rb : 0;
        std::vector<int32_t> eh(2 * (lq + 1) + 4);
        std::vector<uint8_t> z(cig_z_bytes(o, b.qb, b.qe, b.rb, b.re, b.truesc, b.w) + 1);   // exactly what the kernel's planner reserves
        std::vector<uint32_t> cg(lq + lt + 8);
        std::vector<char> mdv(3 * lt + 32);
        // rows: the circular window of the region's tier, as the kernel uses it (tier 4 = full rows)
        const int tier = cig_tier(o, b.qb, b.qe, b.rb, b.re, b.truesc, b.w);
        const int mask = tier == 4 ? 0x7fffffff : (16 << tier) - 1;
        if (tier != 4) eh.assign(2 * (size_t)(16 << tier), 0);
        CigarScratch sc = { eh.data(), 1, mask, z.data(), cg.data(), mdv.data() };
        CigarResult r; r.score = 0;
        bwagpu_cigar_out &out = outs[j];
        memset(&out, 0, sizeof out);
        out.pos = -1; out.nm = -1;
        int rc = cig_reg2aln(o, l_pac, ref_string, reads + off[b.read], b.qb, b.qe, b.rb, b.re, b.truesc, b.w, sc, &r);
        if (rc == 2) {                                   // a later try outgrew the window: the kernel's second pass, full rows
            eh.assign(2 * (lq + 1) + 4, 0);
            sc.eh = eh.data(); sc.eh_mask = 0x7fffffff;
            rc = cig_reg2aln(o, l_pac, ref_string, reads + off[b.read], b.qb, b.qe, b.rb, b.re, b.truesc, b.w, sc, &r);
            ++n_out[2];
        }
        out.ok = rc == 1 ? 1 : 0;
        out.cigar_off = nc; out.md_off = nm;
        if (!out.ok) continue;
        int is_rev;
        out.n_cigar = cig_finish(l_pac, b.rb, b.re, b.qb, b.qe, len[b.read], sc.cigar, r.n_cigar, &out.pos, &is_rev);
        out.is_rev = is_rev; out.nm = r.nm; out.l_md = r.l_md; out.score = r.score;
        for (int k = 0; k < out.n_cigar; ++k, ++nc) if (nc < cigar_cap) cigar[nc] = sc.cigar[k];
        for (int k = 0; k < out.l_md; ++k, ++nm) if (nm < md_cap) md[nm] = sc.md[k];
        if (nm < md_cap) md[nm] = 0;
        ++nm;
    }
    n_out[0] = nc; n_out[1] = nm;
    return (nc <= cigar_cap && nm <= md_cap) ? 0 : -1;
}
