// TEST SCAFFOLDING — runs the per-thread chaining code of the CUDA kernels
// (bwa_mem_scale_b200/csrc/chain_core.cuh) on the host, one read at a time, so that the B-tree /
// introsort replay can be checked against the reference's chains without a GPU.  Never part of
// libbwagpu.so.
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../bwa_mem_scale_b200/csrc/chain_core.cuh"

extern "C" int64_t hostsim_chain_seeds(const bwagpu_chain_opt *opt, int64_t l_pac, int n_seqs, const int64_t *contig_off,
                                       const int32_t *contig_len, const bwagpu_smem *smems, int64_t n_smem,
                                       const int64_t *coords, const int32_t *len, int32_t nseq,
                                       bwagpu_chain *out_chains, bwagpu_seed *out_seeds, int64_t *n_seeds_out)
{
    ChainOpt o = { opt->w, opt->max_chain_gap, opt->max_occ, opt->min_seed_len, opt->min_chain_weight,
                   opt->max_chain_extend, opt->mask_level, opt->drop_ratio };
    RefView rv = { l_pac, n_seqs, contig_off, contig_len, nullptr };
    int64_t nc = 0, ns = 0, cpos = 0, i = 0;
    for (int r = 0; r < nseq; ++r) {
        int64_t s0 = i;
        while (i < n_smem && (int)smems[i].rid == r) ++i;
        const int64_t s1 = i;
        // coordinates of this read
        int64_t ncoord = 0;
        std::vector<int64_t> rows(s1 - s0);
        for (int64_t k = s0; k < s1; ++k) {
            const int64_t s = smems[k].s, step = s > o.max_occ ? s / o.max_occ : 1;
            int64_t c = (s + step - 1) / step; if (c > o.max_occ) c = o.max_occ;
            rows[k - s0] = c; ncoord += c;
        }
        const int64_t c0 = cpos; cpos += ncoord;
        if (s0 == s1 || n_smem <= 1 || len[r] < o.min_seed_len) continue;
        std::vector<bwagpu_seed> sd(ncoord + 1); std::vector<WChain> ch(ncoord + 1);
        std::vector<BtNode> nodes(ncoord / 2 + 8); std::vector<int32_t> ord(ncoord + 1), lst(ncoord + 1);
        ChainScratch sc = { sd.data(), ch.data(), nodes.data(), ord.data(), lst.data() };
        ChainBuilder b; b.init(sc);
        int bb = 0, e = 0, l_rep = 0;
        for (int64_t k = s0; k < s1; ++k) {
            const int sb = (int)smems[k].m, se = (int)smems[k].n + 1;
            if (smems[k].s <= o.max_occ) continue;
            if (sb > e) { l_rep += e - bb; bb = sb; e = se; } else e = e > se ? e : se;
        }
        l_rep += e - bb;
        int64_t c = c0;
        for (int64_t k = s0; k < s1; ++k) {
            bwagpu_seed p; p.qbeg = (int32_t)smems[k].m; p.len = p.score = (int32_t)(smems[k].n + 1 - smems[k].m); p.pad_ = -1;
            for (int64_t x = 0; x < rows[k - s0]; ++x) { p.rbeg = coords[c++]; b.add_seed(o, rv, p); }
        }
        b.traverse();
        const int m = b.filter(o);
        for (int q = 0; q < m; ++q) {
            const WChain &wc = ch[ord[q]];
            bwagpu_chain oc;
            oc.seqid = r; oc.rid = wc.rid; oc.n = wc.n; oc.w = wc.w; oc.kept = wc.kept; oc.first = wc.first; oc.is_alt = wc.is_alt;
            oc.frac_rep = (float)l_rep / len[r]; oc.pos = wc.pos; oc.seed_off = ns;
            out_chains[nc++] = oc;
            for (int si = wc.head; si >= 0; si = sd[si].pad_) { out_seeds[ns] = sd[si]; out_seeds[ns].pad_ = 0; ++ns; }
        }
    }
    *n_seeds_out = ns;
    return nc;
}
