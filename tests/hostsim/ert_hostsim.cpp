// TEST SCAFFOLDING — runs the per-thread ERT seeding code of the CUDA kernels
// (bwa_mem_scale_b200/csrc/ert_core.cuh) on the host, one "thread" (= read) at a time, so that it can be
// checked against the reference tap without a GPU.  Never part of libbwagpu.so.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
static std::vector<uint32_t> *g_touched = nullptr;      // k-mer entries read by the walks (golden fixture maker)
#define ERT_KMER_LOAD(ix, h) ((g_touched ? g_touched->push_back(h) : (void)0), (ix).kmer_offsets[h])
#include "../../bwa_mem_scale_b200/csrc/ert_core.cuh"

struct hs_mem { int32_t rid, start, end, hitcount; int64_t hit_off; };

extern "C" int hostsim_ert_offsets(int ptr_width, int code, int c, int which)
{
    return which ? ert_child_off(ptr_width, (uint8_t)code, c) : ert_leaf_off(ptr_width, (uint8_t)code, c);
}

// same contract as ref_ert_seed (oracle/refwrap/ref_ert_wrap.cpp)
extern "C" int hostsim_ert_seed(const uint64_t *kmer, const uint8_t *mlt, const uint8_t *ref, int64_t l_pac,
                                int min_seed_len, float split_factor, int split_width, int max_mem_intv,
                                const uint8_t *reads, const int64_t *off, const int32_t *len, int nseq,
                                hs_mem *mems, int64_t mem_cap, int64_t *hits_out, int64_t hit_cap,
                                int64_t *mem_off, int64_t *n_out, int per_read_mems, int per_read_hits)
{
    ErtOpt o = {min_seed_len, (int)(min_seed_len * split_factor + .499), split_width, max_mem_intv};
    int64_t nm = 0, nh = 0, bytes = 0;
    for (int l = 0; l < nseq; ++l) {
        mem_off[l] = nm;
        if (len[l] > ERT_MAX_READ) continue;
        int mc = per_read_mems, hc = per_read_hits;
        for (;;) {
            std::vector<ErtMemRec> mr(mc);
            std::vector<int64_t> hv(hc);
            ErtWalker w;
            w.ix.kmer_offsets = kmer; w.ix.mlt = mlt; w.ix.ref = ref; w.ix.l_pac = l_pac;
            w.fwd = reads + off[l]; w.l_seq = len[l];
            w.mems = mr.data(); w.mem_cap = mc; w.hits = hv.data(); w.hit_cap = hc;
            w.seed_read(o);
            if (w.n_mem > mc || w.hit_peak > hc) {                 // the kernel's retry-with-more-room rule
                if (w.n_mem > mc) mc = w.n_mem * 2;
                if (w.hit_peak > hc) hc = w.hit_peak * 2;
                continue;
            }
            bytes += w.n_bytes;
            std::vector<uint64_t> keys(w.n_mem);
            for (int k = 0; k < w.n_mem; ++k) keys[k] = ((uint64_t)mr[k].start << 40) | ((uint64_t)mr[k].end << 20) | (uint64_t)k;
            ErtSort::introsort(keys.data(), w.n_mem);
            for (int k = 0; k < w.n_mem; ++k) {
                const ErtMemRec &r = mr[keys[k] & 0xFFFFF];
                if (nm < mem_cap) { mems[nm].rid = l; mems[nm].start = r.start; mems[nm].end = r.end;
                                    mems[nm].hitcount = r.hitcount; mems[nm].hit_off = nh; }
                ++nm;
                for (int q = 0; q < r.hitcount; ++q, ++nh)
                    if (nh < hit_cap) hits_out[nh] = ert_hit_rbeg(r, hv[r.hitbeg + q], l_pac);
            }
            break;
        }
    }
    mem_off[nseq] = nm;
    n_out[0] = nm; n_out[1] = nh; n_out[2] = bytes;
    return (nm <= mem_cap && nh <= hit_cap) ? 0 : -1;
}

// start / stop logging the k-mer table indices the walks read; stop copies up to cap of them and returns the count
extern "C" void hostsim_ert_log_start(void) { delete g_touched; g_touched = new std::vector<uint32_t>(); }
extern "C" int64_t hostsim_ert_log_stop(uint32_t *out, int64_t cap)
{
    if (!g_touched) return 0;
    const int64_t n = (int64_t)g_touched->size();
    for (int64_t i = 0; i < n && i < cap; ++i) out[i] = (*g_touched)[i];
    delete g_touched; g_touched = nullptr;
    return n;
}

// k-mer table low bits from occurrence counts (csrc/ert_build.cuh): out[i] = lep | hits << 16 (hits capped at 2^40)
#include "../../bwa_mem_scale_b200/csrc/ert_build.cuh"
extern "C" void hostsim_ert_kmer_lep(const bwagpu_fm_index_desc *d, const uint32_t *idx, int64_t n, uint64_t *out)
{
    FmView f;
    f.fma_all = f.fma_last = nullptr;
    f.cp_occ = (const int64_t *)d->cp_occ; f.sa_ms_byte = d->sa_ms_byte; f.sa_ls_word = d->sa_ls_word;
    memcpy(f.count, d->count, sizeof f.count);
    f.sentinel_index = d->sentinel_index; f.ref_seq_len = d->reference_seq_len;
    for (int64_t i = 0; i < n; ++i) {
        int64_t hits;
        const uint32_t lep = ert_kmer_lep(f, idx[i], &hits);
        out[i] = (uint64_t)lep | ((uint64_t)hits << 16);
    }
}
