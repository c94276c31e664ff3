// TEST SCAFFOLDING — single-steps the per-thread seeding code of the CUDA kernels
// (bwa_mem_scale_b200/csrc/fmi_core.cuh) on the host, one "thread" at a time, so that the state
// machine can be checked against the oracle without a GPU.  Never part of libbwagpu.so.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../bwa_mem_scale_b200/csrc/fmi_core.cuh"

struct HostPrev {
    int64_t *buf;
    void reset(int) {}
    void put(int e, const Biv &v, int n) { buf[3 * e] = v.k | ((int64_t)n << 40); buf[3 * e + 1] = v.l; buf[3 * e + 2] = v.s; }
    void get(int e, Biv &v, int &n) const { v.k = buf[3 * e] & ((1LL << 40) - 1); n = (int)(buf[3 * e] >> 40); v.l = buf[3 * e + 1]; v.s = buf[3 * e + 2]; }
};

extern "C" int64_t hostsim_collect_smem(const bwagpu_fm_index_desc *d, const bwagpu_seed_opt *o,
                                        const uint8_t *reads, const int64_t *off, const int32_t *len,
                                        const uint8_t *skip, int32_t nseq, bwagpu_smem *out,
                                        int64_t cap, int per_read_cap, int64_t *n_ext,
                                        const uint32_t *fma_all, const uint32_t *fma_last)
{
    FmView f;
    f.fma_all = fma_all; f.fma_last = fma_last;
    f.cp_occ = (const int64_t *)d->cp_occ; f.sa_ms_byte = d->sa_ms_byte; f.sa_ls_word = d->sa_ls_word;
    memcpy(f.count, d->count, sizeof f.count);
    f.sentinel_index = d->sentinel_index; f.ref_seq_len = d->reference_seq_len;
    SeedOpt so;
    so.min_seed_len = o->min_seed_len; so.split_len = (int)(o->min_seed_len * o->split_factor + .499);
    so.split_width = o->split_width; so.max_mem_intv = o->max_mem_intv;
    int64_t tot = 0;
    std::vector<bwagpu_smem> tmp(per_read_cap);
    for (int i = 0; i < nseq; ++i) {
        if (skip && skip[i]) continue;
        std::vector<int64_t> pv(3 * (size_t)(len[i] + 1));
        SeedMachine<HostPrev> m;
        HostPrev hp; hp.buf = pv.data();
        m.init(reads + off[i], len[i], i, hp, tmp.data(), per_read_cap);
        // same driver shape as the kernel: common transitions inline, the rest through advance()
        bool ready = false;
        for (;;) {
            if (!ready && !m.advance(f, so)) break;
#if SEED_PAIR2
            Biv r2;
            Biv r = fm_bext2(f, m.req, m.req2, m.has2, m.req_a, r2);
            ready = m.step_fast(r, so, &r2);
#else
            Biv r = fm_bext(f, m.req, m.req_a);
            ready = m.step_fast(r, so);
#endif
        }
        if (m.n_out > per_read_cap) return -(int64_t)m.n_out;     // caller retries with more room
        if (n_ext) *n_ext += m.n_ext;
        std::stable_sort(tmp.begin(), tmp.begin() + m.n_out, [](const bwagpu_smem &a, const bwagpu_smem &b) {
            return a.m != b.m ? a.m < b.m : a.n < b.n; });
        for (int e = 0; e < m.n_out; ++e, ++tot) if (tot < cap) out[tot] = tmp[e];
    }
    return tot;
}

extern "C" int64_t hostsim_sa_lookup(const bwagpu_fm_index_desc *d, const bwagpu_smem *sm, int64_t n,
                                     int32_t max_occ, int64_t *coords)
{
    FmView f;
    f.fma_all = f.fma_last = nullptr;
    f.cp_occ = (const int64_t *)d->cp_occ; f.sa_ms_byte = d->sa_ms_byte; f.sa_ls_word = d->sa_ls_word;
    memcpy(f.count, d->count, sizeof f.count);
    f.sentinel_index = d->sentinel_index; f.ref_seq_len = d->reference_seq_len;
    int64_t w = 0;
    for (int64_t i = 0; i < n; ++i) {
        int64_t step;
        int c = fm_sa_rows(sm[i].s, max_occ, &step);
        for (int r = 0; r < c; ++r) coords[w++] = fm_sa_walk(f, sm[i].k + r * step, nullptr);
    }
    return w;
}

// FMA table entries as the build kernels compute them (fma_entry_all / fma_entry_last)
extern "C" void hostsim_fma_entries(const bwagpu_fm_index_desc *d, int which_last, const uint32_t *idx,
                                    int64_t n, uint32_t *out)
{
    FmView f;
    f.fma_all = f.fma_last = nullptr;
    f.cp_occ = (const int64_t *)d->cp_occ; f.sa_ms_byte = d->sa_ms_byte; f.sa_ls_word = d->sa_ls_word;
    memcpy(f.count, d->count, sizeof f.count);
    f.sentinel_index = d->sentinel_index; f.ref_seq_len = d->reference_seq_len;
    for (int64_t i = 0; i < n; ++i) {
        if (which_last) fma_entry_last(f, idx[i], out + 4 * i);
        else fma_entry_all(f, idx[i], out + 32 * i);
    }
}
