// Chain -> alignment-region stage of the C ABI: bwagpu_index_load_ref, bwagpu_chain2aln{,_device}
// (include/bwagpu.h).  Replaces mem_chain2aln_across_reads_V2 (reference src/bwamem.cpp:2810-3745):
// extension-job construction from chains, the 8-bit / 16-bit / scalar kernel routing of
// sortPairsLenExt (:2667-2740), the two-try band doubling with result application (:3218-3618) and
// the containment filter (:3634-3745), all on the device: the reference window of every job is read
// straight from the HBM-resident .0123 text, nothing but chains goes in and regions come out.
//
// Kernels (all one-thread-per-item, integer/byte work bound by HBM latency; the SW itself is
// bsw.cu's kernels):
//   c2a_chain_kernel   per chain: reference span rmax (:2892-2913) clipped to the contig
//                      (bns_fetch_seq_v2, src/bntseq.cpp:484-517), seeds ordered by (score, index)
//   c2a_size_kernel    per region: which jobs it needs and their byte sizes (then 6 prefix sums)
//   c2a_fill_kernel    per region: mem_alnreg_t init, left/right SeqPair, sequence bytes (:2953-3190)
//   c2a_route_kernel   per job: bucket by kernel kind;  c2a_h0_kernel: right h0 = left score (:3423)
//   c2a_apply_kernel   per job: band-retry decision and result application
//   c2a_filter_kernel  per read: the sequential containment filter
#include <cub/device/device_scan.cuh>
#include <stdlib.h>
#include "ctx.cuh"

#define H0_ (-99)

struct C2aOpt { int a, b, o_del, e_del, o_ins, e_ins, pen_clip5, pen_clip3, w, zdrop; };

__device__ __forceinline__ int c2a_max_gap(const C2aOpt &o, int qlen)           // cal_max_gap :94-104
{
    const int l_del = (int)((double)(qlen * o.a - o.o_del) / o.e_del + 1.);
    const int l_ins = (int)((double)(qlen * o.a - o.o_ins) / o.e_ins + 1.);
    int l = l_del > l_ins ? l_del : l_ins;
    l = l > 1 ? l : 1;
    return l < o.w << 1 ? l : o.w << 1;
}

__device__ __forceinline__ int c2a_kind(int len1, int len2, int h0, int a)       // sortPairsLenExt
{
    const int minval = h0 + min(len1, len2) * a;
    if (len1 < 128 && len2 < 128 && minval < 128) return 2;      // 8-bit
    if (len1 < 32768 && len2 < 32768 && minval < 32768) return 1; // 16-bit
    return 0;                                                     // scalar
}

// ---- per chain ------------------------------------------------------------------------------
__global__ void c2a_chain_kernel(C2aOpt o, int64_t l_pac, const int64_t *__restrict__ contig_off,
                                 const int32_t *__restrict__ contig_len, const int32_t *__restrict__ len,
                                 const bwagpu_chain *__restrict__ chains, int64_t n_chains,
                                 const bwagpu_seed *__restrict__ seeds, int64_t *__restrict__ rmax,
                                 uint64_t *__restrict__ keys, uint32_t *__restrict__ srt,
                                 int32_t *__restrict__ chain_of, int64_t *__restrict__ first_chain,
                                 int64_t *__restrict__ last_chain)
{
    const int64_t ci = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (ci >= n_chains) return;
    const bwagpu_chain c = chains[ci];
    const bwagpu_seed *cs = seeds + c.seed_off;
    const int l_query = len[c.seqid];
    if (ci == 0 || chains[ci - 1].seqid != c.seqid) first_chain[c.seqid] = ci;
    if (ci == n_chains - 1 || chains[ci + 1].seqid != c.seqid) last_chain[c.seqid] = ci;
    int64_t r0 = l_pac << 1, r1 = 0;
    for (int i = 0; i < c.n; ++i) {
        const bwagpu_seed t = cs[i];
        const int64_t b = t.rbeg - (t.qbeg + c2a_max_gap(o, t.qbeg));
        const int64_t e = t.rbeg + t.len + ((l_query - t.qbeg - t.len) + c2a_max_gap(o, l_query - t.qbeg - t.len));
        r0 = r0 < b ? r0 : b;
        r1 = r1 > e ? r1 : e;
        keys[c.seed_off + i] = ((uint64_t)(uint32_t)t.score << 32) | (uint32_t)i;
        chain_of[c.seed_off + i] = (int32_t)ci;
    }
    r0 = r0 > 0 ? r0 : 0;
    r1 = r1 < (l_pac << 1) ? r1 : (l_pac << 1);
    if (c.n > 0) {
        const int64_t rb0 = cs[0].rbeg;
        if (r0 < l_pac && l_pac < r1) { if (rb0 < l_pac) r1 = l_pac; else r0 = l_pac; }
        int64_t fb = contig_off[c.rid], fe = fb + contig_len[c.rid];
        if (rb0 >= l_pac) { const int64_t t = fb; fb = (l_pac << 1) - fe; fe = (l_pac << 1) - t; }
        r0 = r0 > fb ? r0 : fb;
        r1 = r1 < fe ? r1 : fe;
    }
    rmax[2 * ci] = r0; rmax[2 * ci + 1] = r1;
    // order the seeds by (score, index) ascending; keys are unique so any sort is the reference's
    uint64_t *k = keys + c.seed_off;
    const int n = c.n;
    if (n <= 24) {
        for (int i = 1; i < n; ++i) {
            const uint64_t v = k[i];
            int j = i - 1;
            while (j >= 0 && k[j] > v) { k[j + 1] = k[j]; --j; }
            k[j + 1] = v;
        }
    } else {                                                     // heapsort
        for (int start = n / 2 - 1; start >= 0; --start) {
            int root = start;
            for (;;) {
                int child = 2 * root + 1;
                if (child >= n) break;
                if (child + 1 < n && k[child] < k[child + 1]) ++child;
                if (k[root] >= k[child]) break;
                const uint64_t t = k[root]; k[root] = k[child]; k[child] = t;
                root = child;
            }
        }
        for (int end = n - 1; end > 0; --end) {
            const uint64_t t0 = k[0]; k[0] = k[end]; k[end] = t0;
            int root = 0;
            for (;;) {
                int child = 2 * root + 1;
                if (child >= end) break;
                if (child + 1 < end && k[child] < k[child + 1]) ++child;
                if (k[root] >= k[child]) break;
                const uint64_t t = k[root]; k[root] = k[child]; k[child] = t;
                root = child;
            }
        }
    }
    for (int i = 0; i < n; ++i) srt[c.seed_off + i] = (uint32_t)k[i];
}

// region ri of chain c (seed_off <= ri < seed_off + n) belongs to the seed srt[seed_off + (n-1) - (ri-seed_off)]
__device__ __forceinline__ bwagpu_seed c2a_seed_of(const bwagpu_chain &c, const bwagpu_seed *seeds,
                                                   const uint32_t *srt, int64_t ri)
{
    const int k = c.n - 1 - (int)(ri - c.seed_off);
    return seeds[c.seed_off + srt[c.seed_off + k]];
}

// sizes[0..5][ri]: has left, has right, left query bytes, left ref bytes, right query bytes, right ref bytes
__global__ void c2a_size_kernel(const int32_t *__restrict__ len, const bwagpu_chain *__restrict__ chains,
                                const bwagpu_seed *__restrict__ seeds, const uint32_t *__restrict__ srt,
                                const int32_t *__restrict__ chain_of, const int64_t *__restrict__ rmax,
                                int64_t n_regs, int64_t *__restrict__ sizes, int64_t stride)
{
    const int64_t ri = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (ri >= n_regs) return;
    const int ci = chain_of[ri];
    const bwagpu_chain c = chains[ci];
    const bwagpu_seed s = c2a_seed_of(c, seeds, srt, ri);
    const int l_query = len[c.seqid];
    const int64_t r0 = rmax[2 * ci], r1 = rmax[2 * ci + 1];
    const bool hl = s.qbeg != 0, hr = s.qbeg + s.len != l_query;
    sizes[ri] = hl; sizes[stride + ri] = hr;
    sizes[2 * stride + ri] = hl ? s.qbeg : 0;
    sizes[3 * stride + ri] = hl ? s.rbeg - r0 : 0;
    sizes[4 * stride + ri] = hr ? l_query - (s.qbeg + s.len) : 0;
    sizes[5 * stride + ri] = hr ? r1 - r0 - (s.rbeg + s.len - r0) : 0;
}

// C2A_DIRECT = 1: the extension jobs do not carry copies of their sequences.  A job names where its query starts in the
// chunk's read buffer and where its window starts in the resident reference text, plus a direction (left extensions walk
// both backwards), and the SW kernels read from there (BswView, bsw.cu).  No sequence arenas (2.3 GB written and read
// again per million reads), no 2^31-byte limit on a chunk's extension sequences, four prefix sums fewer.
#ifndef C2A_DIRECT
#define C2A_DIRECT 1
#endif
__device__ __forceinline__ void c2a_direct_addr(bwagpu_seqpair &sp, int64_t qpos, int64_t rpos, int descending, int *err)
{
    if (qpos < 0 || qpos >= (1LL << 32) || rpos >= (1LL << 56)) *err = 1;     // outside what the job record can name
    if (rpos < 0) rpos = 0;                                                    // empty window at the very start of the text
    sp.idq = (int32_t)(uint32_t)qpos;
    sp.idr = (int32_t)(uint32_t)rpos;
    sp.regid = (int32_t)(((uint32_t)(rpos >> 32) << 8) | (uint32_t)(descending & 1));
}

__global__ void c2a_fill_kernel(C2aOpt o, const uint8_t *__restrict__ ref_string, const uint8_t *__restrict__ reads,
                                const int64_t *__restrict__ off, const int32_t *__restrict__ len,
                                const bwagpu_chain *__restrict__ chains, const bwagpu_seed *__restrict__ seeds,
                                const uint32_t *__restrict__ srt, const int32_t *__restrict__ chain_of,
                                const int64_t *__restrict__ rmax, int64_t n_regs, const int64_t *__restrict__ offs,
                                int64_t stride, bwagpu_alnreg *__restrict__ regs, bwagpu_seqpair *__restrict__ L,
                                bwagpu_seqpair *__restrict__ R, uint8_t *__restrict__ LQ, uint8_t *__restrict__ LR,
                                uint8_t *__restrict__ RQ, uint8_t *__restrict__ RR, int *__restrict__ err)
{
    // one warp per region: every lane derives the same scalars, lane 0 stores the records, all 32
    // lanes copy the sequence bytes (coalesced instead of one serial byte loop per region)
#if C2A_DIRECT
    const int64_t ri = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;          // nothing to copy: one thread per region
    const int lane = 0;
#else
    const int64_t ri = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
#endif
    if (ri >= n_regs) return;
    const int ci = chain_of[ri];
    const bwagpu_chain c = chains[ci];
    const bwagpu_seed s = c2a_seed_of(c, seeds, srt, ri);
    const int l_query = len[c.seqid];
    [[maybe_unused]] const uint8_t *query = reads + off[c.seqid];
    const int64_t r0 = rmax[2 * ci], r1 = rmax[2 * ci + 1];
    [[maybe_unused]] const uint8_t *rseq = ref_string + r0;
    bwagpu_alnreg a;
    a.w = o.w; a.score = a.truesc = -1; a.rid = c.rid; a.frac_rep = c.frac_rep; a.seedlen0 = s.len;
    a.chain = ci; a.seedcov = 0;
    a.rb = a.re = H0_; a.qb = a.qe = H0_;
    if (s.qbeg) {
        bwagpu_seqpair sp;
        sp.h0 = s.len * o.a; sp.seqid = c.seqid; sp.regid = 0;
        sp.id = (int32_t)ri;
        const int64_t tmp = s.rbeg - r0;
#if C2A_DIRECT
        // query[j] = read[qbeg - 1 - j], target[i] = text[rbeg - 1 - i]
        c2a_direct_addr(sp, off[c.seqid] + s.qbeg - 1, s.rbeg - 1, 1, err);
#else
        const int64_t qo = offs[2 * stride + ri], ro = offs[3 * stride + ri];
        sp.idq = (int32_t)qo; sp.idr = (int32_t)ro;
        for (int i = lane; i < s.qbeg; i += 32) LQ[qo + i] = query[s.qbeg - 1 - i];
        for (int64_t i = lane; i < tmp; i += 32) LR[ro + i] = rseq[tmp - 1 - i];
#endif
        sp.len2 = s.qbeg; sp.len1 = (int32_t)tmp;
        sp.score = sp.tle = sp.gtle = sp.qle = sp.gscore = sp.max_off = 0;
        if (lane == 0) L[offs[ri]] = sp;
        a.qb = s.qbeg; a.rb = s.rbeg;
    } else {
        a.score = a.truesc = s.len * o.a; a.qb = 0; a.rb = s.rbeg;
    }
    if (s.qbeg + s.len != l_query) {
        const int64_t qe = s.qbeg + s.len, re = s.rbeg + s.len - r0;
        bwagpu_seqpair sp;
        sp.h0 = H0_; sp.seqid = c.seqid; sp.regid = 0; sp.id = (int32_t)ri;
        sp.len2 = (int32_t)(l_query - qe); sp.len1 = (int32_t)(r1 - r0 - re);
#if C2A_DIRECT
        c2a_direct_addr(sp, off[c.seqid] + qe, r0 + re, 0, err);                // query[j] = read[qe + j], target[i] = text[r0 + re + i]
#else
        const int64_t qo = offs[4 * stride + ri], ro = offs[5 * stride + ri];
        sp.idq = (int32_t)qo; sp.idr = (int32_t)ro;
        for (int i = lane; i < sp.len2; i += 32) RQ[qo + i] = query[qe + i];
        for (int i = lane; i < sp.len1; i += 32) RR[ro + i] = rseq[re + i];
#endif
        sp.score = sp.tle = sp.gtle = sp.qle = sp.gscore = sp.max_off = 0;
        if (lane == 0) R[offs[stride + ri]] = sp;
        a.qe = (int32_t)qe; a.re = r0 + re;
    } else {
        a.qe = l_query; a.re = s.rbeg + s.len;
        if (a.rb != H0_ && a.qb != H0_) {
            const bwagpu_seed *cs = seeds + c.seed_off;
            int cov = 0;
            for (int i = 0; i < c.n; ++i) {
                const bwagpu_seed t = cs[i];
                if (t.qbeg >= a.qb && t.qbeg + t.len <= a.qe && t.rbeg >= a.rb && t.rbeg + t.len <= a.re) cov += t.len;
            }
            a.seedcov = cov;
        }
    }
    if (lane == 0) regs[ri] = a;
}

// right jobs start from the score the left extension reached (:3423-3428)
__global__ void c2a_h0_kernel(bwagpu_seqpair *__restrict__ R, int64_t n, const bwagpu_alnreg *__restrict__ regs)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) R[i].h0 = regs[R[i].id].score;
}

// bucket the jobs by kernel kind: out[kind] gets a compacted copy (order within a bucket is free:
// results do not depend on lane-mates)
__global__ void c2a_route_kernel(C2aOpt o, const bwagpu_seqpair *__restrict__ in, int64_t n,
                                 bwagpu_seqpair *__restrict__ out0, bwagpu_seqpair *__restrict__ out1,
                                 bwagpu_seqpair *__restrict__ out2, int *__restrict__ cnt)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int kind = -1;
    bwagpu_seqpair sp;
    if (i < n) { sp = in[i]; kind = c2a_kind(sp.len1, sp.len2, sp.h0, o.a); }
    for (int k = 0; k < 3; ++k) {
        const unsigned m = __ballot_sync(0xffffffffu, kind == k);
        if (!m) continue;
        const int leader = __ffs(m) - 1, lane = threadIdx.x & 31;
        int base = 0;
        if (lane == leader) base = atomicAdd(&cnt[k], __popc(m));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (kind == k) {
            bwagpu_seqpair *dst = k == 0 ? out0 : k == 1 ? out1 : out2;
            dst[base + __popc(m & ((1u << lane) - 1))] = sp;
        }
    }
}

// band-retry decision + result application for one finished SW job (:3240-3283 left, :3458-3492 right)
__global__ void c2a_apply_kernel(C2aOpt o, int left, int w, int last_try, const bwagpu_seqpair *__restrict__ pairs,
                                 int n, bwagpu_alnreg *__restrict__ regs, const bwagpu_chain *__restrict__ chains,
                                 const bwagpu_seed *__restrict__ seeds, const int32_t *__restrict__ len,
                                 bwagpu_seqpair *__restrict__ retry, int *__restrict__ n_retry)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const bwagpu_seqpair sp = pairs[i];
    bwagpu_alnreg a = regs[sp.id];
    const int prev = a.score;
    a.score = sp.score;
    if (a.score == prev || sp.max_off < (w >> 1) + (w >> 2) || last_try) {
        if (left) {
            if (sp.gscore <= 0 || sp.gscore <= a.score - o.pen_clip5) { a.qb -= sp.qle; a.rb -= sp.tle; a.truesc = a.score; }
            else { a.qb = 0; a.rb -= sp.gtle; a.truesc = sp.gscore; }
        } else {
            if (sp.gscore <= 0 || sp.gscore <= a.score - o.pen_clip3) { a.qe += sp.qle; a.re += sp.tle; a.truesc += a.score - sp.h0; }
            else { a.qe = len[sp.seqid]; a.re += sp.gtle; a.truesc += sp.gscore - sp.h0; }
        }
        a.w = max(a.w, w);
        if (a.rb != H0_ && a.qb != H0_ && a.qe != H0_ && a.re != H0_) {
            const bwagpu_chain c = chains[a.chain];
            const bwagpu_seed *cs = seeds + c.seed_off;
            int cov = 0;
            for (int k = 0; k < c.n; ++k) {
                const bwagpu_seed t = cs[k];
                if (t.qbeg >= a.qb && t.qbeg + t.len <= a.qe && t.rbeg >= a.rb && t.rbeg + t.len <= a.re) cov += t.len;
            }
            a.seedcov = cov;
        }
        regs[sp.id] = a;
    } else {
        regs[sp.id].score = a.score;
        retry[atomicAdd(n_retry, 1)] = sp;
    }
}

// reads with more regions than this are filtered by a whole warp (c2a_filter_warp_kernel)
#define C2A_FILTER_WARP_MIN 48

// The containment filter for reads with many regions (repeats): one read per warp.  The scan over
// the regions created so far (:3652-3690) is the quadratic part; 32 of them are classified at a time
// (purged / not containing -> counts / close to the seed's diagonal -> stop) and the reference's
// sequential rules -- stop at the first "close" region, or once `lim` live regions have been seen --
// are resolved with ballots and a prefix population count.
__global__ void __launch_bounds__(128)
c2a_filter_warp_kernel(C2aOpt o, const int32_t *__restrict__ len, int32_t nseq,
                       const bwagpu_chain *__restrict__ chains, const bwagpu_seed *__restrict__ seeds,
                       const uint32_t *__restrict__ srt, const int64_t *__restrict__ first_chain,
                       const int64_t *__restrict__ last_chain, bwagpu_alnreg *__restrict__ regs,
                       uint8_t *__restrict__ purged)
{
    const int l = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (l >= nseq) return;
    const int64_t c0 = first_chain[l], c1 = last_chain[l];
    if (c0 < 0) return;
    const int64_t rbase = chains[c0].seed_off;
    const int64_t n_regs = chains[c1].seed_off + chains[c1].n - rbase;
    if (n_regs <= C2A_FILTER_WARP_MIN) return;
    const int l_query = len[l];
    bwagpu_alnreg *av = regs + rbase;
    int lim = 0;
    for (int64_t ci = c0; ci <= c1; ++ci) {
        const bwagpu_chain c = chains[ci];
        const bwagpu_seed *cs = seeds + c.seed_off;
        const uint32_t *srt2 = srt + c.seed_off;
        uint8_t *pg = purged + c.seed_off;
        for (int k = c.n - 1; k >= 0; --k) {
            const bwagpu_seed s = cs[srt2[k]];
            int v = 0;
            bool hit = false;                                   // left the scan through a `break`
            for (int64_t base = 0; base < n_regs && v < lim && !hit; base += 32) {
                const int64_t i = base + lane;
                int cls = 0;                                    // 0 = skipped (purged / past the end), 1 = counts, 2 = stop
                if (i < n_regs) {
                    const bwagpu_alnreg a = av[i];
                    if (!(a.qb == -1 && a.qe == -1)) {
                        cls = 1;
                        if (!(s.rbeg < a.rb || s.rbeg + s.len > a.re || s.qbeg < a.qb || s.qbeg + s.len > a.qe) &&
                            !(s.len - a.seedlen0 > .1 * l_query)) {
                            int qd = s.qbeg - a.qb; int64_t rd = s.rbeg - a.rb;
                            int max_gap = c2a_max_gap(o, qd < rd ? qd : (int)rd);
                            int w = max_gap < a.w ? max_gap : a.w;
                            if (qd - rd < w && rd - qd < w) cls = 2;
                            else {
                                qd = a.qe - (s.qbeg + s.len); rd = a.re - (s.rbeg + s.len);
                                max_gap = c2a_max_gap(o, qd < rd ? qd : (int)rd);
                                w = max_gap < a.w ? max_gap : a.w;
                                if (qd - rd < w && rd - qd < w) cls = 2;
                            }
                        }
                    }
                }
                const unsigned m_cnt = __ballot_sync(0xffffffffu, cls == 1);
                const unsigned m_stop = __ballot_sync(0xffffffffu, cls == 2);
                // entry `lane` is examined only while the count of live regions seen before it is < lim
                const int before = v + __popc(m_cnt & ((1u << lane) - 1u));
                const unsigned m_seen = __ballot_sync(0xffffffffu, before < lim);     // a prefix of the lanes
                const unsigned m_hit = m_stop & m_seen;
                if (m_hit) {
                    const int at = __ffs(m_hit) - 1;
                    v += __popc(m_cnt & ((1u << at) - 1u));
                    hit = true;
                } else {
                    v += __popc(m_cnt & m_seen);
                }
            }
            if (v < lim) {          // (also when the scan ran off the end of the list, as in the reference)
                int vv = c.n;
                if (lane == 0) {
                    for (vv = k + 1; vv < c.n; ++vv) {
                        if (pg[vv]) continue;
                        const bwagpu_seed t = cs[srt2[vv]];
                        if (t.len < s.len * .95) continue;
                        if (s.qbeg <= t.qbeg && s.qbeg + s.len - t.qbeg >= s.len >> 2 && t.qbeg - s.qbeg != t.rbeg - s.rbeg) break;
                        if (t.qbeg <= s.qbeg && t.qbeg + t.len - s.qbeg >= s.len >> 2 && s.qbeg - t.qbeg != s.rbeg - t.rbeg) break;
                    }
                }
                vv = __shfl_sync(0xffffffffu, vv, 0);
                if (vv == c.n) {
                    if (lane == 0) {
                        bwagpu_alnreg *ar = &regs[c.seed_off + (c.n - 1 - k)];
                        ar->qb = -1; ar->qe = -1;
                        pg[k] = 1;
                    }
                    __syncwarp();
                    continue;
                }
            }
            lim++;
        }
    }
}

// the containment filter, one read per thread (:3634-3745)
__global__ void c2a_filter_kernel(C2aOpt o, const int32_t *__restrict__ len, int32_t nseq,
                                  const bwagpu_chain *__restrict__ chains, const bwagpu_seed *__restrict__ seeds,
                                  const uint32_t *__restrict__ srt, const int64_t *__restrict__ first_chain,
                                  const int64_t *__restrict__ last_chain, bwagpu_alnreg *__restrict__ regs,
                                  uint8_t *__restrict__ purged)
{
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= nseq) return;
    const int64_t c0 = first_chain[l], c1 = last_chain[l];
    if (c0 < 0) return;
    const int l_query = len[l];
    const int64_t rbase = chains[c0].seed_off;
    const int64_t n_regs = chains[c1].seed_off + chains[c1].n - rbase;
    if (n_regs > C2A_FILTER_WARP_MIN) return;            // c2a_filter_warp_kernel takes this read
    bwagpu_alnreg *av = regs + rbase;
    int lim = 0;
    for (int64_t ci = c0; ci <= c1; ++ci) {
        const bwagpu_chain c = chains[ci];
        const bwagpu_seed *cs = seeds + c.seed_off;
        const uint32_t *srt2 = srt + c.seed_off;
        uint8_t *pg = purged + c.seed_off;
        for (int k = c.n - 1; k >= 0; --k) {
            const bwagpu_seed s = cs[srt2[k]];
            int64_t i; int v = 0;
            for (i = 0; i < n_regs && v < lim; ++i) {
                const int64_t prb = av[i].rb, pre = av[i].re;
                const int pqb = av[i].qb, pqe = av[i].qe;
                if (pqb == -1 && pqe == -1) continue;
                if (s.rbeg < prb || s.rbeg + s.len > pre || s.qbeg < pqb || s.qbeg + s.len > pqe) { v++; continue; }
                if (s.len - av[i].seedlen0 > .1 * l_query) { v++; continue; }
                const int pw = av[i].w;
                int qd = s.qbeg - pqb; int64_t rd = s.rbeg - prb;
                int max_gap = c2a_max_gap(o, qd < rd ? qd : (int)rd);
                int w = max_gap < pw ? max_gap : pw;
                if (qd - rd < w && rd - qd < w) break;
                qd = pqe - (s.qbeg + s.len); rd = pre - (s.rbeg + s.len);
                max_gap = c2a_max_gap(o, qd < rd ? qd : (int)rd);
                w = max_gap < pw ? max_gap : pw;
                if (qd - rd < w && rd - qd < w) break;
                v++;
            }
            if (v < lim) {
                for (v = k + 1; v < c.n; ++v) {
                    if (pg[v]) continue;
                    const bwagpu_seed t = cs[srt2[v]];
                    if (t.len < s.len * .95) continue;
                    if (s.qbeg <= t.qbeg && s.qbeg + s.len - t.qbeg >= s.len >> 2 && t.qbeg - s.qbeg != t.rbeg - s.rbeg) break;
                    if (t.qbeg <= s.qbeg && t.qbeg + t.len - s.qbeg >= s.len >> 2 && s.qbeg - t.qbeg != s.rbeg - t.rbeg) break;
                }
                if (v == c.n) {
                    bwagpu_alnreg *ar = &regs[c.seed_off + (c.n - 1 - k)];
                    ar->qb = -1; ar->qe = -1;
                    pg[k] = 1;
                    continue;
                }
            }
            lim++;
        }
    }
}

// ---- host side ------------------------------------------------------------------------------
extern "C" int bwagpu_index_load_ref(bwagpu_ctx *ctx, const uint8_t *ref_string, int64_t l_pac, int32_t n_seqs,
                                     const int64_t *contig_off, const int32_t *contig_len, const uint8_t *contig_is_alt)
{
    BWAGPU_OWNS_INDEX(ctx);
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!ref_string || l_pac <= 0 || n_seqs <= 0 || !contig_off || !contig_len)
        BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "index_load_ref: incomplete reference descriptor");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    if (ctx->d_ref_string) { cudaFree(ctx->d_ref_string); cudaFree(ctx->d_contig_off); cudaFree(ctx->d_contig_len); cudaFree(ctx->d_contig_alt); }
    ctx->d_ref_string = ctx->d_contig_off = ctx->d_contig_len = ctx->d_contig_alt = nullptr;
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_ref_string, (size_t)(2 * l_pac) + 64));
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_contig_off, (size_t)n_seqs * 8));
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_contig_len, (size_t)n_seqs * 4));
    BWAGPU_CUDA(ctx, cudaMalloc(&ctx->d_contig_alt, (size_t)n_seqs));
    if (contig_is_alt) BWAGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_contig_alt, contig_is_alt, (size_t)n_seqs, cudaMemcpyHostToDevice, ctx->stream));
    else BWAGPU_CUDA(ctx, cudaMemsetAsync(ctx->d_contig_alt, 0, (size_t)n_seqs, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_ref_string, ref_string, (size_t)(2 * l_pac), cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_contig_off, contig_off, (size_t)n_seqs * 8, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_contig_len, contig_len, (size_t)n_seqs * 4, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    ctx->l_pac = l_pac; ctx->n_contigs = n_seqs;
    return BWAGPU_OK;
}

static int c2a_scan(bwagpu_ctx *ctx, const int64_t *in, int64_t *out, int64_t n)
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

#define C2A_GRID(n) (unsigned)(((n) + 127) / 128), 128, 0, ctx->stream

// one side: route by kind, then per kind two band tries
static int c2a_side(bwagpu_ctx *ctx, const C2aOpt &o, const bwagpu_ext_opt *eo, int left, bwagpu_seqpair *d_pairs,
                    int64_t n, const uint8_t *d_rbuf, int64_t rbytes, const uint8_t *d_qbuf, int64_t qbytes,
                    bwagpu_alnreg *d_regs, const bwagpu_chain *d_chains, const bwagpu_seed *d_seeds,
                    const int32_t *d_len, bwagpu_c2a_result *res)
{
    if (n == 0) return BWAGPU_OK;
    int rc;
    bwagpu_seqpair *bk[3], *retry; int *d_cnt;
    for (int k = 0; k < 3; ++k)
        if ((rc = bwagpu_scratch(ctx, SCR_C2A_BK0 + k, (n + 1) * (int64_t)sizeof(bwagpu_seqpair), (void **)&bk[k]))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_RETRY, (n + 1) * (int64_t)sizeof(bwagpu_seqpair), (void **)&retry))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_CNT, 64, (void **)&d_cnt))) return rc;
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_cnt, 0, 64, ctx->stream));
    c2a_route_kernel<<<C2A_GRID(n)>>>(o, d_pairs, n, bk[0], bk[1], bk[2], d_cnt);
    ctx->launches++;
    int h_cnt[4];
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(h_cnt, d_cnt, 16, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    res->n_pairs_scalar += h_cnt[0]; res->n_pairs16 += h_cnt[1]; res->n_pairs8 += h_cnt[2];
    bwagpu_bsw_params P = { eo->o_del, eo->e_del, eo->o_ins, eo->e_ins, eo->zdrop, left ? eo->pen_clip5 : eo->pen_clip3,
                            eo->a, eo->b };
    static const int kinds[3] = { BWAGPU_BSW_SCALAR, BWAGPU_BSW_16BIT, BWAGPU_BSW_8BIT };
    for (int k = 0; k < 3; ++k) {
        int nump = h_cnt[k];
        bwagpu_seqpair *cur = bk[k], *nxt = retry;
        for (int t = 0; t < 2 && nump > 0; ++t) {
            const int w = o.w << t;
#if C2A_DIRECT
            (void)rbytes; (void)qbytes;
            if ((rc = bwagpu_bsw_scores_direct(ctx, &P, kinds[k], cur, d_rbuf, d_qbuf, nump, w))) return rc;
#else
            if ((rc = bwagpu_bsw_get_scores_device(ctx, &P, kinds[k], cur, d_rbuf, rbytes, d_qbuf, qbytes, nump, w))) return rc;
#endif
            BWAGPU_CUDA(ctx, cudaMemsetAsync(d_cnt + 8, 0, 4, ctx->stream));
            c2a_apply_kernel<<<C2A_GRID(nump)>>>(o, left, w, t == 1, cur, nump, d_regs, d_chains, d_seeds, d_len, nxt, d_cnt + 8);
            ctx->launches++;
            int nr = 0;
            BWAGPU_CUDA(ctx, cudaMemcpyAsync(&nr, d_cnt + 8, 4, cudaMemcpyDeviceToHost, ctx->stream));
            BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
            if (left) res->n_retry_left += nr; else res->n_retry_right += nr;
            nump = nr;
            bwagpu_seqpair *tmp = cur; cur = nxt; nxt = tmp;
        }
    }
    BWAGPU_CUDA(ctx, cudaGetLastError());
    return BWAGPU_OK;
}

extern "C" int bwagpu_chain2aln_device(bwagpu_ctx *ctx, const bwagpu_ext_opt *eo, const uint8_t *d_reads,
                                       const int64_t *d_off, const int32_t *d_len, int32_t nseq,
                                       const bwagpu_chain *d_chains, int64_t n_chains, const bwagpu_seed *d_seeds,
                                       int64_t n_seeds, bwagpu_c2a_result *res)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (!eo || !res) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain2aln: NULL options/result");
    if (!ctx->d_ref_string) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain2aln: no reference loaded (bwagpu_index_load_ref)");
    if (nseq < 0 || n_chains < 0 || n_seeds < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain2aln: negative count");
    if (eo->e_del <= 0 || eo->e_ins <= 0 || eo->w < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain2aln: bad penalties / band");
    if (n_seeds >= (1LL << 31)) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "chain2aln: more than 2^31 seeds in one batch");
    memset(res, 0, sizeof *res);
    if (n_chains == 0 || n_seeds == 0) return BWAGPU_OK;
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    int64_t cells0 = 0;
    if (bwagpu_bsw_cells(ctx, 0, &cells0)) return BWAGPU_ERR_CUDA;
    C2aOpt o = { eo->a, eo->b, eo->o_del, eo->e_del, eo->o_ins, eo->e_ins, eo->pen_clip5, eo->pen_clip3, eo->w, eo->zdrop };
    int rc;
    int64_t *rmax, *first_chain, *last_chain, *sizes, *offs; uint64_t *keys; uint32_t *srt; int32_t *chain_of;
    bwagpu_alnreg *regs; uint8_t *purged;
    const int64_t stride = n_seeds + 1;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_RMAX, n_chains * 16, (void **)&rmax))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_FIRST, (int64_t)(nseq + 1) * 16, (void **)&first_chain))) return rc;
    last_chain = first_chain + nseq + 1;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_KEYS, n_seeds * 8, (void **)&keys))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_SRT, n_seeds * 4, (void **)&srt))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_CHAINOF, n_seeds * 4, (void **)&chain_of))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_SIZES, 6 * stride * 8, (void **)&sizes))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_OFFS, 6 * stride * 8, (void **)&offs))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_REGS, n_seeds * (int64_t)sizeof(bwagpu_alnreg), (void **)&regs))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_PURGED, n_seeds, (void **)&purged))) return rc;
    BWAGPU_CUDA(ctx, cudaMemsetAsync(first_chain, 0xFF, (size_t)(nseq + 1) * 16, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(purged, 0, (size_t)n_seeds, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemsetAsync(sizes, 0, (size_t)(6 * stride * 8), ctx->stream));
    BWAGPU_KERNEL_BEGIN(ctx);
    c2a_chain_kernel<<<C2A_GRID(n_chains)>>>(o, ctx->l_pac, (const int64_t *)ctx->d_contig_off, (const int32_t *)ctx->d_contig_len,
                                             d_len, d_chains, n_chains, d_seeds, rmax, keys, srt, chain_of, first_chain, last_chain);
    c2a_size_kernel<<<C2A_GRID(n_seeds)>>>(d_len, d_chains, d_seeds, srt, chain_of, rmax, n_seeds, sizes, stride);
    ctx->launches++;
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_C2A);
    const int n_scan = C2A_DIRECT ? 2 : 6;            // job counts only, or job counts and the four arena sizes
    for (int k = 0; k < n_scan; ++k)
        if ((rc = c2a_scan(ctx, sizes + k * stride, offs + k * stride, stride))) return rc;
    int64_t tot[6] = {0, 0, 0, 0, 0, 0};
    for (int k = 0; k < n_scan; ++k)
        BWAGPU_CUDA(ctx, cudaMemcpyAsync(&tot[k], offs + k * stride + n_seeds, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    for (int k = 2; k < n_scan; ++k)
        if (tot[k] >= (1LL << 31)) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "chain2aln: extension sequences exceed 2 GiB; split the batch");
    if (tot[0] >= (1LL << 31) || tot[1] >= (1LL << 31)) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "chain2aln: more than 2^31 extension jobs; split the batch");
    bwagpu_seqpair *L, *R; uint8_t *LQ = nullptr, *LR = nullptr, *RQ = nullptr, *RR = nullptr; int *d_err;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_L, (tot[0] + 1) * (int64_t)sizeof(bwagpu_seqpair), (void **)&L))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_R, (tot[1] + 1) * (int64_t)sizeof(bwagpu_seqpair), (void **)&R))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_MISC3, 64, (void **)&d_err))) return rc;
    BWAGPU_CUDA(ctx, cudaMemsetAsync(d_err, 0, 4, ctx->stream));
#if !C2A_DIRECT
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_LQ, tot[2] + 64, (void **)&LQ))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_LR, tot[3] + 64, (void **)&LR))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_RQ, tot[4] + 64, (void **)&RQ))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_RR, tot[5] + 64, (void **)&RR))) return rc;
#endif
    BWAGPU_KERNEL_BEGIN(ctx);
    c2a_fill_kernel<<<C2A_GRID(C2A_DIRECT ? n_seeds : n_seeds * 32)>>>(o, (const uint8_t *)ctx->d_ref_string, d_reads, d_off, d_len, d_chains,
                                           d_seeds, srt, chain_of, rmax, n_seeds, offs, stride, regs, L, R, LQ, LR, RQ, RR, d_err);
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_C2A);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    int h_err = 0;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(&h_err, d_err, 4, cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    if (h_err) BWAGPU_FAIL(ctx, BWAGPU_ERR_RANGE, "chain2aln: read buffer of 4 GiB or more; split the batch");
    res->n_left = tot[0]; res->n_right = tot[1];
    res->seq_bytes = tot[2] + tot[3] + tot[4] + tot[5];
    // direct mode: the jobs read from the chunk's reads and from the reference text themselves
    const uint8_t *text = (const uint8_t *)ctx->d_ref_string;
    if ((rc = c2a_side(ctx, o, eo, 1, L, tot[0], C2A_DIRECT ? text : LR, tot[3], C2A_DIRECT ? d_reads : LQ, tot[2], regs, d_chains, d_seeds, d_len, res))) return rc;
    if (tot[1] > 0) { c2a_h0_kernel<<<C2A_GRID(tot[1])>>>(R, tot[1], regs); ctx->launches++; }
    if ((rc = c2a_side(ctx, o, eo, 0, R, tot[1], C2A_DIRECT ? text : RR, tot[5], C2A_DIRECT ? d_reads : RQ, tot[4], regs, d_chains, d_seeds, d_len, res))) return rc;
    BWAGPU_KERNEL_BEGIN(ctx);
    c2a_filter_kernel<<<C2A_GRID(nseq)>>>(o, d_len, nseq, d_chains, d_seeds, srt, first_chain, last_chain, regs, purged);
    c2a_filter_warp_kernel<<<(unsigned)(((int64_t)nseq * 32 + 127) / 128), 128, 0, ctx->stream>>>(
        o, d_len, nseq, d_chains, d_seeds, srt, first_chain, last_chain, regs, purged);
    ctx->launches++;
    BWAGPU_KERNEL_END(ctx, BWAGPU_KERN_C2A);
    BWAGPU_CUDA(ctx, cudaGetLastError());
    res->d_regs = regs; res->n_regs = n_seeds;
    int64_t cells1 = 0;
    if (bwagpu_bsw_cells(ctx, 0, &cells1)) return BWAGPU_ERR_CUDA;
    res->sw_cells = cells1 - cells0;
    return BWAGPU_OK;
}

extern "C" int bwagpu_chain2aln(bwagpu_ctx *ctx, const bwagpu_ext_opt *eo, const uint8_t *enc_reads, const int64_t *off,
                                const int32_t *len, int32_t nseq, const bwagpu_chain *chains, int64_t n_chains,
                                const bwagpu_seed *seeds, int64_t n_seeds, bwagpu_alnreg *regs)
{
    if (!ctx) return BWAGPU_ERR_ARG;
    if (nseq < 0 || n_chains < 0 || n_seeds < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain2aln: negative count");
    if (n_chains == 0 || n_seeds == 0) return BWAGPU_OK;
    if (!enc_reads || !off || !len || !chains || !seeds || !regs) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain2aln: NULL buffer");
    BWAGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    int64_t bytes = 0, so = 0;
    for (int i = 0; i < nseq; ++i) {
        if (len[i] < 0 || off[i] < 0) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain2aln: read %d has a negative length/offset", i);
        if (off[i] + len[i] > bytes) bytes = off[i] + len[i];
    }
    for (int64_t c = 0; c < n_chains; ++c) {                      // layout contract of the flattened chain_ar
        if (chains[c].seqid < 0 || chains[c].seqid >= nseq || chains[c].n < 0 || chains[c].seed_off != so ||
            (c > 0 && chains[c].seqid < chains[c - 1].seqid) || chains[c].rid < 0 || chains[c].rid >= ctx->n_contigs)
            BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain2aln: chain %lld breaks the layout contract (read order, contiguous seeds, rid)", (long long)c);
        so += chains[c].n;
    }
    if (so != n_seeds) BWAGPU_FAIL(ctx, BWAGPU_ERR_ARG, "chain2aln: chains reference %lld seeds, n_seeds = %lld", (long long)so, (long long)n_seeds);
    uint8_t *d_reads; int64_t *d_off; int32_t *d_len; bwagpu_chain *d_chains; bwagpu_seed *d_seeds;
    int rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_READS, bytes + 16, (void **)&d_reads))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_OFF, (int64_t)nseq * 8, (void **)&d_off))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_SEED_LEN, (int64_t)nseq * 4, (void **)&d_len))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_CHAINS, n_chains * (int64_t)sizeof(bwagpu_chain), (void **)&d_chains))) return rc;
    if ((rc = bwagpu_scratch(ctx, SCR_C2A_SEEDS, n_seeds * (int64_t)sizeof(bwagpu_seed), (void **)&d_seeds))) return rc;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_reads, enc_reads, (size_t)bytes, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_off, off, (size_t)nseq * 8, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_len, len, (size_t)nseq * 4, cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_chains, chains, (size_t)n_chains * sizeof(bwagpu_chain), cudaMemcpyHostToDevice, ctx->stream));
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(d_seeds, seeds, (size_t)n_seeds * sizeof(bwagpu_seed), cudaMemcpyHostToDevice, ctx->stream));
    bwagpu_c2a_result res;
    if ((rc = bwagpu_chain2aln_device(ctx, eo, d_reads, d_off, d_len, nseq, d_chains, n_chains, d_seeds, n_seeds, &res))) return rc;
    BWAGPU_CUDA(ctx, cudaMemcpyAsync(regs, res.d_regs, (size_t)n_seeds * sizeof(bwagpu_alnreg), cudaMemcpyDeviceToHost, ctx->stream));
    BWAGPU_CUDA(ctx, bwagpu_stream_sync(ctx));
    return BWAGPU_OK;
}
