// TEST INFRASTRUCTURE ONLY — never linked into the product library.
//
// extern "C" tap into the UNMODIFIED reference's ERT seeding (src/ertseeding.cpp), linked against the
// same reference objects as ref_api_wrap.cpp:
//   ref_ert_open / ref_ert_close : maps <prefix>.kmer_table / .mlt_table / .0123 as FMI_search::load_index
//                                  with useErt does (src/FMI_search.cpp:1190-1245); the 8 GiB k-mer table is
//                                  mmap'ed so only the touched pages are read
//   ref_ert_seed                 : the per-read seeding body of mem_kernel1_core_ert (src/bwamem.cpp:1124-1193):
//                                  get_seeds[_prefix] -> reseed[_prefix] -> last -> ks_introsort(mem_smem_sort_lt),
//                                  calling the reference's own functions; hits are reported as the coordinates
//                                  mem_chain_new derives from them (src/bwamem.cpp:991-998)
//   ref_ert_tables               : the reference's leaf_table / code_table (src/ertseeding.cpp:25-430)
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <cstdio>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include "bwamem.h"
#include "ertseeding.h"

extern unsigned char leaf_table[256 * 3][4];
extern unsigned char code_table[256 * 3][4];

#define smem_lt_2(a, b) ((a).start == (b).start ? (a).end < (b).end : (a).start < (b).start)
KSORT_INIT(tap_smem_sort_lt, mem_t, smem_lt_2)          // = KSORT_INIT(mem_smem_sort_lt, ...) src/bwamem.cpp:72-74

struct ErtTap {
    int borrowed;                 // tables are caller-owned arrays (ref_ert_open_mem)
    uint64_t *kmer; size_t kmer_bytes;
    uint8_t *mlt; size_t mlt_bytes;
    uint8_t *ref; size_t ref_bytes;
    bntseq_t bns;
};

static void *map_file(const char *prefix, const char *ext, size_t *bytes)
{
    char path[4096];
    snprintf(path, sizeof path, "%s%s", prefix, ext);
    int fd = open(path, O_RDONLY);
    if (fd < 0) return nullptr;
    struct stat st;
    fstat(fd, &st);
    void *p = mmap(nullptr, st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    *bytes = st.st_size;
    return p == MAP_FAILED ? nullptr : p;
}

extern "C" {

struct tap_mem { int32_t rid, start, end, hitcount; int64_t hit_off; };

void *ref_ert_open(const char *prefix)
{
    ErtTap *t = (ErtTap *)calloc(1, sizeof(ErtTap));
    t->kmer = (uint64_t *)map_file(prefix, ".kmer_table", &t->kmer_bytes);
    t->mlt = (uint8_t *)map_file(prefix, ".mlt_table", &t->mlt_bytes);
    t->ref = (uint8_t *)map_file(prefix, ".0123", &t->ref_bytes);
    if (!t->kmer || !t->mlt || !t->ref) { free(t); return nullptr; }
    t->bns.l_pac = (int64_t)(t->ref_bytes / 2);
    return t;
}

// same over caller-owned arrays: dense k-mer table, tree blob, .0123 text (kept alive by the caller)
void *ref_ert_open_mem(uint64_t *kmer, uint8_t *mlt, int64_t mlt_bytes, uint8_t *ref, int64_t ref_bytes)
{
    ErtTap *t = (ErtTap *)calloc(1, sizeof(ErtTap));
    t->borrowed = 1;
    t->kmer = kmer; t->mlt = mlt; t->mlt_bytes = (size_t)mlt_bytes; t->ref = ref; t->ref_bytes = (size_t)ref_bytes;
    t->bns.l_pac = ref_bytes / 2;
    return t;
}

void ref_ert_close(void *h)
{
    ErtTap *t = (ErtTap *)h;
    if (t->borrowed) { free(t); return; }
    munmap(t->kmer, t->kmer_bytes); munmap(t->mlt, t->mlt_bytes); munmap(t->ref, t->ref_bytes);
    free(t);
}

const uint64_t *ref_ert_kmer(void *h) { return ((ErtTap *)h)->kmer; }
const uint8_t *ref_ert_mlt(void *h, int64_t *bytes) { *bytes = (int64_t)((ErtTap *)h)->mlt_bytes; return ((ErtTap *)h)->mlt; }
const unsigned char *ref_ert_tables(int which) { return which ? &code_table[0][0] : &leaf_table[0][0]; }

// reads: concatenated codes (0-3, >= 4 = N), off/len per read.  Per read the sorted MEMs go to mems[]
// (hit_off = index of the MEM's first coordinate in hits[]) and the coordinates of all their hits to hits[].
// Returns 0, or -1 when a capacity was too small (n_out then holds the sizes needed).
int ref_ert_seed(void *h, int min_seed_len, float split_factor, int split_width, int max_mem_intv,
                 const uint8_t *reads, const int64_t *off, const int32_t *len, int nseq,
                 tap_mem *mems, int64_t mem_cap, int64_t *hits_out, int64_t hit_cap, int64_t *mem_off, int64_t *n_out)
{
    ErtTap *t = (ErtTap *)h;
    mem_v smems; kv_init(smems);
    u64v hits; kv_init(hits);
    kv_resize(mem_t, smems, 4096);
    kv_resize(uint64_t, hits, 1 << 16);
    int64_t nm = 0, nh = 0;
    const int split_len = (int)(min_seed_len * split_factor + .499);
    for (int l = 0; l < nseq; ++l) {
        mem_off[l] = nm;
        const int n = len[l];
        if (n > READ_LEN) continue;
        uint8_t seq[READ_LEN + 1], rcq[READ_LEN + 1];
        int hasN = 0;
        for (int i = 0; i < n; ++i) {
            seq[i] = reads[off[l] + i] < 4 ? reads[off[l] + i] : 4;
            hasN |= seq[i] == 4;
            rcq[n - i - 1] = seq[i] < 4 ? 3 - seq[i] : 4;
        }
        seq[n] = 0; rcq[n] = 0;
        index_aux_t iaux;
        memset(&iaux, 0, sizeof iaux);
        iaux.kmer_offsets = t->kmer; iaux.mlt_table = t->mlt; iaux.bns = &t->bns; iaux.pac = nullptr;
        iaux.ref_string = t->ref;
        read_aux_t raux;
        memset(&raux, 0, sizeof raux);
        raux.min_seed_len = min_seed_len;
        raux.l_seq = n;
        raux.read_name = (char *)"r";
        raux.unpacked_queue_buf = seq;
        raux.unpacked_rc_queue_buf = rcq;
        smems.n = 0; hits.n = 0;
        if (hasN) get_seeds(&iaux, &raux, &smems, &hits); else get_seeds_prefix(&iaux, &raux, &smems, &hits);
        const int old_n = (int)smems.n;
        for (int i = 0; i < old_n; ++i) {
            const int qbeg = smems.a[i].start, qend = smems.a[i].end;
            if ((qend - qbeg) < split_len || smems.a[i].hitcount > split_width) continue;
            pivot_t pt = smems.a[i].pt;
            if (hasN) reseed(&iaux, &raux, &smems, (qbeg + qend) >> 1, smems.a[i].hitcount + 1, &pt, &hits);
            else reseed_prefix(&iaux, &raux, &smems, (qbeg + qend) >> 1, smems.a[i].hitcount + 1, &pt, &hits);
        }
        last(&iaux, &raux, &smems, max_mem_intv, &hits);
        ks_introsort(tap_smem_sort_lt, smems.n, smems.a);
        for (size_t i = 0; i < smems.n; ++i) {
            const mem_t *p = &smems.a[i];
            if (nm < mem_cap) { mems[nm].rid = l; mems[nm].start = p->start; mems[nm].end = p->end;
                                mems[nm].hitcount = p->hitcount; mems[nm].hit_off = nh; }
            ++nm;
            const int slen = p->end - p->start;
            for (int k = 0; k < p->hitcount; ++k, ++nh) {
                if (nh >= hit_cap) continue;
                const uint64_t hv = hits.a[p->hitbeg + k];
                hits_out[nh] = (p->forward || p->fetch_leaves) ? (int64_t)hv
                             : (int64_t)((t->bns.l_pac << 1) - (hv + slen - p->end_correction));
            }
        }
    }
    mem_off[nseq] = nm;
    n_out[0] = nm; n_out[1] = nh;
    kv_destroy(smems); kv_destroy(hits);
    return (nm <= mem_cap && nh <= hit_cap) ? 0 : -1;
}

}  // extern "C"
