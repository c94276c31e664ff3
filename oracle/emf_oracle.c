/* TEST INFRASTRUCTURE ONLY — CPU restatement of the EMF (exact-match filter) probe, SURVEY.md §8 a12.
 *
 * Follows the reference's find_perfect_match_entry (src/perfect_map.cpp:638-659),
 * __find_perfect_match_entry (:583-629) and seedmatch_further (:528-581) over the table layout of
 * src/perfect.h:81-199, one read at a time, scalar.  Parity status: PINNED — tests/test_oracle_emf.py
 * compares it with the reference's own find_perfect_match_entry (oracle/_ref/libref_api_emf_sse41.so)
 * on tables written by the reference's `perfect-index`, and with tests/golden/emf_golden.npz.
 *
 * The reference compares 8 bytes at a time with byte-swaps (perfect.h:246-300); that is a plain
 * lexicographic comparison of the two oriented strings, which is what is written here.
 */
#include <string.h>
#include "oracle.h"

#define EMF_NO_ENTRY 0xFFFFFFFFu

/* base j of the string at p (length len) read forward (fw != 0) or as its reverse complement */
static inline int emf_base(const uint8_t *p, int len, int fw, int j)
{
    return fw ? p[j] : 3 - (p[len - 1 - j] & 3);
}

/* ____seedcmp, perfect.h:246-312 */
static int emf_cmp(const uint8_t *a, int afw, const uint8_t *b, int bfw, int len)
{
    for (int j = 0; j < len; ++j) {
        int x = emf_base(a, len, afw, j), y = emf_base(b, len, bfw, j);
        if (x != y) return x > y ? 1 : -1;
    }
    return 0;
}

/* __compare_fw_rc, perfect.h:345-351: only the first half of each orientation is looked at */
int orc_emf_fw_less(const uint8_t *seed, int len)
{
    int half = (len + 1) / 2;
    return emf_cmp(seed, 1, seed + (len - half), 0, half) <= 0;
}

static inline uint64_t emf_fmix64(uint64_t k)   /* perfect.h:425-433 (MurmurHash3 finaliser) */
{
    k ^= k >> 33; k *= 0xff51afd7ed558ccdULL;
    k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL;
    k ^= k >> 33;
    return k;
}

/* __get_hash_idx_fw / __get_hash_idx_rc, perfect.h:466-706: XOR of the 64-bit words that hold 32
 * bases each of the oriented string, first base in the top bits; a last partial word is right-aligned */
uint64_t orc_emf_hash(const uint8_t *seed, int len, int fw_less)
{
    uint64_t h = 0, s = 0;
    int j = 0;
    for (; j + 32 <= len; j += 32) {
        s = 0;
        for (int t = 0; t < 32; ++t) s = (s << 2) | (uint64_t)(emf_base(seed, len, fw_less, j + t) & 3);
        h ^= s;
    }
    if (j < len) {
        s = 0;
        for (; j < len; ++j) s = (s << 2) | (uint64_t)(emf_base(seed, len, fw_less, j) & 3);
        h ^= s;
    }
    return emf_fmix64(h);
}

static const orc_emf_entry_t *emf_get(const orc_emf_table_t *pt, uint32_t key)   /* get_seed_entry, perfect.h:749-755 */
{
    return key < pt->num_seed_load ? &pt->seed_table[key] : NULL;
}

/* __seedmatch_further, perfect.h:475-491 (the bounds tests are the reference's, literally) */
static int emf_further(const orc_emf_table_t *pt, uint32_t loc, const uint8_t *seed, int is_rev, int len)
{
    const int L = pt->seed_len, extra = len - L;
    if (!is_rev) {
        if ((uint64_t)loc + (uint64_t)extra >= pt->seq_len) return 0;
        return memcmp(pt->ref_string + loc + L, seed + L, (size_t)extra) == 0;
    }
    if (loc < (uint32_t)extra) return 0;
    for (int j = 0; j < extra; ++j)
        if (pt->ref_string[loc - extra + j] != 3 - (seed[len - 1 - j] & 3)) return 0;
    return 1;
}

/* multi-location list of an entry: GET_MULTI_FW_AND_RC, perfect.h:157-173 */
static void emf_multi(const uint32_t *lt, uint32_t ml, uint32_t *nfw, const uint32_t **fw, uint32_t *nrc, const uint32_t **rc)
{
    if (lt[ml] & 0x80000000u) {
        uint32_t st = lt[ml] & 0x7FFFFFFFu;
        *nfw = lt[st]; *nrc = lt[st + 1]; *fw = lt + st + 2; *rc = lt + st + 2 + *nfw;
    } else {
        *nfw = (lt[ml] >> 16) & 0xFFFF; *nrc = lt[ml] & 0xFFFF; *fw = lt + ml + 1; *rc = lt + ml + 1 + *nfw;
    }
}

int orc_emf_probe1(const orc_emf_table_t *pt, const uint8_t *seed, int len, uint32_t *out_flags, uint32_t *out_loc)
{
    if (!pt || len < pt->seed_len) return 0;                        /* FIND_PERFECT_NO_TABLE  :641-653 */
    int n = 0;
    for (int i = 0; i < len; ++i) n |= seed[i] & 0xC;               /* seed_with_N :631-636 */
    if (n) return 1;                                                /* FIND_PERFECT_WITH_N */
    const int L = pt->seed_len;
    const int fw_less = orc_emf_fw_less(seed, L);
    uint32_t idx = (uint32_t)(orc_emf_hash(seed, L, fw_less) % pt->num_seed_entry);
    const orc_emf_entry_t *ent = emf_get(pt, idx);
    if (!ent || ent->location == EMF_NO_ENTRY || (ent->flags & 2)) return 2;   /* is_hash_matched_entry */
    while (ent) {
        int cmp = emf_cmp(pt->ref_string + ent->location, ent->flags & 1, seed, fw_less, L);
        if (cmp == 0) {
            int is_rev = ((int)(ent->flags & 1) == fw_less) ? 0 : 1;
            uint32_t loc = EMF_NO_ENTRY;
            if (len == L) loc = ent->location;
            else if (emf_further(pt, ent->location, seed, is_rev, len)) loc = ent->location;
            else if (ent->flags >> 2) {
                uint32_t nfw, nrc; const uint32_t *fw, *rc;
                emf_multi(pt->loc_table, ent->flags >> 2, &nfw, &fw, &nrc, &rc);
                for (uint32_t i = 0; i < nfw && loc == EMF_NO_ENTRY; ++i)
                    if (emf_further(pt, fw[i], seed, is_rev, len)) loc = fw[i];
                if (loc == EMF_NO_ENTRY) {
                    is_rev = !is_rev;
                    for (uint32_t i = 0; i < nrc && loc == EMF_NO_ENTRY; ++i)
                        if (emf_further(pt, rc[i], seed, is_rev, len)) loc = rc[i];
                }
            }
            if (loc == EMF_NO_ENTRY) return 5;                      /* FIND_PERFECT_SEED_ONLY_MATCHED */
            *out_loc = loc;
            *out_flags = is_rev ? (ent->flags | 2u | 1u) : ((ent->flags & ~2u) | 1u);
            return is_rev ? 4 : 3;                                  /* RC_MATCHED : FW_MATCHED */
        }
        ent = emf_get(pt, cmp > 0 ? ent->left : ent->right);
    }
    return 2;                                                       /* FIND_PERFECT_NOT_MATCHED */
}

void orc_emf_probe(const orc_emf_table_t *pt, const uint8_t *reads, const int64_t *off, const int32_t *len,
                   int32_t nseq, int32_t *ret, uint32_t *perfect /* [nseq][2] = flags, location */)
{
    for (int32_t i = 0; i < nseq; ++i) {
        perfect[2 * i] = perfect[2 * i + 1] = 0;                    /* seq->perfect.exist starts at 0 */
        ret[i] = orc_emf_probe1(pt, reads + off[i], len[i], &perfect[2 * i], &perfect[2 * i + 1]);
    }
}
