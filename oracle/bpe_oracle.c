/* bpe_oracle.c — plain-C CPU restatement of the reference's BPE hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may link or call this file, and
 * only as the checker.  The product (bpe_transformer_b200/) never falls back to it.
 *
 * What it restates (reference = milasd/BPE-Transformer, paths relative to /root/reference):
 *   orc_scan     the GPT-2 split pattern (bpe_transformer/settings.py:8) as regex.finditer applies it
 *                (tokenization/preprocessing/pretokenization.py:238-252) after regex.split on the special
 *                tokens (:211-235) — written as the sequential scanner of SURVEY.md Appendix A.1, with the
 *                code-point classes of the installed `regex` module (include/bpe_unicode_tables.h).
 *   orc_count    the pretoken Counter (pretokenization.py:171-208).
 *   orc_train    BPETrainer._initialize_pairs_cache / _merge_tokens / _merge_pair / _get_adjacency_pairs
 *                (tokenization/bpe_trainer.py:169-426) in the heap-free form of SURVEY.md Appendix A.2:
 *                argmax over (count, (bytes_a, bytes_b)) among pairs with count > 0 that still occur.
 *   orc_encode   BPETokenizer._encode_pretokens / _find_pair_in_merges (tokenization/bpe_tokenizer.py:209-290).
 *
 * The scanner is deliberately a different formulation from the CUDA kernel (sequential token-by-token
 * here, a local per-byte rule there) and from oracle/bpe_ref_port.py (which calls `regex` itself), so
 * the three check one another.  Parity pinned: tests/test_oracle_pinned.py.
 *
 * Build: gcc -O2 -shared -fPIC -I include -o oracle/liboracle.so oracle/bpe_oracle.c
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "bpe_unicode_tables.h"

#define ORC_OK 0
#define ORC_ERR_UTF8 1
#define ORC_ERR_NOMEM 2
#define ORC_ERR_MISSING_ID 3   /* a merge whose joined bytes have no id was applied (reference: KeyError) */
#define ORC_ERR_CAP 4

#define ORC_CUT_AFTER_NEWLINE 1u

enum { C_O = 0, C_L = 1, C_N = 2, C_S = 3 };

static int cp_class(uint32_t cp) {
    uint32_t b = cp >> BPE_UC_BLOCK_SHIFT;
    if (b >= BPE_UC_N_STAGE1) return C_O;
    uint32_t w = BPE_UC_STAGE2[BPE_UC_STAGE1[b] * BPE_UC_WORDS_PER_BLOCK + ((cp & ((1u << BPE_UC_BLOCK_SHIFT) - 1)) >> 4)];
    return (int)((w >> (2 * (cp & 15))) & 3);
}

/* Strict UTF-8 (what CPython's decoder accepts).  Returns the sequence length, 0 if invalid. */
static int utf8_at(const uint8_t *t, uint64_t i, uint64_t end, uint32_t *cp) {
    uint8_t b0 = t[i];
    if (b0 < 0x80) { *cp = b0; return 1; }
    if (b0 < 0xC2) return 0;
    if (b0 < 0xE0) {
        if (i + 1 >= end || (t[i + 1] & 0xC0) != 0x80) return 0;
        *cp = ((uint32_t)(b0 & 0x1F) << 6) | (t[i + 1] & 0x3F);
        return 2;
    }
    if (b0 < 0xF0) {
        if (i + 2 >= end || (t[i + 1] & 0xC0) != 0x80 || (t[i + 2] & 0xC0) != 0x80) return 0;
        uint32_t c = ((uint32_t)(b0 & 0x0F) << 12) | ((uint32_t)(t[i + 1] & 0x3F) << 6) | (t[i + 2] & 0x3F);
        if (c < 0x800 || (c >= 0xD800 && c <= 0xDFFF)) return 0;
        *cp = c;
        return 3;
    }
    if (b0 < 0xF5) {
        if (i + 3 >= end || (t[i + 1] & 0xC0) != 0x80 || (t[i + 2] & 0xC0) != 0x80 || (t[i + 3] & 0xC0) != 0x80) return 0;
        uint32_t c = ((uint32_t)(b0 & 0x07) << 18) | ((uint32_t)(t[i + 1] & 0x3F) << 12) | ((uint32_t)(t[i + 2] & 0x3F) << 6) | (t[i + 3] & 0x3F);
        if (c < 0x10000 || c > 0x10FFFF) return 0;
        *cp = c;
        return 4;
    }
    return 0;
}

static inline void set_bit(uint32_t *m, uint64_t i) { m[i >> 5] |= 1u << (i & 31); }
static inline int get_bit(const uint32_t *m, uint64_t i) { return (m[i >> 5] >> (i & 31)) & 1; }

/* One segment [lo, hi) that contains no special token and no cut: SURVEY.md Appendix A.1 rules 1-3. */
static int scan_segment(const uint8_t *t, uint64_t lo, uint64_t hi, uint32_t *start) {
    uint64_t i = lo;
    while (i < hi) {
        set_bit(start, i);
        /* rule 1: contraction, only at a token start */
        if (t[i] == '\'') {
            if (i + 1 < hi && (t[i + 1] == 's' || t[i + 1] == 'd' || t[i + 1] == 'm' || t[i + 1] == 't')) { i += 2; continue; }
            if (i + 2 < hi && ((t[i + 1] == 'l' && t[i + 2] == 'l') || (t[i + 1] == 'v' && t[i + 2] == 'e') || (t[i + 1] == 'r' && t[i + 2] == 'e'))) { i += 3; continue; }
        }
        /* rule 2: optional single U+0020, then a run of one class among letter / number / other */
        uint64_t j = (t[i] == ' ') ? i + 1 : i;
        if (j < hi) {
            uint32_t cp;
            int n = utf8_at(t, j, hi, &cp);
            if (!n) return ORC_ERR_UTF8;
            int c = cp_class(cp);
            if (c != C_S) {
                uint64_t k = j + n;
                while (k < hi) {
                    int m = utf8_at(t, k, hi, &cp);
                    if (!m) return ORC_ERR_UTF8;
                    if (cp_class(cp) != c) break;
                    k += m;
                }
                i = k;
                continue;
            }
        }
        /* rule 3: whitespace run; leave its last character behind when something non-white follows */
        uint64_t k = i, last = 0, count = 0;
        while (k < hi) {
            uint32_t cp;
            int m = utf8_at(t, k, hi, &cp);
            if (!m) return ORC_ERR_UTF8;
            if (cp_class(cp) != C_S) break;
            last = (uint64_t)m; k += m; count++;
        }
        i = (k == hi || count == 1) ? k : k - last;
    }
    return ORC_OK;
}

/* A piece [lo, hi) between two cuts: split on special tokens (priority order = the order given, which
 * the host makes longest-first like pretokenization.py:228), leftmost match, non-overlapping. */
static int scan_piece(const uint8_t *t, uint64_t lo, uint64_t hi, const uint8_t *sp, const uint32_t *sp_off, int n_sp, uint32_t *start) {
    uint64_t seg = lo, p = lo;
    while (p < hi) {
        int hit = -1;
        for (int s = 0; s < n_sp && hit < 0; s++) {
            uint32_t L = sp_off[s + 1] - sp_off[s];
            if (L && p + L <= hi && memcmp(t + p, sp + sp_off[s], L) == 0) hit = s;
        }
        if (hit < 0) { p++; continue; }
        int rc = scan_segment(t, seg, p, start);
        if (rc) return rc;
        set_bit(start, p);                       /* the special token itself is a token */
        p += sp_off[hit + 1] - sp_off[hit];
        seg = p;
    }
    return scan_segment(t, seg, hi, start);
}

/* Token-start bitmask of text[0, n).  `cuts` = sorted positions with a segment boundary in front of them
 * (encode_iterable's newline cuts, bpe_tokenizer.py:368-379); flag CUT_AFTER_NEWLINE = cut after every
 * '\n'.  start must hold (n + 31) / 32 words; it is cleared here. */
int orc_scan(const uint8_t *t, uint64_t n, const uint8_t *sp, const uint32_t *sp_off, int n_sp,
             const uint64_t *cuts, uint64_t n_cuts, uint32_t flags, uint32_t *start) {
    memset(start, 0, ((n + 31) / 32) * sizeof(uint32_t));
    uint64_t lo = 0, ci = 0;
    while (lo < n) {
        uint64_t hi = n;
        while (ci < n_cuts && cuts[ci] <= lo) ci++;
        if (ci < n_cuts && cuts[ci] < hi) hi = cuts[ci];
        if (flags & ORC_CUT_AFTER_NEWLINE) {
            const uint8_t *nl = (const uint8_t *)memchr(t + lo, '\n', hi - lo);
            if (nl) hi = (uint64_t)(nl - t) + 1;
        }
        int rc = scan_piece(t, lo, hi, sp, sp_off, n_sp, start);
        if (rc) return rc;
        lo = hi;
    }
    return ORC_OK;
}

/* ---------------------------------------------------------------------------------------------- */
/* pretoken Counter                                                                                */
/* ---------------------------------------------------------------------------------------------- */
static uint64_t fnv(const uint8_t *p, uint64_t n) {
    uint64_t h = 1469598103934665603ull;
    for (uint64_t i = 0; i < n; i++) { h ^= p[i]; h *= 1099511628211ull; }
    return h ? h : 1;
}

static int is_special(const uint8_t *t, uint64_t at, uint64_t len, const uint8_t *sp, const uint32_t *sp_off, int n_sp) {
    for (int s = 0; s < n_sp; s++)
        if (sp_off[s + 1] - sp_off[s] == len && len && memcmp(t + at, sp + sp_off[s], len) == 0) return s;
    return -1;
}

/* Distinct pretokens of text given its start mask, special tokens left out (training mode), in order of
 * first occurrence.  Returns the number found (or -1 when cap is too small); out_off/out_len/out_cnt
 * hold cap entries. */
int64_t orc_count(const uint8_t *t, uint64_t n, const uint32_t *start, const uint8_t *sp, const uint32_t *sp_off, int n_sp,
                  uint64_t *out_off, uint32_t *out_len, uint64_t *out_cnt, uint64_t cap) {
    uint64_t tcap = 1024;
    while (tcap < cap * 2) tcap <<= 1;
    int64_t *slot = (int64_t *)malloc(tcap * sizeof(int64_t));
    uint64_t *hs = (uint64_t *)malloc(tcap * sizeof(uint64_t));
    if (!slot || !hs) return -2;
    memset(slot, 0xFF, tcap * sizeof(int64_t));
    uint64_t found = 0, i = 0;
    while (i < n) {
        uint64_t e = i + 1;
        while (e < n && !get_bit(start, e)) e++;
        if (get_bit(start, i) && is_special(t, i, e - i, sp, sp_off, n_sp) < 0) {
            uint64_t h = fnv(t + i, e - i), k = h & (tcap - 1);
            for (;;) {
                if (slot[k] < 0) {
                    if (found == cap) { free(slot); free(hs); return -1; }
                    slot[k] = (int64_t)found; hs[k] = h;
                    out_off[found] = i; out_len[found] = (uint32_t)(e - i); out_cnt[found] = 1; found++;
                    break;
                }
                uint64_t u = (uint64_t)slot[k];
                if (hs[k] == h && out_len[u] == e - i && memcmp(t + out_off[u], t + i, e - i) == 0) { out_cnt[u]++; break; }
                k = (k + 1) & (tcap - 1);
            }
        }
        i = e;
    }
    free(slot); free(hs);
    return (int64_t)found;
}

/* ---------------------------------------------------------------------------------------------- */
/* trainer                                                                */
/* ---------------------------------------------------------------------------------------------- */
typedef struct { uint64_t key; int64_t cnt; uint8_t dead; } pair_t;
typedef struct { pair_t *e; uint64_t cap, used; } pmap_t;

static pair_t *pm_get(pmap_t *m, uint32_t a, uint32_t b) {
    uint64_t key = (((uint64_t)a << 32) | b) + 1;     /* 0 = empty */
    uint64_t k = (key * 0x9E3779B97F4A7C15ull) >> 20 & (m->cap - 1);
    for (;;) {
        if (m->e[k].key == key) return &m->e[k];
        if (m->e[k].key == 0) { m->e[k].key = key; m->e[k].cnt = 0; m->e[k].dead = 0; m->used++; return &m->e[k]; }
        k = (k + 1) & (m->cap - 1);
    }
}
static int pm_grow(pmap_t *m) {
    pmap_t big = {(pair_t *)calloc(m->cap * 2, sizeof(pair_t)), m->cap * 2, 0};
    if (!big.e) return 1;
    for (uint64_t k = 0; k < m->cap; k++)
        if (m->e[k].key) { pair_t *d = pm_get(&big, (uint32_t)((m->e[k].key - 1) >> 32), (uint32_t)(m->e[k].key - 1)); d->cnt = m->e[k].cnt; d->dead = m->e[k].dead; }
    free(m->e); *m = big; return 0;
}

static int bytes_cmp(const uint8_t *a, uint32_t na, const uint8_t *b, uint32_t nb) {
    uint32_t n = na < nb ? na : nb;
    int c = memcmp(a, b, n);
    if (c) return c;
    return (na > nb) - (na < nb);
}

/* words: CSR (w_off[n_words + 1], sym[]) of initial symbol ids, count per word.  Token byte strings of
 * the initial vocabulary (ids 0 .. first_new_id - 1) come in tok_blob / tok_off.  Learns merges until the
 * vocabulary holds vocab_size entries or nothing qualifies.  Exact ties on (count, bytes, bytes) between
 * different id pairs are unpinned in the reference (heap order); here the larger (a, b) id pair wins. */
int orc_train(uint64_t n_words, const uint64_t *w_off, const uint32_t *sym_in, const uint64_t *w_cnt,
              const uint8_t *tok_blob, const uint32_t *tok_off, uint32_t first_new_id, uint32_t vocab_size,
              uint32_t *merge_a, uint32_t *merge_b, uint32_t *n_merges_out) {
    uint64_t n_sym = w_off[n_words];
    uint32_t *sym = (uint32_t *)malloc((n_sym + 1) * sizeof(uint32_t));
    uint32_t *len = (uint32_t *)malloc((n_words + 1) * sizeof(uint32_t));
    uint32_t *tmp = NULL; uint32_t tmp_cap = 0;
    uint32_t max_tok = vocab_size > first_new_id ? vocab_size : first_new_id;
    uint64_t blob_cap = (uint64_t)tok_off[first_new_id] + 64;
    uint8_t *blob = (uint8_t *)malloc(blob_cap);
    uint64_t *toff = (uint64_t *)malloc((max_tok + 1) * sizeof(uint64_t));
    pmap_t pm = {(pair_t *)calloc(1 << 16, sizeof(pair_t)), 1 << 16, 0};
    if (!sym || !len || !blob || !toff || !pm.e) return ORC_ERR_NOMEM;
    memcpy(sym, sym_in, n_sym * sizeof(uint32_t));
    memcpy(blob, tok_blob, tok_off[first_new_id]);
    for (uint32_t k = 0; k <= first_new_id; k++) toff[k] = tok_off[k];
    for (uint64_t w = 0; w < n_words; w++) {
        len[w] = (uint32_t)(w_off[w + 1] - w_off[w]);
        if (len[w] > tmp_cap) tmp_cap = len[w];
        for (uint32_t k = 0; k + 1 < len[w]; k++) {               /* bpe_trainer.py:188-199 */
            if (pm.used * 2 > pm.cap && pm_grow(&pm)) return ORC_ERR_NOMEM;
            pm_get(&pm, sym[w_off[w] + k], sym[w_off[w] + k + 1])->cnt += (int64_t)w_cnt[w];
        }
    }
    tmp = (uint32_t *)malloc((tmp_cap + 1) * sizeof(uint32_t));
    uint32_t next_id = first_new_id, n_merges = 0;
    while (next_id < vocab_size) {
        /* argmax over live pairs with count > 0 (bpe_trainer.py:36-67 ordering) */
        pair_t *best = NULL; uint32_t ba = 0, bb = 0;
        for (uint64_t k = 0; k < pm.cap; k++) {
            pair_t *p = &pm.e[k];
            if (!p->key || p->dead || p->cnt <= 0) continue;
            uint32_t a = (uint32_t)((p->key - 1) >> 32), b = (uint32_t)(p->key - 1);
            int better;
            if (!best) better = 1;
            else if (p->cnt != best->cnt) better = p->cnt > best->cnt;
            else {
                int c = bytes_cmp(blob + toff[a], (uint32_t)(toff[a + 1] - toff[a]), blob + toff[ba], (uint32_t)(toff[ba + 1] - toff[ba]));
                if (!c) c = bytes_cmp(blob + toff[b], (uint32_t)(toff[b + 1] - toff[b]), blob + toff[bb], (uint32_t)(toff[bb + 1] - toff[bb]));
                if (!c) c = (a > ba) - (a < ba);
                if (!c) c = (b > bb) - (b < bb);
                better = c > 0;
            }
            if (better) { best = p; ba = a; bb = b; }
        }
        if (!best) break;
        uint64_t best_key = best->key;
        int merged_any = 0;
        for (uint64_t w = 0; w < n_words; w++) {
            uint32_t L = len[w];
            uint32_t *s = sym + w_off[w];
            int has = 0;
            for (uint32_t k = 0; k + 1 < L; k++) if (s[k] == ba && s[k + 1] == bb) { has = 1; break; }
            if (!has) continue;
            merged_any = 1;
            int64_t c = (int64_t)w_cnt[w];
            uint32_t m = 0, j = 0;
            while (j < L) {                                        /* bpe_trainer.py:400-420 */
                if (j + 1 < L && s[j] == ba && s[j + 1] == bb) {
                    if (pm.used * 2 + 8 > pm.cap && pm_grow(&pm)) return ORC_ERR_NOMEM;
                    if (j >= 1) pm_get(&pm, s[j - 1], s[j])->cnt -= c;          /* neighbours from the ORIGINAL word */
                    if (j + 2 < L) pm_get(&pm, s[j + 1], s[j + 2])->cnt -= c;
                    tmp[m++] = next_id; j += 2;
                } else tmp[m++] = s[j++];
            }
            for (uint32_t k = 0; k < m; k++) {                     /* bpe_trainer.py:349-357 */
                if (tmp[k] != next_id) continue;
                if (pm.used * 2 + 8 > pm.cap && pm_grow(&pm)) return ORC_ERR_NOMEM;
                if (k >= 1) pm_get(&pm, tmp[k - 1], tmp[k])->cnt += c;
                if (k + 1 < m) pm_get(&pm, tmp[k], tmp[k + 1])->cnt += c;
            }
            memcpy(s, tmp, m * sizeof(uint32_t));
            len[w] = m;
        }
        /* the winner never occurs again: retire it (the reference pops it from the heap) */
        pm_get(&pm, (uint32_t)((best_key - 1) >> 32), (uint32_t)(best_key - 1))->dead = 1;
        if (!merged_any) continue;                                 /* count > 0 but no occurrence: no merge */
        uint64_t la = toff[ba + 1] - toff[ba], lb = toff[bb + 1] - toff[bb];
        if (toff[next_id] + la + lb > blob_cap) {
            blob_cap = (toff[next_id] + la + lb) * 2;
            blob = (uint8_t *)realloc(blob, blob_cap);
            if (!blob) return ORC_ERR_NOMEM;
        }
        memcpy(blob + toff[next_id], blob + toff[ba], la);
        memcpy(blob + toff[next_id] + la, blob + toff[bb], lb);
        toff[next_id + 1] = toff[next_id] + la + lb;
        merge_a[n_merges] = ba; merge_b[n_merges] = bb; n_merges++;
        next_id++;
    }
    *n_merges_out = n_merges;
    free(sym); free(len); free(tmp); free(blob); free(toff); free(pm.e);
    return ORC_OK;
}

/* ---------------------------------------------------------------------------------------------- */
/* encoder                                                                */
/* ---------------------------------------------------------------------------------------------- */
typedef struct { uint64_t key; uint32_t rank, merged; } mrg_t;

/* Encodes text given its start mask.  byte_to_id[256]; merge table rows (a, b, merged) in rank order over
 * canonical ids (the host resolves bytes -> id exactly as bpe_tokenizer.py:55,85 do); merged == 0xFFFFFFFF
 * marks "joined bytes not in the vocabulary".  special_id[s] = id of special s.  Writes ids to out. */
int orc_encode(const uint8_t *t, uint64_t n, const uint32_t *start, const uint8_t *sp, const uint32_t *sp_off, int n_sp,
               const uint32_t *special_id, const uint32_t *byte_to_id, const uint32_t *m_a, const uint32_t *m_b,
               const uint32_t *m_id, uint32_t n_merges, uint32_t *out, uint64_t cap, uint64_t *n_out) {
    uint64_t tcap = 1024;
    while (tcap < (uint64_t)n_merges * 2 + 2) tcap <<= 1;
    mrg_t *tab = (mrg_t *)calloc(tcap, sizeof(mrg_t));
    if (!tab) return ORC_ERR_NOMEM;
    for (uint32_t r = 0; r < n_merges; r++) {                      /* later duplicate wins, like the dict */
        uint64_t key = (((uint64_t)m_a[r] << 32) | m_b[r]) + 1, k = (key * 0x9E3779B97F4A7C15ull) >> 20 & (tcap - 1);
        while (tab[k].key && tab[k].key != key) k = (k + 1) & (tcap - 1);
        tab[k].key = key; tab[k].rank = r; tab[k].merged = m_id[r];
    }
    uint64_t w = 0, i = 0;
    uint32_t *ids = NULL; uint64_t ids_cap = 0;
    int rc = ORC_OK;
    while (i < n && rc == ORC_OK) {
        uint64_t e = i + 1;
        while (e < n && !get_bit(start, e)) e++;
        if (!get_bit(start, i)) { i = e; continue; }
        int s = is_special(t, i, e - i, sp, sp_off, n_sp);
        if (s >= 0) {
            if (w >= cap) { rc = ORC_ERR_CAP; break; }
            out[w++] = special_id[s]; i = e; continue;
        }
        uint64_t L = e - i;
        if (L > ids_cap) { ids_cap = L * 2; ids = (uint32_t *)realloc(ids, ids_cap * sizeof(uint32_t)); }
        for (uint64_t k = 0; k < L; k++) ids[k] = byte_to_id[t[i + k]];
        while (L >= 2) {                                           /* bpe_tokenizer.py:257-290 */
            uint32_t best = 0xFFFFFFFFu, merged = 0; uint64_t at = 0;
            for (uint64_t k = 0; k + 1 < L; k++) {
                uint64_t key = (((uint64_t)ids[k] << 32) | ids[k + 1]) + 1, q = (key * 0x9E3779B97F4A7C15ull) >> 20 & (tcap - 1);
                while (tab[q].key && tab[q].key != key) q = (q + 1) & (tcap - 1);
                if (tab[q].key && tab[q].rank < best) { best = tab[q].rank; merged = tab[q].merged; at = k; }
            }
            if (best == 0xFFFFFFFFu) break;
            if (merged == 0xFFFFFFFFu) { rc = ORC_ERR_MISSING_ID; break; }
            ids[at] = merged;
            memmove(ids + at + 1, ids + at + 2, (L - at - 2) * sizeof(uint32_t));
            L--;
        }
        if (w + L > cap) { rc = ORC_ERR_CAP; break; }
        memcpy(out + w, ids, L * sizeof(uint32_t));
        w += L; i = e;
    }
    *n_out = w;
    free(ids); free(tab);
    return rc;
}
