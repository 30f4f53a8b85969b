/* Deterministic synthetic corpus generator (there is no network for TinyStories / OpenWebText).
 *
 * Integer-only arithmetic (no libm) so the same (style, seed, bytes) gives the same file on any box.
 * Shapes follow SURVEY.md §8(d):
 *   style 0 "tinystories": documents of 8-25 short sentences, Zipf-distributed words from a ~24k lexicon,
 *           contractions, quotes, repeated punctuation, "\n\n" paragraph breaks, ~0.3 % non-ASCII,
 *           every document terminated by "\n<|endoftext|>\n".
 *   style 1 "owt": heavier tail (fresh rare words, URLs, numbers, mixed case, Cyrillic / CJK snippets,
 *           long whitespace and punctuation runs), documents separated by "<|endoftext|>".
 * Every document is generated from (seed, document index) alone, so the output does not depend on the
 * number of threads.  The file ends at a document end, at the first document that reaches `target`.
 *
 * Build: gcc -O2 -shared -fPIC -pthread -o tools/libgencorpus.so tools/gen_corpus.c
 *        gcc -O2 -pthread -DGEN_MAIN -o tools/gen_corpus tools/gen_corpus.c
 */
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef struct { uint64_t s; } rng_t;
static inline uint64_t rng_next(rng_t *r) {
    uint64_t z = (r->s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline uint32_t rng_below(rng_t *r, uint32_t n) { return (uint32_t)(((rng_next(r) >> 32) * (uint64_t)n) >> 32); }

/* ---- lexicon ---------------------------------------------------------------------------------- */
#define LEX_MAX 200000
typedef struct {
    int n;
    char *pool;
    uint32_t *off;      /* n+1 */
    uint64_t *cum;      /* cumulative Zipf weights, n */
    uint64_t total;
} lex_t;

static const char *ONSET[] = {"b","c","d","f","g","h","j","k","l","m","n","p","r","s","t","v","w","y","z","br","ch","cl","cr","dr","fl","fr","gl","gr","pl","pr","sh","sl","sm","sn","sp","st","str","sw","th","tr","wh","qu",""};
static const char *NUCLEUS[] = {"a","e","i","o","u","ai","ea","ee","oo","ou","ie","oa","ay","ow"};
static const char *CODA[] = {"","","","n","r","s","t","l","d","m","ng","nd","st","ck","ll","ss","rt","nt","sh","th","p","g","ve","ly","er","ed","ing"};
#define NEL(a) ((int)(sizeof(a) / sizeof((a)[0])))
static const char *COMMON[] = {"the","and","a","to","was","he","she","it","they","said","in","with","her","his","you","that","of","had","day","very","so","but","on","little","big","time","were","is","one","there","not","be","all","for","girl","boy","mom","dad","saw","play","went","friend","happy","wanted","named","them","up","could","what","then","did","at","an","as","out","like","have","their","when","this","can","do","from","upon","once","would","fun","back","new","see","go","lived","dog","cat","bird","tree","ball","park","home","made","loved","too","asked","good","got","we","into","just","again","help","away","him","looked","came","found","took","thought","felt","always","my","me","are","no","your","will","know","how","who","now","come","make","want","some","more","down","around","after","other","over","because","smiled","started","something","together","both","toy","best","every","even","until"};

static uint64_t isqrt64(uint64_t x) {
    uint64_t r = 0, bit = (uint64_t)1 << 62;
    while (bit > x) bit >>= 2;
    while (bit) { if (x >= r + bit) { x -= r + bit; r = (r >> 1) + bit; } else r >>= 1; bit >>= 2; }
    return r;
}

static int make_word(rng_t *r, char *dst, int syll) {
    int n = 0;
    for (int s = 0; s < syll; s++) {
        const char *a = ONSET[rng_below(r, NEL(ONSET))], *b = NUCLEUS[rng_below(r, NEL(NUCLEUS))], *c = CODA[rng_below(r, NEL(CODA))];
        n += sprintf(dst + n, "%s%s%s", a, b, (s == syll - 1 || rng_below(r, 4) == 0) ? c : "");
    }
    return n;
}

static void lex_build(lex_t *L, int n, uint64_t seed, int shift) {
    rng_t r = {seed ^ 0xABCDEF12345ull};
    L->n = n;
    L->pool = (char *)malloc((size_t)n * 24);
    L->off = (uint32_t *)malloc(sizeof(uint32_t) * (n + 1));
    L->cum = (uint64_t *)malloc(sizeof(uint64_t) * n);
    uint32_t at = 0;
    uint64_t acc = 0;
    for (int k = 0; k < n; k++) {
        L->off[k] = at;
        if (k < NEL(COMMON)) {
            at += (uint32_t)sprintf(L->pool + at, "%s", COMMON[k]);
        } else {
            int syll = 1 + (k > 1500) + (k > 20000 && rng_below(&r, 2)) + (k > 60000 && rng_below(&r, 3) == 0);
            at += (uint32_t)make_word(&r, L->pool + at, syll);
        }
        uint64_t kk = (uint64_t)(k + shift);                  /* Zipf, exponent 1.25, shifted; integer only */
        uint64_t root4 = isqrt64(isqrt64(kk << 32) << 16);    /* kk^(1/4) * 2^16 */
        acc += (((uint64_t)1 << 60) / kk) / root4;
        L->cum[k] = acc;
    }
    L->off[n] = at;
    L->total = acc;
}

static inline int lex_pick(const lex_t *L, rng_t *r) {
    uint64_t u = rng_next(r) % L->total;
    int lo = 0, hi = L->n - 1;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (L->cum[mid] > u) hi = mid; else lo = mid + 1; }
    return lo;
}

/* ---- document writers ------------------------------------------------------------------------- */
typedef struct { char *p; size_t n, cap; } buf_t;
static inline void put(buf_t *b, const char *s, size_t k) {
    if (b->n + k + 64 > b->cap) { b->cap = (b->cap + k) * 2 + 4096; b->p = (char *)realloc(b->p, b->cap); }
    memcpy(b->p + b->n, s, k); b->n += k;
}
static inline void puts_(buf_t *b, const char *s) { put(b, s, strlen(s)); }

static const char *CONTR[] = {"'s","'t","'ll","'ve","'re","'d","'m"};
static const char *ACCENT[] = {"caf\xc3\xa9","na\xc3\xafve","\xc3\xbc" "ber","jalape\xc3\xb1o","fa\xc3\xa7" "ade","\xe2\x80\x9cyes\xe2\x80\x9d","\xf0\x9f\x99\x83","d\xc3\xa9j\xc3\xa0","\xe2\x80\x94","se\xc3\xb1or"};
static const char *CYR[] = {"\xd0\xbf\xd1\x80\xd0\xb8\xd0\xb2\xd0\xb5\xd1\x82","\xd0\xbc\xd0\xb8\xd1\x80","\xd0\xb4\xd0\xbe\xd0\xbc","\xd0\xa0\xd0\xbe\xd1\x81\xd1\x81\xd0\xb8\xd1\x8f","\xd1\x81\xd0\xbb\xd0\xbe\xd0\xb2\xd0\xbe"};
static const char *CJK[] = {"\xe4\xbd\xa0\xe5\xa5\xbd","\xe4\xb8\x96\xe7\x95\x8c","\xe6\x9d\xb1\xe4\xba\xac","\xe3\x81\x93\xe3\x82\x93\xe3\x81\xab\xe3\x81\xa1\xe3\x81\xaf","\xed\x95\x9c\xea\xb5\xad","\xe3\x80\x80","\xef\xbc\x91\xef\xbc\x92"};

static void put_word(buf_t *b, const lex_t *L, rng_t *r, int cap_first, int style) {
    char tmp[96];
    int k = lex_pick(L, r);
    int n = (int)(L->off[k + 1] - L->off[k]);
    memcpy(tmp, L->pool + L->off[k], n);
    if (style == 1 && rng_below(r, 300) == 0) {          /* fresh rare word: unbounded tail */
        rng_t w = {rng_next(r)};
        n = make_word(&w, tmp, 2 + rng_below(r, 3));
    }
    if (cap_first && tmp[0] >= 'a' && tmp[0] <= 'z') tmp[0] -= 32;
    if (style == 1 && rng_below(r, 200) == 0) for (int i = 0; i < n; i++) if (tmp[i] >= 'a' && tmp[i] <= 'z') tmp[i] -= 32;
    put(b, tmp, n);
    if (rng_below(r, 60) == 0) puts_(b, CONTR[rng_below(r, NEL(CONTR))]);
}

static void put_sentence(buf_t *b, const lex_t *L, rng_t *r, int style) {
    int words = 4 + rng_below(r, style ? 22 : 11);
    int quoted = rng_below(r, 9) == 0;
    if (quoted) puts_(b, "\"");
    for (int w = 0; w < words; w++) {
        if (w) {
            uint32_t c = rng_below(r, 1000);
            if (c < 70) puts_(b, ", ");
            else if (style == 1 && c < 74) puts_(b, "  ");
            else if (style == 1 && c < 76) puts_(b, " \t");
            else if (c < 78 + 4 * style) puts_(b, "; ");
            else if (c < 81 + 4 * style) puts_(b, " - ");
            else puts_(b, " ");
        }
        uint32_t kind = rng_below(r, 1000);
        if (kind < 3 + 5 * style) { puts_(b, ACCENT[rng_below(r, NEL(ACCENT))]); continue; }
        if (kind < 8 + 30 * style) {                       /* number */
            char t[32]; int n = sprintf(t, "%u", rng_below(r, kind & 1 ? 100 : 100000));
            if (style == 1 && rng_below(r, 4) == 0) n += sprintf(t + n, ".%u", rng_below(r, 100));
            if (style == 1 && rng_below(r, 6) == 0) n += sprintf(t + n, "%%");
            put(b, t, n); continue;
        }
        if (style == 1) {
            if (kind < 44) { puts_(b, CYR[rng_below(r, NEL(CYR))]); continue; }
            if (kind < 50) { puts_(b, CJK[rng_below(r, NEL(CJK))]); if (rng_below(r, 2)) puts_(b, CJK[rng_below(r, NEL(CJK))]); continue; }
            if (kind < 56) {                                 /* URL */
                char t[160]; rng_t w = {rng_next(r)}; int n = sprintf(t, "http%s://www.", rng_below(r, 2) ? "s" : "");
                n += make_word(&w, t + n, 2); n += sprintf(t + n, ".%s/", rng_below(r, 2) ? "com" : "org");
                n += make_word(&w, t + n, 2); n += sprintf(t + n, "?id=%u&x=%u", rng_below(r, 1000000), rng_below(r, 99));
                put(b, t, n); continue;
            }
            if (kind < 58) { int run = 3 + rng_below(r, rng_below(r, 50) == 0 ? 3000 : 40); const char *ch = rng_below(r, 2) ? "-" : "="; for (int i = 0; i < run; i++) puts_(b, ch); continue; }
            if (kind < 59) { int run = 2 + rng_below(r, rng_below(r, 30) == 0 ? 2000 : 12); for (int i = 0; i < run; i++) puts_(b, " "); puts_(b, "x"); continue; }
        }
        put_word(b, L, r, w == 0, style);
    }
    uint32_t e = rng_below(r, 100);
    puts_(b, e < 78 ? "." : e < 88 ? "!" : e < 95 ? "?" : e < 98 ? "..." : "!!!");
    if (quoted) puts_(b, "\"");
}

static void put_document(buf_t *b, const lex_t *L, uint64_t seed, uint64_t doc, int style) {
    rng_t h = {doc ^ 0x5851F42D4C957F2Dull};               /* decorrelate neighbouring documents */
    rng_t r = {rng_next(&h) ^ (seed * 0x100000001B3ull)};
    r.s = rng_next(&r);
    int sentences = style ? 5 + (int)rng_below(&r, 120) : 8 + (int)rng_below(&r, 18);
    for (int s = 0; s < sentences; s++) {
        if (s) puts_(b, rng_below(&r, 6) == 0 ? "\n\n" : (style == 1 && rng_below(&r, 25) == 0) ? "\n" : " ");
        put_sentence(b, L, &r, style);
    }
    puts_(b, style ? "<|endoftext|>" : "\n<|endoftext|>\n");
}

/* ---- parallel driver -------------------------------------------------------------------------- */
#define UNIT_DOCS 256
typedef struct {
    const lex_t *L; uint64_t seed; int style;
    uint64_t base, n_units;            /* this round covers global units [base, base + n_units) */
    buf_t *units; uint32_t (*doc_len)[UNIT_DOCS];
    volatile uint64_t *next;
} job_t;

static void *worker(void *arg) {
    job_t *j = (job_t *)arg;
    for (;;) {
        uint64_t u = __sync_fetch_and_add(j->next, 1);
        if (u >= j->n_units) return NULL;
        buf_t *b = &j->units[u];
        for (int d = 0; d < UNIT_DOCS; d++) {
            size_t before = b->n;
            put_document(b, j->L, j->seed, (j->base + u) * UNIT_DOCS + d, j->style);
            j->doc_len[u][d] = (uint32_t)(b->n - before);
        }
    }
}

/* Writes the corpus to `path`; returns the number of bytes written (0 on failure). */
uint64_t gen_corpus(int style, uint64_t seed, uint64_t target, const char *path, int n_threads) {
    lex_t L;
    lex_build(&L, style ? 180000 : 24000, seed, style ? 6 : 3);
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 64) n_threads = 64;
    FILE *f = fopen(path, "wb");
    if (!f) return 0;
    uint64_t written = 0, base = 0;
    const uint64_t avg_doc = style ? 8000 : 1100;
    int done = 0;
    while (!done) {                        /* rounds of units until the target is reached */
        uint64_t remaining = target > written ? target - written : 1;
        uint64_t n_units = remaining / (avg_doc * UNIT_DOCS) + 1;
        if (n_units > 4096) n_units = 4096;
        buf_t *units = (buf_t *)calloc(n_units, sizeof(buf_t));
        uint32_t(*doc_len)[UNIT_DOCS] = (uint32_t(*)[UNIT_DOCS])calloc(n_units, sizeof(*doc_len));
        volatile uint64_t next = 0;
        job_t job = {&L, seed, style, base, n_units, units, doc_len, &next};
        pthread_t th[64];
        for (int t = 0; t < n_threads; t++) pthread_create(&th[t], NULL, worker, &job);
        for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
        for (uint64_t u = 0; u < n_units && !done; u++) {
            size_t take = 0;
            for (int d = 0; d < UNIT_DOCS; d++) {
                take += doc_len[u][d];
                if (written + take >= target) { done = 1; break; }
            }
            if (fwrite(units[u].p, 1, take, f) != take) { fclose(f); return 0; }
            written += take;
        }
        for (uint64_t u = 0; u < n_units; u++) free(units[u].p);
        free(units); free(doc_len);
        base += n_units;
    }
    fclose(f);
    free(L.pool); free(L.off); free(L.cum);
    return written;
}

#ifdef GEN_MAIN
int main(int argc, char **argv) {
    if (argc < 5) { fprintf(stderr, "usage: %s style(0|1) seed bytes out [threads]\n", argv[0]); return 2; }
    uint64_t n = gen_corpus(atoi(argv[1]), strtoull(argv[2], 0, 10), strtoull(argv[3], 0, 10), argv[4], argc > 5 ? atoi(argv[5]) : 8);
    printf("%llu\n", (unsigned long long)n);
    return n ? 0 : 1;
}
#endif
