// scan.cu — token-start bitmask of a UTF-8 text (the GPT-2 split pattern as a local, per-byte rule).
//
// Replaces regex.split on the special tokens + regex.finditer(PAT) of the reference
// (bpe_transformer/tokenization/preprocessing/pretokenization.py:211-252, pattern bpe_transformer/settings.py:8).
// The pattern needs no sequential state: whether a token starts at byte i depends on the classes of the
// characters from 4 before to 2 after i.  So every thread owns 32 bytes, classifies
// them SIMD-within-a-register (4 bytes per 32-bit op), reduces the classes to 32-bit masks and evaluates the
// rule with ~30 bitwise ops for all 32 bytes at once.  Rare things leave the bit-parallel path:
//   * a byte that may begin a special token  -> exact match + leftmost-non-overlapping resolution (scalar)
//   * non-ASCII bytes                        -> decode, two-stage class table, strict UTF-8 validation
//   * apostrophes                            -> contraction rule 's 'd 'm 't 'll 've 're (scalar)
// HBM traffic: N bytes read, N/8 bytes written.  Grid = persistent, a multiple of 148 SMs.
#include "common.cuh"

#define SCAN_THREADS 256
#define SCAN_HALO 2  // chunks in front of a tile that are searched for special tokens reaching into it

namespace {

struct ScanParams {
    const uint8_t* text;
    uint64_t n;
    const uint32_t* cutmask;  // may be null; bit i = boundary in front of byte i
    uint32_t flags;
    uint32_t* start;
    uint32_t* status;
    ClassTables cls;
    uint64_t n_tiles;
};

// ---- SWAR byte classification: each helper returns 0x80 in every byte that qualifies -----------------------
__device__ __forceinline__ uint32_t sw_ge(uint32_t y, uint32_t c) { return y + (0x80u - c) * 0x01010101u; }  // bit7: y >= c
__device__ __forceinline__ uint32_t sw_range(uint32_t y, uint32_t lo, uint32_t hi) { return sw_ge(y, lo) & ~sw_ge(y, hi + 1u); }
__device__ __forceinline__ uint32_t sw_eq(uint32_t y, uint32_t c) { return ~((y ^ (c * 0x01010101u)) + 0x7F7F7F7Fu); }
// 0x80-per-byte flags -> 4-bit nibble (bit k = byte k)
__device__ __forceinline__ uint32_t sw_collapse(uint32_t m) { return (((m >> 7) & 0x01010101u) * 0x00204081u >> 21) & 0xFu; }

struct WordClass {
    uint32_t L, N, S, SP, AP, NL, HI, FB;  // 0x80 flags (garbage in the low 7 bits of each byte until collapsed)
};

__device__ __forceinline__ WordClass classify_word(uint32_t w, uint32_t fb0, uint32_t fb1) {
    WordClass c;
    uint32_t hi = w & 0x80808080u;
    uint32_t y = w & 0x7F7F7F7Fu;
    uint32_t ascii = ~hi;
    c.HI = hi;
    c.L = sw_range(y | 0x20202020u, 'a', 'z') & ascii;
    c.N = sw_range(y, '0', '9') & ascii;
    c.SP = sw_eq(y, ' ') & ascii;
    c.S = (sw_range(y, 9, 13) & ascii) | c.SP;
    c.AP = sw_eq(y, '\'') & ascii;
    c.NL = sw_eq(y, '\n') & ascii;
    c.FB = (sw_eq(y, fb0) | sw_eq(y, fb1)) & ascii;
    return c;
}

// ---- scalar helpers for the rare paths ------------------------------------------------------------------------
// Strict UTF-8 decode of the character that contains byte `pos`.  Returns false if that byte is not part of
// a valid sequence inside [0, n).  *lead = position of the first byte, *len = sequence length.
__device__ bool char_containing(const uint8_t* __restrict__ t, uint64_t pos, uint64_t n, uint64_t* lead, uint32_t* len, uint32_t* cp) {
    uint64_t p = pos;
    for (int k = 0; k < 3 && p > 0 && (t[p] & 0xC0) == 0x80; k++) p--;
    uint32_t b0 = t[p];
    uint32_t L, c;
    if (b0 < 0x80) { L = 1; c = b0; }
    else if (b0 < 0xC2) return false;
    else if (b0 < 0xE0) { L = 2; c = b0 & 0x1F; }
    else if (b0 < 0xF0) { L = 3; c = b0 & 0x0F; }
    else if (b0 < 0xF5) { L = 4; c = b0 & 0x07; }
    else return false;
    if (p + L > n || p + L <= pos) return false;
    for (uint32_t k = 1; k < L; k++) {
        uint32_t b = t[p + k];
        if ((b & 0xC0) != 0x80) return false;
        c = (c << 6) | (b & 0x3F);
    }
    if (L == 3 && (c < 0x800 || (c >= 0xD800 && c <= 0xDFFF))) return false;
    if (L == 4 && (c < 0x10000 || c > 0x10FFFF)) return false;
    *lead = p; *len = L; *cp = c;
    return true;
}

__device__ __forceinline__ uint32_t ascii_class(uint32_t b) {
    if (((b | 0x20) - 'a') < 26u) return CLS_L;
    if ((b - '0') < 10u) return CLS_N;
    if (b == ' ' || (b - 9u) < 5u) return CLS_S;
    return CLS_O;
}

// Class of the character containing byte pos (CLS_O when the bytes are not valid UTF-8: the status word
// reports that separately).
__device__ uint32_t class_containing(const ScanParams& P, uint64_t pos) {
    uint32_t b = P.text[pos];
    if (b < 0x80) return ascii_class(b);
    uint64_t lead; uint32_t len, cp;
    if (!char_containing(P.text, pos, P.n, &lead, &len, &cp)) return CLS_O;
    return cp_class(P.cls, cp);
}

struct Tile {
    uint64_t byte0;        // first byte of the tile's core
    const uint32_t* wall;  // shared: words for chunks [-SCAN_HALO, SCAN_THREADS + 1)
};

__device__ __forceinline__ bool wall_bit(const Tile& T, int64_t rel) {  // rel = byte position relative to the tile core
    int64_t idx = rel + SCAN_HALO * 32;
    return (T.wall[idx >> 5] >> (idx & 31)) & 1u;
}

// Is there a segment boundary directly in front of byte `pos`?  (text start / end, after a special token,
// before a special token, at a caller cut, after a newline in newline mode)
__device__ bool boundary_before(const ScanParams& P, const Tile& T, uint64_t pos) {
    if (pos == 0 || pos >= P.n) return true;
    int64_t rel = (int64_t)pos - (int64_t)T.byte0;
    if (wall_bit(T, rel) || wall_bit(T, rel - 1)) return true;
    if (P.cutmask && ((P.cutmask[pos >> 5] >> (pos & 31)) & 1u)) return true;
    if ((P.flags & BPE_CUT_AFTER_NEWLINE) && P.text[pos - 1] == '\n') return true;
    return false;
}

// First position > p at which a special token starting at p must end at the latest: the next cut.
__device__ uint64_t piece_limit(const ScanParams& P, uint64_t p, int max_len) {
    uint64_t lim = p + (uint64_t)max_len < P.n ? p + (uint64_t)max_len : P.n;
    if (P.flags & BPE_CUT_AFTER_NEWLINE)
        for (uint64_t q = p; q < lim; q++)
            if (P.text[q] == '\n') { lim = q + 1; break; }
    if (P.cutmask)
        for (uint64_t q = p + 1; q < lim; q++)
            if ((P.cutmask[q >> 5] >> (q & 31)) & 1u) { lim = q; break; }
    return lim;
}

__device__ uint32_t special_len_at(const ScanParams& P, const Specials& sp, uint64_t p) {
    uint8_t b0 = P.text[p];
    if (!((sp.first_byte_mask[b0 >> 5] >> (b0 & 31)) & 1u)) return 0;
    int s = special_at(sp, P.text, p, piece_limit(P, p, sp.max_len));
    return s < 0 ? 0u : sp.off[s + 1] - sp.off[s];
}

// regex.split scans left to right and continues after each match, so a candidate is a real match iff the
// scan arrives exactly at it.  A candidate with no other candidate in the max_len-1 bytes before it cannot
// be covered: walk back to such a chain head, then replay the scan from there.
__device__ bool special_is_real(const ScanParams& P, const Specials& sp, uint64_t p) {
    uint64_t head = p, j = p;
    while (j > 0) {
        j--;
        if (head - j >= (uint64_t)sp.max_len) break;
        if (special_len_at(P, sp, j)) head = j;
    }
    uint64_t pos = head;
    while (pos < p) {
        uint32_t L = special_len_at(P, sp, pos);
        pos += L ? L : 1u;
    }
    return pos == p;
}

__device__ void mark_special(const ScanParams& P, const Specials& sp, uint64_t tile0, uint32_t* wall, uint32_t* sps, uint64_t p) {
    uint32_t L = special_len_at(P, sp, p);
    if (!L || !special_is_real(P, sp, p)) return;
    int64_t rel = (int64_t)p - (int64_t)tile0 + SCAN_HALO * 32;
    int64_t hi = (int64_t)(SCAN_THREADS + SCAN_HALO + 1) * 32;
    atomicOr(&sps[rel >> 5], 1u << (rel & 31));
    for (int64_t q = rel; q < rel + (int64_t)L && q < hi; q++) atomicOr(&wall[q >> 5], 1u << (q & 31));
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_kernel(const ScanParams P, const __grid_constant__ Specials sp) {
    __shared__ uint32_t wall[SCAN_THREADS + SCAN_HALO + 1];  // bytes covered by a special token
    __shared__ uint32_t sps[SCAN_THREADS + SCAN_HALO + 1];   // first byte of a special token
    const uint32_t tid = threadIdx.x;
    const uint32_t fb0 = sp.n ? sp.blob[sp.off[0]] : 0xFFu;  // up to two distinct ASCII first bytes go the fast way
    uint32_t fb1 = fb0;
    bool fb_generic = false;  // otherwise: test every byte against the 256-bit set
    for (int s = 0; s < sp.n; s++) {
        uint32_t b = sp.blob[sp.off[s]];
        if (b >= 0x80) fb_generic = true;
        else if (b != fb0 && fb1 == fb0) fb1 = b;
        else if (b != fb0 && b != fb1) fb_generic = true;
    }
    const bool nlmode = P.flags & BPE_CUT_AFTER_NEWLINE;

    for (uint64_t tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
        const uint64_t tile0 = tile * (uint64_t)(SCAN_THREADS * 32);
        const uint64_t c0 = tile0 + (uint64_t)tid * 32;  // first byte of this thread's chunk
        for (uint32_t k = tid; k < SCAN_THREADS + SCAN_HALO + 1; k += SCAN_THREADS) { wall[k] = 0; sps[k] = 0; }

        // ---- load 32 bytes + 4 before + 4 after, classify ------------------------------------------------
        uint32_t w[10];
        {
            const uint4* q = reinterpret_cast<const uint4*>(P.text + c0);
            uint4 a = make_uint4(0, 0, 0, 0), b = a;
            if (c0 < P.n) a = __ldg(q);
            if (c0 + 16 < P.n) b = __ldg(q + 1);
            w[1] = a.x; w[2] = a.y; w[3] = a.z; w[4] = a.w; w[5] = b.x; w[6] = b.y; w[7] = b.z; w[8] = b.w;
            w[0] = (c0 >= 4 && c0 - 4 < P.n) ? __ldg(reinterpret_cast<const uint32_t*>(P.text + c0 - 4)) : 0u;
            w[9] = (c0 + 32 < P.n) ? __ldg(reinterpret_cast<const uint32_t*>(P.text + c0 + 32)) : 0u;
        }
        uint32_t L = 0, N = 0, S = 0, SP = 0, AP = 0, NL = 0, HI = 0, FB = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            WordClass c = classify_word(w[k + 1], fb0, fb1);
            L |= sw_collapse(c.L) << (4 * k);
            N |= sw_collapse(c.N) << (4 * k);
            S |= sw_collapse(c.S) << (4 * k);
            SP |= sw_collapse(c.SP) << (4 * k);
            AP |= sw_collapse(c.AP) << (4 * k);
            NL |= sw_collapse(c.NL) << (4 * k);
            HI |= sw_collapse(c.HI) << (4 * k);
            FB |= sw_collapse(c.FB) << (4 * k);
        }
        WordClass cp_ = classify_word(w[0], fb0, fb1), cn_ = classify_word(w[9], fb0, fb1);
        // previous four bytes in bits 28..31, next four bytes in bits 0..3
        uint32_t Lp = sw_collapse(cp_.L) << 28, Np = sw_collapse(cp_.N) << 28, Sp = sw_collapse(cp_.S) << 28, SPp = sw_collapse(cp_.SP) << 28;
        uint32_t APp = sw_collapse(cp_.AP) << 28, NLp = sw_collapse(cp_.NL) << 28, HIp = sw_collapse(cp_.HI) << 28;
        uint32_t Ln = sw_collapse(cn_.L), Nn = sw_collapse(cn_.N), Sn = sw_collapse(cn_.S), HIn = sw_collapse(cn_.HI);

        // bytes at or beyond n do not exist
        uint32_t valid = (c0 >= P.n) ? 0u : (P.n - c0 >= 32 ? 0xFFFFFFFFu : ((1u << (uint32_t)(P.n - c0)) - 1u));
        uint32_t validn = (c0 + 32 >= P.n) ? 0u : (P.n - c0 - 32 >= 4 ? 0xFu : ((1u << (uint32_t)(P.n - c0 - 32)) - 1u));
        if (fb_generic) {
            FB = 0;
            for (uint32_t b = 0; b < 32; b++) {
                uint32_t x = (w[1 + (b >> 2)] >> (8 * (b & 3))) & 0xFF;
                if ((sp.first_byte_mask[x >> 5] >> (x & 31)) & 1u) FB |= 1u << b;
            }
        }
        FB &= valid;
        __syncthreads();  // wall / sps cleared

        // ---- phase 1: special tokens that start in this chunk (or, for a few threads, in the halo chunks) ----
        if (sp.n) {
            for (uint32_t m = FB; m; m &= m - 1) mark_special(P, sp, tile0, wall, sps, c0 + (__ffs(m) - 1));
            if (tid <= SCAN_HALO) {
                int64_t chunk = (tid < SCAN_HALO) ? -(int64_t)(tid + 1) : (int64_t)SCAN_THREADS;
                int64_t b0 = (int64_t)tile0 + chunk * 32;
                for (int64_t p = b0; p < b0 + 32; p++)
                    if (p >= 0 && (uint64_t)p < P.n) {
                        uint8_t x = P.text[p];
                        if ((sp.first_byte_mask[x >> 5] >> (x & 31)) & 1u) mark_special(P, sp, tile0, wall, sps, (uint64_t)p);
                    }
            }
        }
        __syncthreads();

        if (c0 < P.n) {
            Tile T{tile0, wall};
            uint32_t W = wall[tid + SCAN_HALO] | ~valid;
            uint32_t Wp = (c0 == 0) ? 0xFFFFFFFFu : wall[tid + SCAN_HALO - 1];
            uint32_t Wn = wall[tid + SCAN_HALO + 1] | ~validn;
            uint32_t SPS = sps[tid + SCAN_HALO];
            uint32_t CONT = 0;

            // ---- non-ASCII: decode, look the class up, validate ---------------------------------------------
            if ((HI & valid) | HIp | (HIn & validn)) {
                for (uint32_t m = HI & valid; m; m &= m - 1) {
                    uint32_t b = __ffs(m) - 1;
                    uint64_t lead; uint32_t len, cp;
                    if (!char_containing(P.text, c0 + b, P.n, &lead, &len, &cp)) { set_status(P.status, ST_INVALID_UTF8); continue; }
                    uint32_t k = cp_class(P.cls, cp);
                    if (k == CLS_L) L |= 1u << b; else if (k == CLS_N) N |= 1u << b; else if (k == CLS_S) S |= 1u << b;
                    if (lead != c0 + b) CONT |= 1u << b;
                }
                if (HIp >> 31) {  // only the class of the byte right in front matters to the bit-parallel rule
                    uint32_t k = class_containing(P, c0 - 1);
                    if (k == CLS_L) Lp |= 1u << 31; else if (k == CLS_N) Np |= 1u << 31; else if (k == CLS_S) Sp |= 1u << 31;
                }
                if (HIn & validn & 1u) {
                    uint32_t k = class_containing(P, c0 + 32);
                    if (k == CLS_L) Ln |= 1u; else if (k == CLS_N) Nn |= 1u; else if (k == CLS_S) Sn |= 1u;
                }
            }

            // ---- the rule, 32 bytes at a time ------------------------------------------------------------------
            uint32_t cut = P.cutmask ? __ldg(P.cutmask + (c0 >> 5)) : 0u;
            uint32_t cutn = P.cutmask ? __ldg(P.cutmask + (c0 >> 5) + 1) : 0u;
            uint32_t PN = ((W << 1) | (Wp >> 31)) | cut;                  // nothing in front of this byte in its segment
            uint32_t NN = ((W >> 1) | (Wn << 31)) | ((cut >> 1) | (cutn << 31));  // nothing behind it
            if (nlmode) { PN |= (NL << 1) | (NLp >> 31); NN |= NL; }
            L &= ~W; N &= ~W; S &= ~W; SP &= ~W;
            uint32_t O = ~(L | N | S | W);
            uint32_t pL = (L << 1) | (Lp >> 31), pN = (N << 1) | (Np >> 31), pS = (S << 1) | (Sp >> 31), pSP = (SP << 1) | (SPp >> 31);
            uint32_t pO = ~(pL | pN | pS);  // only read where the previous byte exists and is no wall
            uint32_t nonS = L | N | O;
            uint32_t nNonS = ((nonS >> 1) | ((~Sn & 1u) << 31)) & ~NN;   // the next byte exists and is not whitespace
            uint32_t diff = (L & ~pL) | (N & ~pN) | (O & ~pO);
            uint32_t startNS = nonS & (PN | (~pSP & (pS | diff)));
            uint32_t startS = S & (PN | ~pS | nNonS);
            uint32_t start = (startNS | startS) & ~CONT;

            // multi-byte whitespace (NBSP, U+2000.., U+3000): "followed by non-whitespace" looks past the whole character
            for (uint32_t m = S & HI & ~CONT & ~start; m; m &= m - 1) {
                uint32_t b = __ffs(m) - 1;
                uint64_t lead; uint32_t len, cp;
                if (!char_containing(P.text, c0 + b, P.n, &lead, &len, &cp)) continue;
                uint64_t nx = c0 + b + len;
                if (!boundary_before(P, T, nx) && class_containing(P, nx) != CLS_S) start |= 1u << b;
            }

            // ---- contractions: an apostrophe that starts a token swallows s d m t / ll ve re --------------------------
            uint32_t apc = AP & ~W;
            uint64_t aph = APp >> 29;  // apostrophes 3, 2, 1 bytes in front of the chunk -> bits 0..2
            if (apc | (uint32_t)aph) {
                uint64_t cand = ((uint64_t)apc << 3) | aph;  // bit k <-> byte c0 + k - 3
                for (; cand; cand &= cand - 1) {
                    int64_t a = (int64_t)c0 + (__ffsll((long long)cand) - 1) - 3;
                    if (a < 0 || (uint64_t)a >= P.n) continue;
                    uint64_t ua = (uint64_t)a;
                    if (wall_bit(T, a - (int64_t)tile0)) continue;
                    bool starts = boundary_before(P, T, ua);
                    if (!starts) {
                        uint32_t pb = P.text[ua - 1];
                        uint32_t k = class_containing(P, ua - 1);
                        starts = (k == CLS_L) || (k == CLS_N) || (k == CLS_S && pb != ' ');
                    }
                    if (!starts || boundary_before(P, T, ua + 1)) continue;
                    uint32_t b1 = P.text[ua + 1];
                    uint64_t after = 0;  // first byte behind the contraction
                    if (b1 == 's' || b1 == 'd' || b1 == 'm' || b1 == 't') after = ua + 2;
                    else if (!boundary_before(P, T, ua + 2)) {
                        uint32_t b2 = P.text[ua + 2];
                        if ((b1 == 'l' && b2 == 'l') || (b1 == 'v' && b2 == 'e') || (b1 == 'r' && b2 == 'e')) after = ua + 3;
                    }
                    if (!after) continue;
                    if (ua + 1 >= c0 && ua + 1 < c0 + 32) start &= ~(1u << (uint32_t)(ua + 1 - c0));  // the letter belongs to the apostrophe
                    if (after >= c0 && after < c0 + 32 && !boundary_before(P, T, after) && class_containing(P, after) == CLS_L)
                        start |= 1u << (uint32_t)(after - c0);  // a letter run begins behind the contraction
                }
            }
            start = (start & ~W) | SPS;
            P.start[c0 >> 5] = start;
        }
        __syncthreads();  // wall / sps are re-cleared by the next iteration
    }
}

__global__ void cuts_to_mask_kernel(const uint64_t* __restrict__ cuts, uint64_t n_cuts, uint64_t n, uint32_t* __restrict__ mask) {
    uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (i < n_cuts) {
        uint64_t p = cuts[i];
        if (p < n) atomicOr(&mask[p >> 5], 1u << (p & 31));
    }
}

}  // namespace

int launch_scan(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint32_t* d_cutmask, uint32_t flags, uint32_t* d_start, cudaStream_t s) {
    uint64_t n_words = n / 32 + 2;
    CUDA_TRY(ctx, cudaMemsetAsync(d_start + (n / 32), 0, 2 * sizeof(uint32_t), s));
    (void)n_words;
    if (n == 0) return BPE_OK;
    ScanParams P;
    P.text = d_text; P.n = n; P.cutmask = d_cutmask; P.flags = flags; P.start = d_start; P.status = ctx->d_status; P.cls = ctx->cls;
    P.n_tiles = (n + SCAN_THREADS * 32 - 1) / (SCAN_THREADS * 32);
    uint64_t grid = P.n_tiles < (uint64_t)BPE_N_SM * 8 ? P.n_tiles : (uint64_t)BPE_N_SM * 8;
    KernelTimer t(ctx, "scan_kernel", s);
    scan_kernel<<<(unsigned)grid, SCAN_THREADS, 0, s>>>(P, ctx->sp);
    CUDA_TRY(ctx, cudaGetLastError());
    return BPE_OK;
}

int launch_cuts_to_mask(bpe_ctx* ctx, const uint64_t* d_cuts, uint64_t n_cuts, uint64_t n, uint32_t* d_cutmask, cudaStream_t s) {
    CUDA_TRY(ctx, cudaMemsetAsync(d_cutmask, 0, (n / 32 + 2) * sizeof(uint32_t), s));
    if (n_cuts) {
        cuts_to_mask_kernel<<<(unsigned)((n_cuts + 255) / 256), 256, 0, s>>>(d_cuts, n_cuts, n, d_cutmask);
        CUDA_TRY(ctx, cudaGetLastError());
    }
    return BPE_OK;
}
