// pretok.cuh — the GPT-2 split pattern as a local per-byte rule, in the form both the mask kernel (scan.cu) and the
// fused one-pass kernels (fused.cu) use.
//
// Replaces regex.split on the special tokens + regex.finditer(PAT) of the reference
// (bpe_transformer/tokenization/preprocessing/pretokenization.py:211-252, pattern bpe_transformer/settings.py:8).
// Whether a token starts at byte i depends only on the classes of the characters from 4 before to 2 after i
// (SURVEY.md Appendix A.1), so a thread that owns 32 bytes can decide all 32 positions with ~40 bitwise operations once it
// knows the classes of its neighbours' edge bytes.  scan.cu gets those through shared memory (a CTA per 8 KiB tile);
// the WARP WINDOW below gets them through shuffles: a warp covers 1 KiB, lane L owns bytes [32 L, 32 L + 32) of it, lanes
// 0, 1 and 31 are ghosts (classified like the others so that the 29 core lanes see exact neighbours; their own results
// are not used), and nothing leaves the registers — which is what lets hashing, counting and id lookup be fused behind it.
#pragma once
#include "common.cuh"

#define PT_FULL 0xFFFFFFFFu
#define PT_LG 2                       // ghost lanes in front of the core (special tokens of up to 60 bytes reach across two chunks)
#define PT_RG 1                       // ghost lane behind it
#define PT_CORE (32 - PT_LG - PT_RG)  // 29 core lanes
#define PT_WIN_BYTES (PT_CORE * 32)   // 928 bytes of text per window
#define PT_RELIABLE 24                // bits 0..24 of the right ghost's mask depend on no byte beyond the window

struct PretokParams {
    const uint8_t* text;
    uint64_t n;
    const uint32_t* cutmask;  // may be null; bit i = boundary in front of byte i
    uint32_t flags;
    uint32_t* status;
    ClassTables cls;
};

// ---- SWAR byte classification: each helper returns 0x80 in every byte that qualifies -----------------------
__device__ __forceinline__ uint32_t sw_ge(uint32_t y, uint32_t c) { return y + (0x80u - c) * 0x01010101u; }  // bit7: y >= c
__device__ __forceinline__ uint32_t sw_range(uint32_t y, uint32_t lo, uint32_t hi) { return sw_ge(y, lo) & ~sw_ge(y, hi + 1u); }
__device__ __forceinline__ uint32_t sw_eq(uint32_t y, uint32_t c) { return ~((y ^ (c * 0x01010101u)) + 0x7F7F7F7Fu); }
// 0x80-per-byte flags -> 4-bit nibble (bit k = byte k)
__device__ __forceinline__ uint32_t sw_collapse(uint32_t m) { return (((m >> 7) & 0x01010101u) * 0x00204081u >> 21) & 0xFu; }

// ---- scalar helpers for the rare paths ------------------------------------------------------------------------
// Strict UTF-8 decode of the character that contains byte `pos`.  Returns false if that byte is not part of
// a valid sequence inside [0, n).  *lead = position of the first byte, *len = sequence length.
// The rarely taken helpers below are real calls in the one-pass kernel (its hot loop must stay inside the instruction
// cache) and inlined in the mask kernel (scan.cu defines PRETOK_SLOW before including this file: a call under its CTA barriers
// makes every warp wait for the slowest one's round trip).
#ifndef PRETOK_SLOW
#define PRETOK_SLOW __noinline__
#endif
static __device__ PRETOK_SLOW bool char_containing(const uint8_t* __restrict__ t, uint64_t pos, uint64_t n, uint64_t* lead, uint32_t* len, uint32_t* cp) {
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
static __device__ PRETOK_SLOW uint32_t class_containing(const PretokParams& P, uint64_t pos) {
    uint32_t b = P.text[pos];
    if (b < 0x80) return ascii_class(b);
    uint64_t lead; uint32_t len, cp;
    if (!char_containing(P.text, pos, P.n, &lead, &len, &cp)) return CLS_O;
    return cp_class(P.cls, cp);
}

// Length of the special token (priority order) that matches the text at p and ends before the next cut; 0 if none.
// The bytes are compared first, eight at a time; only a match looks at the cuts: a newline inside the token (when every
// newline is a cut — known from the token itself) or a caller's cut at p + 1 .. p + len - 1 splits it, so it is no match.
static __device__ PRETOK_SLOW uint32_t special_len_at(const PretokParams& P, const Specials& sp, uint64_t p) {
    uint8_t b0 = P.text[p];
    if (!((sp.first_byte_mask[b0 >> 5] >> (b0 & 31)) & 1u)) return 0;
    for (int s = 0; s < sp.n; s++) {
        const uint32_t len = sp.off[s + 1] - sp.off[s];
        if (p + len > P.n || !special_equals(sp, s, P.text, p, len)) continue;
        if ((P.flags & BPE_CUT_AFTER_NEWLINE) && ((sp.newline_inside >> s) & 1u)) continue;
        if (P.cutmask) {
            bool cut = false;
            for (uint64_t q = p + 1; q < p + len && !cut; q++) cut = (P.cutmask[q >> 5] >> (q & 31)) & 1u;
            if (cut) continue;
        }
        return len;
    }
    return 0;
}

// regex.split scans left to right and continues after each match, so a candidate is a real match iff the
// scan arrives exactly at it.  A candidate with no other candidate in the max_len-1 bytes before it cannot
// be covered: walk back to such a chain head, then replay the scan from there.
static __device__ PRETOK_SLOW bool special_is_real(const PretokParams& P, const Specials& sp, uint64_t p) {
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

// Up to two distinct ASCII first bytes of the special tokens go the SWAR way; anything else tests every byte against
// the 256-bit set.
struct SpecialFirst {
    uint32_t fb0, fb1;
    bool generic;
};

__device__ __forceinline__ SpecialFirst special_first(const Specials& sp) {
    SpecialFirst f;
    f.fb0 = sp.n ? sp.blob[sp.off[0]] : 0xFFu;
    f.fb1 = f.fb0;
    f.generic = false;
    for (int s = 0; s < sp.n; s++) {
        uint32_t b = sp.blob[sp.off[s]];
        if (b >= 0x80) f.generic = true;
        else if (b != f.fb0 && f.fb1 == f.fb0) f.fb1 = b;
        else if (b != f.fb0 && b != f.fb1) f.generic = true;
    }
    return f;
}

// ---- phase A: 32 bytes -> class masks -----------------------------------------------------------------------------
struct ChunkClasses {
    uint32_t L, N, S, SP, NL, AP, CONT, HI, FB;
};

// Two words (8 bytes) per trip and the eight words rotated by two after each trip — four trips put them back where they
// were — so that the loop body exists once: the fused kernels' hot loop has to fit the instruction cache, and indexing
// the register array with the trip count would push it to local memory.
// (COMPACT = false unrolls the four trips: faster where the instruction cache is no concern, as in scan.cu's small kernel.)
template <bool NLMODE, bool COMPACT = true>
__device__ __forceinline__ ChunkClasses classify_chunk(const PretokParams& P, const Specials& sp, const SpecialFirst& sf, uint32_t (&w)[8],
                                                       uint64_t c0, uint32_t valid) {
    uint32_t L = 0, N = 0, S = 0, SP = 0, NL = 0, AP = 0, HI = 0, FB = 0;
#pragma unroll(COMPACT ? 1 : 4)
    for (int trip = 0; trip < 4; trip++) {
        uint32_t l8 = 0, n8 = 0, s8 = 0, sp8 = 0, nl8 = 0, ap8 = 0, hi8 = 0, fb8 = 0;
#pragma unroll
        for (int k = 0; k < 2; k++) {
            const uint32_t hi = w[k] & 0x80808080u, y = w[k] & 0x7F7F7F7Fu, ascii = ~hi;
            const uint32_t cL = sw_range(y | 0x20202020u, 'a', 'z') & ascii;
            const uint32_t cN = sw_range(y, '0', '9') & ascii;
            const uint32_t cSP = sw_eq(y, ' ') & ascii;
            const uint32_t cS = (sw_range(y, 9, 13) & ascii) | cSP;
            l8 |= sw_collapse(cL) << (4 * k);
            n8 |= sw_collapse(cN) << (4 * k);
            s8 |= sw_collapse(cS) << (4 * k);
            sp8 |= sw_collapse(cSP) << (4 * k);
            ap8 |= sw_collapse(sw_eq(y, '\'') & ascii) << (4 * k);
            hi8 |= sw_collapse(hi) << (4 * k);  // non-ASCII bytes: handled one by one below
            if (NLMODE) nl8 |= sw_collapse(sw_eq(y, '\n') & ascii) << (4 * k);
            if (sp.n) fb8 |= sw_collapse((sw_eq(y, sf.fb0) | sw_eq(y, sf.fb1)) & ascii) << (4 * k);
        }
        L = (L >> 8) | (l8 << 24); N = (N >> 8) | (n8 << 24); S = (S >> 8) | (s8 << 24); SP = (SP >> 8) | (sp8 << 24);
        AP = (AP >> 8) | (ap8 << 24); HI = (HI >> 8) | (hi8 << 24);
        if (NLMODE) NL = (NL >> 8) | (nl8 << 24);
        if (sp.n) FB = (FB >> 8) | (fb8 << 24);
        const uint32_t t0 = w[0], t1 = w[1];
        w[0] = w[2]; w[1] = w[3]; w[2] = w[4]; w[3] = w[5]; w[4] = w[6]; w[5] = w[7]; w[6] = t0; w[7] = t1;
    }
    L &= valid; N &= valid; S &= valid; SP &= valid; AP &= valid; HI &= valid; NL &= valid;
    if (sf.generic) {
        FB = 0;
#pragma unroll 1
        for (uint32_t b = 0; b < 32 && c0 + b < P.n; b++) {
            const uint32_t x = P.text[c0 + b];
            if ((sp.first_byte_mask[x >> 5] >> (x & 31)) & 1u) FB |= 1u << b;
        }
    }
    FB &= valid;
    // non-ASCII bytes one by one: decode, class table, strict validation
    uint32_t CONT = 0;
    for (uint32_t m = HI; m; m &= m - 1) {
        const uint32_t b = __ffs(m) - 1;
        uint64_t lead; uint32_t len, cp;
        if (!char_containing(P.text, c0 + b, P.n, &lead, &len, &cp)) { set_status(P.status, ST_INVALID_UTF8); continue; }
        const uint32_t k = cp_class(P.cls, cp);
        if (k == CLS_L) L |= 1u << b; else if (k == CLS_N) N |= 1u << b; else if (k == CLS_S) S |= 1u << b;
        if (lead != c0 + b) CONT |= 1u << b;
    }
    ChunkClasses c;
    c.L = L; c.N = N; c.S = S; c.SP = SP; c.NL = NL; c.AP = AP; c.CONT = CONT; c.HI = HI; c.FB = FB;
    return c;
}

// Classes of a chunk's last / first four bytes, packed for the neighbour: 4 bits each of L, N, S, SP, NL, AP.
__device__ __forceinline__ uint32_t edge_hi(const ChunkClasses& c) {
    return (c.L >> 28) | ((c.N >> 28) << 4) | ((c.S >> 28) << 8) | ((c.SP >> 28) << 12) | ((c.NL >> 28) << 16) | ((c.AP >> 28) << 20);
}
__device__ __forceinline__ uint32_t edge_lo(const ChunkClasses& c) {
    return (c.L & 15u) | ((c.N & 15u) << 4) | ((c.S & 15u) << 8) | ((c.SP & 15u) << 12) | ((c.NL & 15u) << 16) | ((c.AP & 15u) << 20);
}

// ---- phase B: one special-token candidate at bit b of a chunk --------------------------------------------------------
// `before` = candidate first bytes in the 64 bytes in front of the chunk (bit 63 = the byte just before it), only
// meaningful when have_before.  Returns the token's length, 0 if it is not a match of regex.split's scan.
__device__ __forceinline__ uint32_t special_match(const PretokParams& P, const Specials& sp, uint64_t c0, uint32_t b, uint32_t FB, uint64_t before,
                                                  bool have_before) {
    const uint64_t p = c0 + b;
    const uint32_t len = special_len_at(P, sp, p);
    if (!len) return 0;
    // another candidate within max_len - 1 bytes in front of p could cover it: then replay the scan exactly
    bool lonely = false;
    if (have_before) {
        const uint64_t own = FB & ((1u << b) - 1u);
        const uint32_t reach = (uint32_t)sp.max_len - 1;  // <= 59
        uint64_t near_prev = reach > b ? before >> (64 - (reach - b)) : 0;  // the top (reach - b) bits
        uint64_t near_own = reach >= b ? own : own >> (b - reach);
        lonely = !(near_prev | near_own);
    }
    if (!lonely && !special_is_real(P, sp, p)) return 0;
    return len;
}

// ---- phase C: the split pattern over 32 bytes given the neighbours' edges and the special-token walls ------------------
// c: classes of this chunk; ph / nl: edge_hi of the chunk in front, edge_lo of the chunk behind; wall / Wp / Wn_raw: bytes
// covered by special tokens in this chunk and its neighbours; sps: special tokens that start in this chunk.
// Returns the token-start mask of the chunk (special-token starts included, bytes inside special tokens excluded).
template <bool NLMODE>
__device__ __forceinline__ uint32_t split_chunk(const PretokParams& P, const ChunkClasses& c, uint64_t chunk, uint64_t c0, uint32_t valid, uint32_t ph_in,
                                                uint32_t nl, uint32_t wall, uint32_t Wp_in, uint32_t Wn_raw, uint32_t sps) {
    const bool first = chunk == 0;
    const uint32_t ph = first ? 0u : ph_in;
    const uint32_t W = wall | ~valid;
    const uint32_t Wp = first ? 0xFFFFFFFFu : Wp_in;
    const uint64_t n_next = (chunk + 1) * 32;
    const uint32_t validn = n_next >= P.n ? 0u : (P.n - n_next >= 32 ? 0xFFFFFFFFu : ((1u << (uint32_t)(P.n - n_next)) - 1u));
    const uint32_t Wn = Wn_raw | ~validn;
    uint32_t L = c.L & ~W, N = c.N & ~W, S = c.S & ~W, SP = c.SP & ~W, AP = c.AP & ~W;
    // 40-bit views: bit 0..3 = the four bytes in front of the chunk, 4..35 = the chunk, 36..39 = the four behind it
    auto wide = [](uint32_t prev4, uint32_t core, uint32_t next4) { return (uint64_t)prev4 | ((uint64_t)core << 4) | ((uint64_t)next4 << 36); };
    const uint64_t W64 = wide(Wp >> 28, W, Wn & 15u);
    const uint64_t L64 = wide(ph & 15u, L, nl & 15u) & ~W64, N64 = wide((ph >> 4) & 15u, N, (nl >> 4) & 15u) & ~W64;
    const uint64_t S64 = wide((ph >> 8) & 15u, S, (nl >> 8) & 15u) & ~W64, SP64 = wide((ph >> 12) & 15u, SP, (nl >> 12) & 15u) & ~W64;
    uint64_t PN64 = (W64 << 1) | W64;  // bit j: a segment boundary between bytes j - 1 and j (a wall on either side, a cut)
    if (first) PN64 |= 0x1Full;
    if (P.cutmask) {
        const uint32_t cut = __ldg(P.cutmask + (c0 >> 5)), cutp = first ? 0u : __ldg(P.cutmask + (c0 >> 5) - 1), cutn = __ldg(P.cutmask + (c0 >> 5) + 1);
        PN64 |= wide(cutp >> 28, cut, cutn & 15u);
    }
    if (NLMODE) PN64 |= wide((ph >> 16) & 15u, c.NL, (nl >> 16) & 15u) << 1;
    const uint32_t PN = (uint32_t)(PN64 >> 4), NN = (uint32_t)(PN64 >> 5);
    const uint32_t pL = (uint32_t)(L64 >> 3), pN = (uint32_t)(N64 >> 3), pS = (uint32_t)(S64 >> 3), pSP = (uint32_t)(SP64 >> 3);
    const uint32_t O = ~(L | N | S | W);
    const uint32_t pO = ~(pL | pN | pS);  // only read where the previous byte exists and is no wall
    const uint32_t nonS = L | N | O;
    const uint32_t nNonS = ~(uint32_t)(S64 >> 5) & ~NN;  // the next byte exists and is not whitespace
    const uint32_t diff = (L & ~pL) | (N & ~pN) | (O & ~pO);
    const uint32_t startNS = nonS & (PN | (~pSP & (pS | diff)));
    const uint32_t startS = S & (PN | ~pS | nNonS);
    uint32_t start = (startNS | startS) & ~c.CONT;

    // multi-byte whitespace (NBSP, U+2000.., U+3000): "followed by non-whitespace" looks past the whole character
    for (uint32_t m = S & c.HI & ~c.CONT & ~start; m; m &= m - 1) {
        const uint32_t b = __ffs(m) - 1;
        uint64_t lead; uint32_t len, cp;
        if (!char_containing(P.text, c0 + b, P.n, &lead, &len, &cp)) continue;
        const uint32_t nx = b + len + 4;  // 40-bit index of the next character
        if (!((PN64 >> nx) & 1u) && c0 + b + len < P.n && class_containing(P, c0 + b + len) != CLS_S) start |= 1u << b;
    }

    // contractions: an apostrophe that starts a token swallows s d m t / ll ve re
    uint64_t AP64 = (uint64_t)AP << 4;
    for (uint32_t m = (ph >> 20) & 14u; m; m &= m - 1) AP64 |= 1ull << (__ffs(m) - 1);  // apostrophes 3, 2, 1 bytes in front
    AP64 &= ~W64;
    for (; AP64; AP64 &= AP64 - 1) {
        const uint32_t a = __ffsll((long long)AP64) - 1;  // 1 .. 35
        const bool starts = ((PN64 >> a) & 1u) || (((L64 | N64 | (S64 & ~SP64)) >> (a - 1)) & 1u);
        if (!starts || ((PN64 >> (a + 1)) & 1u)) continue;
        const uint64_t pos = c0 + a - 4;  // byte position of the apostrophe
        const uint32_t b1 = P.text[pos + 1];
        uint32_t after = 0;  // 40-bit index of the first byte behind the contraction
        if (b1 == 's' || b1 == 'd' || b1 == 'm' || b1 == 't') after = a + 2;
        else if (!((PN64 >> (a + 2)) & 1u)) {
            const uint32_t b2 = P.text[pos + 2];
            if ((b1 == 'l' && b2 == 'l') || (b1 == 'v' && b2 == 'e') || (b1 == 'r' && b2 == 'e')) after = a + 3;
        }
        if (!after) continue;
        if (a + 1 >= 4 && a + 1 < 36) start &= ~(1u << (a + 1 - 4));  // the letter belongs to the apostrophe
        if (after >= 4 && after < 36 && !((PN64 >> after) & 1u) && ((L64 >> after) & 1u)) start |= 1u << (after - 4);
    }
    return (start & ~W) | sps;
}

// ---- the warp window ----------------------------------------------------------------------------------------------------
// Lane L of the calling warp holds w[8] = the 32 bytes of chunk `chunk` (zeros when the chunk lies outside the text).
// Returns the lane's token-start mask; exact for lanes PT_LG .. 30, and for bits 0 .. PT_RELIABLE of lane 31.
// *sps_out = special tokens that start in the chunk.  All 32 lanes must call.
template <bool NLMODE>
__device__ __forceinline__ uint32_t window_split(const PretokParams& P, const Specials& sp, const SpecialFirst& sf, int64_t chunk, uint32_t (&w)[8],
                                                 uint32_t lane, uint32_t* sps_out) {
    const bool exists = chunk >= 0 && (uint64_t)chunk * 32 < P.n;
    const uint64_t c0 = exists ? (uint64_t)chunk * 32 : 0;
    const uint32_t valid = !exists ? 0u : (P.n - c0 >= 32 ? 0xFFFFFFFFu : ((1u << (uint32_t)(P.n - c0)) - 1u));
    ChunkClasses c = classify_chunk<NLMODE>(P, sp, sf, w, c0, valid);
    const uint32_t ehi = edge_hi(c), elo = edge_lo(c);
    const uint32_t ph = __shfl_up_sync(PT_FULL, ehi, 1), nl = __shfl_down_sync(PT_FULL, elo, 1);
    uint32_t wall = 0, spill1 = 0, spill2 = 0, sps = 0;
    if (sp.n) {  // warp-uniform
        const uint32_t fb1 = __shfl_up_sync(PT_FULL, c.FB, 1), fb2 = __shfl_up_sync(PT_FULL, c.FB, 2);
        const uint64_t before = (uint64_t)fb2 | ((uint64_t)fb1 << 32);
        for (uint32_t m = c.FB; m; m &= m - 1) {
            const uint32_t b = __ffs(m) - 1;
            const uint32_t len = special_match(P, sp, c0, b, c.FB, before, lane >= 2);
            if (!len) continue;
            sps |= 1u << b;
            const uint64_t bits = ((1ull << len) - 1ull);  // len <= 60: spans at most three chunks
            const uint64_t lo64 = bits << b;
            wall |= (uint32_t)lo64;
            if (b + len > 32) spill1 |= (uint32_t)(lo64 >> 32);
            if (b + len > 64) spill2 |= (uint32_t)(bits >> (64 - b));
        }
        const uint32_t t1 = __shfl_up_sync(PT_FULL, spill1, 1), t2 = __shfl_up_sync(PT_FULL, spill2, 2);
        if (lane >= 1) wall |= t1;
        if (lane >= 2) wall |= t2;
    }
    const uint32_t Wp = __shfl_up_sync(PT_FULL, wall, 1), Wn = __shfl_down_sync(PT_FULL, wall, 1);
    *sps_out = sps;
    if (!exists) return 0u;
    return split_chunk<NLMODE>(P, c, (uint64_t)chunk, c0, valid, lane ? ph : 0u, lane < 31 ? nl : 0u, wall, lane ? Wp : 0u, lane < 31 ? Wn : 0u, sps);
}
