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

#define SCAN_LG 2                                    // ghost chunks in front of a tile's core (special tokens up to 60 bytes reach in)
#define SCAN_RG 1                                    // ghost chunk behind it
#define SCAN_CORE (SCAN_THREADS - SCAN_LG - SCAN_RG)  // chunks of 32 bytes a tile produces

// A tile is SCAN_THREADS chunks of 32 bytes, one per thread; the first two and the last one are ghosts: they are
// classified like the others (so that their neighbours can read the classes next to their own bytes from shared
// memory instead of classifying extra words) but their result belongs to the neighbouring tile.
template <bool NLMODE>
__global__ void __launch_bounds__(SCAN_THREADS) scan_kernel(const ScanParams P, const __grid_constant__ Specials sp) {
    __shared__ uint32_t wall[SCAN_THREADS + 2];  // bytes covered by a special token, per chunk of the tile
    __shared__ uint32_t sps[SCAN_THREADS + 2];   // first byte of a special token
    __shared__ uint32_t e_hi[SCAN_THREADS];      // classes of a chunk's last four bytes: 4 bits each of L, N, S, SP, NL, RARE
    __shared__ uint32_t e_lo[SCAN_THREADS];      // ... of its first four bytes
    __shared__ uint32_t fbm[SCAN_THREADS];       // bytes that can begin a special token
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

    for (uint64_t tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
        const int64_t chunk = (int64_t)tile * SCAN_CORE - SCAN_LG + tid;  // may lie in front of the text (first tile)
        const bool exists = chunk >= 0 && (uint64_t)chunk * 32 < P.n;
        const uint64_t c0 = exists ? (uint64_t)chunk * 32 : 0;
        const uint32_t valid = !exists ? 0u : (P.n - c0 >= 32 ? 0xFFFFFFFFu : ((1u << (uint32_t)(P.n - c0)) - 1u));
        wall[tid] = 0; sps[tid] = 0;
        if (tid < 2) { wall[SCAN_THREADS + tid] = 0; sps[SCAN_THREADS + tid] = 0; }

        // ---- phase A: 32 bytes -> class masks -------------------------------------------------------------------
        uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (exists) {
            const uint4* q = reinterpret_cast<const uint4*>(P.text + c0);
            uint4 a = __ldg(q), b = make_uint4(0, 0, 0, 0);
            if (c0 + 16 < P.n) b = __ldg(q + 1);
            w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
        }
        uint32_t L = 0, N = 0, S = 0, SP = 0, NL = 0, RARE = 0, FB = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint32_t hi = w[k] & 0x80808080u, y = w[k] & 0x7F7F7F7Fu, ascii = ~hi;
            const uint32_t cL = sw_range(y | 0x20202020u, 'a', 'z') & ascii;
            const uint32_t cN = sw_range(y, '0', '9') & ascii;
            const uint32_t cSP = sw_eq(y, ' ') & ascii;
            const uint32_t cS = (sw_range(y, 9, 13) & ascii) | cSP;
            const uint32_t cRare = (sw_eq(y, '\'') & ascii) | hi;  // apostrophes and non-ASCII bytes: handled one by one
            L |= sw_collapse(cL) << (4 * k);
            N |= sw_collapse(cN) << (4 * k);
            S |= sw_collapse(cS) << (4 * k);
            SP |= sw_collapse(cSP) << (4 * k);
            RARE |= sw_collapse(cRare) << (4 * k);
            if (NLMODE) NL |= sw_collapse(sw_eq(y, '\n') & ascii) << (4 * k);
            if (sp.n) FB |= sw_collapse((sw_eq(y, fb0) | sw_eq(y, fb1)) & ascii) << (4 * k);
        }
        L &= valid; N &= valid; S &= valid; SP &= valid; RARE &= valid; NL &= valid;
        if (fb_generic) {
            FB = 0;
            for (uint32_t b = 0; b < 32; b++) {
                uint32_t x = (w[b >> 2] >> (8 * (b & 3))) & 0xFF;
                if ((sp.first_byte_mask[x >> 5] >> (x & 31)) & 1u) FB |= 1u << b;
            }
        }
        FB &= valid;
        // rare bytes one by one: non-ASCII (decode, class table, strict validation) and apostrophes
        uint32_t AP = 0, CONT = 0, HI = 0;
        for (uint32_t m = RARE; m; m &= m - 1) {
            const uint32_t b = __ffs(m) - 1;
            const uint32_t byte = (w[b >> 2] >> (8 * (b & 3))) & 0xFF;
            if (byte == '\'') { AP |= 1u << b; continue; }
            HI |= 1u << b;
            uint64_t lead; uint32_t len, cp;
            if (!char_containing(P.text, c0 + b, P.n, &lead, &len, &cp)) { set_status(P.status, ST_INVALID_UTF8); continue; }
            const uint32_t k = cp_class(P.cls, cp);
            if (k == CLS_L) L |= 1u << b; else if (k == CLS_N) N |= 1u << b; else if (k == CLS_S) S |= 1u << b;
            if (lead != c0 + b) CONT |= 1u << b;
        }
        e_hi[tid] = (L >> 28) | ((N >> 28) << 4) | ((S >> 28) << 8) | ((SP >> 28) << 12) | ((NL >> 28) << 16) | ((AP >> 28) << 20);
        e_lo[tid] = (L & 15u) | ((N & 15u) << 4) | ((S & 15u) << 8) | ((SP & 15u) << 12) | ((NL & 15u) << 16) | ((AP & 15u) << 20);
        fbm[tid] = FB;
        __syncthreads();

        // ---- phase B: special tokens that start in this chunk ---------------------------------------------------------
        if (sp.n) {
            for (uint32_t m = FB; m; m &= m - 1) {
                const uint32_t b = __ffs(m) - 1;
                const uint64_t p = c0 + b;
                const uint32_t len = special_len_at(P, sp, p);
                if (!len) continue;
                // another candidate within max_len - 1 bytes in front of p could cover it: then replay the scan exactly
                bool lonely = false;
                if (tid >= 2) {
                    const uint64_t before = ((uint64_t)fbm[tid - 2]) | ((uint64_t)fbm[tid - 1] << 32);  // 64 bytes in front of the chunk
                    const uint64_t own = FB & ((1u << b) - 1u);
                    // bits of the 96-byte window [c0 - 64, c0 + 32) that lie in [p - (max_len - 1), p)
                    const uint32_t reach = (uint32_t)sp.max_len - 1;  // <= 59
                    uint64_t near_prev = reach > b ? before >> (64 - (reach - b)) : 0;  // the top (reach - b) bits
                    uint64_t near_own = reach >= b ? own : own >> (b - reach);
                    lonely = !(near_prev | near_own);
                }
                if (!lonely && !special_is_real(P, sp, p)) continue;
                atomicOr(&sps[tid], 1u << b);
                const uint64_t bits = ((1ull << len) - 1ull);  // len <= 60: spans at most three chunks
                const uint64_t lo64 = bits << b;
                atomicOr(&wall[tid], (uint32_t)lo64);
                if (b + len > 32) atomicOr(&wall[tid + 1], (uint32_t)(lo64 >> 32));
                if (b + len > 64) atomicOr(&wall[tid + 2], (uint32_t)(bits >> (64 - b)));
            }
        }
        __syncthreads();

        // ---- phase C: the split pattern, 32 bytes at a time -------------------------------------------------------------------
        if (exists && tid >= SCAN_LG && tid < SCAN_THREADS - SCAN_RG) {
            const bool first = chunk == 0;
            const uint32_t ph = first ? 0u : e_hi[tid - 1], nl = e_lo[tid + 1];
            const uint32_t W = wall[tid] | ~valid;
            const uint32_t Wp = first ? 0xFFFFFFFFu : wall[tid - 1];
            const uint64_t n_next = (uint64_t)(chunk + 1) * 32;
            const uint32_t validn = n_next >= P.n ? 0u : (P.n - n_next >= 32 ? 0xFFFFFFFFu : ((1u << (uint32_t)(P.n - n_next)) - 1u));
            const uint32_t Wn = wall[tid + 1] | ~validn;
            L &= ~W; N &= ~W; S &= ~W; SP &= ~W; AP &= ~W;
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
            if (NLMODE) PN64 |= wide((ph >> 16) & 15u, NL, (nl >> 16) & 15u) << 1;
            const uint32_t PN = (uint32_t)(PN64 >> 4), NN = (uint32_t)(PN64 >> 5);
            const uint32_t pL = (uint32_t)(L64 >> 3), pN = (uint32_t)(N64 >> 3), pS = (uint32_t)(S64 >> 3), pSP = (uint32_t)(SP64 >> 3);
            const uint32_t O = ~(L | N | S | W);
            const uint32_t pO = ~(pL | pN | pS);  // only read where the previous byte exists and is no wall
            const uint32_t nonS = L | N | O;
            const uint32_t nNonS = ~(uint32_t)(S64 >> 5) & ~NN;  // the next byte exists and is not whitespace
            const uint32_t diff = (L & ~pL) | (N & ~pN) | (O & ~pO);
            const uint32_t startNS = nonS & (PN | (~pSP & (pS | diff)));
            const uint32_t startS = S & (PN | ~pS | nNonS);
            uint32_t start = (startNS | startS) & ~CONT;

            // multi-byte whitespace (NBSP, U+2000.., U+3000): "followed by non-whitespace" looks past the whole character
            for (uint32_t m = S & HI & ~CONT & ~start; m; m &= m - 1) {
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
            P.start[c0 >> 5] = (start & ~W) | sps[tid];
        }
        __syncthreads();  // shared arrays are rewritten by the next tile
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
    CUDA_TRY(ctx, cudaMemsetAsync(d_start + (n / 32), 0, 2 * sizeof(uint32_t), s));
    if (n == 0) return BPE_OK;
    ScanParams P;
    P.text = d_text; P.n = n; P.cutmask = d_cutmask; P.flags = flags; P.start = d_start; P.status = ctx->d_status; P.cls = ctx->cls;
    P.n_tiles = (n + SCAN_CORE * 32 - 1) / (SCAN_CORE * 32);
    uint64_t grid = P.n_tiles < (uint64_t)BPE_N_SM * 8 ? P.n_tiles : (uint64_t)BPE_N_SM * 8;
    KernelTimer t(ctx, "scan_kernel", s);
    if (flags & BPE_CUT_AFTER_NEWLINE) scan_kernel<true><<<(unsigned)grid, SCAN_THREADS, 0, s>>>(P, ctx->sp);
    else scan_kernel<false><<<(unsigned)grid, SCAN_THREADS, 0, s>>>(P, ctx->sp);
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
