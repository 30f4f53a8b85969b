// fused.cu — one pass over the text that splits it into pretokens, hashes every pretoken and counts it: the reference's
// pretokenize_chunk + the fold over workers (bpe_transformer/tokenization/preprocessing/pretokenization.py:73-111, 171-252)
// without the token-start mask ever going to HBM and back.
//
// A warp takes windows of 928 bytes (pretok.cuh: 29 core lanes of 32 bytes + 3 ghost lanes).  Per window:
//   1. The window's 1 KiB arrives in shared memory by ONE bulk copy (cp.async.bulk, the TMA engine) that was issued while
//      the previous window was being worked on; two buffers per warp take turns.
//   2. window_split: the token-start mask of every lane's 32 bytes, in registers.
//   3. The token starts of the 29 core lanes are written, in text order, as 16-bit offsets to a per-warp list in shared
//      memory (the end of a token is the next entry).
//   4. The list is worked off 32 tokens at a time with all lanes busy: 64-bit key from the token's bytes in the shared-
//      memory copy of the window (eight bytes per step; most pretokens need one), then the key's home slot in the CTA's
//      shared-memory table — where the hot pretokens of Zipf-distributed text sit, because they arrived first — one
//      8-byte load, one compare, one shared atomic.  Tokens that are not there go, compacted, to a second ring that is worked
//      off 32 at a time: further probing, admission of new keys (the global table is told about the pretoken and its
//      bytes), or a direct update of the global table when the CTA's table has no room.
//   5. At the end the CTA adds its table to the global one.
// A token that reaches beyond the window's right ghost lane lets its warp walk on (find_token_end) until it ends.
// HBM traffic: N bytes of text read; everything else lives in shared memory / L2.
#include "pretok.cuh"

#ifndef FC_WARPS
#define FC_WARPS 20
#endif
#define FC_THREADS (FC_WARPS * 32)
#define FC_SLOTS 8192            // per-CTA table: 8 B key + 4 B count
#define FC_ADMIT (FC_SLOTS / 2)  // new keys are admitted while fewer than this many are in the table
#define FC_PROBES 8
#define FC_LIST 1024             // token starts of one window (at most 29 * 32 + 1)
#define FC_SECOND 64             // second ring: 31 left over + 32 new
#define FC_STAGE_BYTES 1056      // a window (1024) plus room for the 8-byte reads behind a token that ends at its last byte

namespace {

struct FusedParams : PretokParams {
    uint32_t* start;           // optional: the token-start mask is written too (tests, callers that want it)
    uint64_t n_windows;
    uint64_t windows_per_cta;  // a CTA owns a contiguous range: positions relative to its first byte fit 32 bits
    UniqueTable ut;
};

struct WarpArea {
    uint4 stage[2][FC_STAGE_BYTES / 16];  // two windows of text: one being worked on, one arriving
    unsigned short list[FC_LIST + 32];     // offsets (from the window's first byte) of the token starts, in text order
    uint4 second[FC_SECOND];               // {key lo, key hi, position relative to the CTA's range, length}
    unsigned long long bar[2];             // mbarriers the bulk copies signal
};

struct FusedShared {
    unsigned long long key[FC_SLOTS];
    unsigned int cnt[FC_SLOTS];
    WarpArea wa[FC_WARPS];
    unsigned long long sp_key[BPE_MAX_SPECIALS];
    unsigned int used;
};

__device__ __forceinline__ void load_chunk(const PretokParams& P, int64_t chunk, uint32_t (&w)[8]) {
    uint4 a = make_uint4(0, 0, 0, 0), b = a;
    if (chunk >= 0 && (uint64_t)chunk * 32 < P.n) {  // the buffer is padded (BPE_TEXT_PAD): both halves are readable
        const uint4* q = reinterpret_cast<const uint4*>(P.text + (uint64_t)chunk * 32);
        a = __ldg(q);
        b = __ldg(q + 1);
    }
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
}

// ---- staging a window through shared memory: 1-D bulk copy (cp.async.bulk, the TMA engine) + mbarrier -------------------
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_addr(bar)), "r"(parity) : "memory");
}

// One lane: bring the bytes of window `win` that exist (the text plus its readable padding) into `stage`; the copy
// completes on `bar`.  The window spans [win * 928 - 64, win * 928 + 960); what lies outside the text stays as it was —
// nothing looks at those bytes.
__device__ __forceinline__ void stage_window(const PretokParams& P, uint64_t win, uint4* stage, unsigned long long* bar) {
    const int64_t wb = (int64_t)win * PT_WIN_BYTES - 32 * PT_LG;
    const int64_t lo = wb < 0 ? 0 : wb;
    const int64_t limit = (int64_t)((P.n + 48) & ~15ull);  // inside the BPE_TEXT_PAD bytes behind the text, a multiple of 16
    const int64_t hi = wb + 1024 < limit ? wb + 1024 : limit;
    const uint32_t bytes = (uint32_t)(hi - lo);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_addr(reinterpret_cast<unsigned char*>(stage) + (lo - wb))),
                 "l"(P.text + lo), "r"(bytes), "r"(smem_addr(bar))
                 : "memory");
}

// Key of the `len` bytes at offset `off` of a staged window (off + len <= 1024): hash_key() with the bytes read from shared
// memory.  Up to 16 bytes — nearly every pretoken — in straight-line code that all lanes run together: three aligned 8-byte
// loads, funnel shifts to the token's first byte, a mask behind its last one, one or two mixing steps.
__device__ __forceinline__ uint64_t hash_key_staged(const uint4* stage, uint32_t off, uint32_t len) {
    const unsigned char* base = reinterpret_cast<const unsigned char*>(stage);
    const uint2* q = reinterpret_cast<const uint2*>(base + (off & ~7u));
    const uint2 a = q[0], b = q[1], c = q[2];
    const bool up = (off & 4u) != 0;           // the token starts in the upper half of the first 8 bytes
    const uint32_t d0 = up ? a.y : a.x, d1 = up ? b.x : a.y, d2 = up ? b.y : b.x, d3 = up ? c.x : b.y, d4 = up ? c.y : c.x;
    const uint32_t sh = (off & 3u) * 8u;
    uint32_t w0 = __funnelshift_r(d0, d1, sh), w1 = __funnelshift_r(d1, d2, sh), w2 = __funnelshift_r(d2, d3, sh), w3 = __funnelshift_r(d3, d4, sh);
    uint64_t h = key_seed(len);
    if (len <= 16) {
        // zero the bytes behind the token: byte j (0..15) survives iff j < len.  Branch-free: a clamped funnel shift of
        // 0xFFFFFFFF by 32 - 8 * (bytes of the token that lie in the word).
        const int L = (int)len;
        const uint32_t k0 = __funnelshift_rc(0xFFFFFFFFu, 0u, 32u - 8u * (uint32_t)min(L, 4));
        const uint32_t k1 = __funnelshift_rc(0xFFFFFFFFu, 0u, 32u - 8u * (uint32_t)min(max(L - 4, 0), 4));
        const uint32_t k2 = __funnelshift_rc(0xFFFFFFFFu, 0u, 32u - 8u * (uint32_t)min(max(L - 8, 0), 4));
        const uint32_t k3 = __funnelshift_rc(0xFFFFFFFFu, 0u, 32u - 8u * (uint32_t)min(max(L - 12, 0), 4));
        w0 &= k0; w1 &= k1; w2 &= k2; w3 &= k3;
        h = mix64(h, w0 | ((uint64_t)w1 << 32));
        if (len > 8) h = mix64(h, w2 | ((uint64_t)w3 << 32));
        return key_finish(h);
    }
    // longer tokens: the general loop, eight bytes per step
    h = mix64(h, w0 | ((uint64_t)w1 << 32));
    h = mix64(h, w2 | ((uint64_t)w3 << 32));
    const uint32_t shift = (off & 7u) * 8u;
    const unsigned long long* q8 = reinterpret_cast<const unsigned long long*>(base + (off & ~7u)) + 2;
    unsigned long long lo = q8[0];
#pragma unroll 1
    for (uint32_t o = 16; o < len;) {
        const unsigned long long hi = q8[1];
        const uint64_t w = shift ? (lo >> shift) | (hi << (64u - shift)) : lo;
        if (o + 8 <= len) {
            h = mix64(h, w);
            o += 8;
        } else {
            h = mix64(h, w & ((1ull << ((len - o) * 8)) - 1ull));
            break;
        }
        lo = hi;
        q8++;
    }
    return key_finish(h);
}

// The first token start at or behind byte `chunk` * 32 (or the end of the text): where a token that is still open at the end
// of a window's right ghost lane ends.  Walks on window by window; warp-uniform; all lanes call.
template <bool NLMODE>
__device__ __noinline__ uint64_t find_token_end(const PretokParams& P, const Specials& sp, int64_t chunk) {
    const uint32_t lane = threadIdx.x & 31;
    const SpecialFirst sf = special_first(sp);
    for (;;) {
        const int64_t mine = chunk - PT_LG + lane;  // a window whose first core lane is `chunk`: lanes 2 .. 30 are exact
        uint32_t w[8];
        load_chunk(P, mine, w);
        uint32_t sps;
        uint32_t mask = window_split<NLMODE>(P, sp, sf, mine, w, lane, &sps);
        const uint64_t c0 = mine >= 0 ? (uint64_t)mine * 32 : 0;
        if (mine >= 0 && P.n >= c0 && P.n - c0 < 32) mask |= 1u << (uint32_t)(P.n - c0);  // the end of the text ends the token
        const bool core = lane >= PT_LG && lane < 32 - PT_RG;
        const uint32_t stops = __ballot_sync(PT_FULL, core && mask != 0);
        if (stops) {
            const uint32_t l = (uint32_t)__ffs(stops) - 1u;
            const uint32_t m = __shfl_sync(PT_FULL, mask, l);
            return (uint64_t)(chunk - PT_LG + l) * 32 + ((uint32_t)__ffs(m) - 1u);
        }
        chunk += PT_CORE;
        if ((uint64_t)chunk * 32 > P.n + 32) return P.n;  // cannot happen (the end of the text stops the walk); keeps a bug from spinning
    }
}

__device__ __forceinline__ bool unique_bump_key(const UniqueTable& t, uint64_t key, uint64_t c) {
    uint64_t slot = (key * 0xD6E8FEB86659FD93ull >> 17) & t.cap_mask;
    for (int probe = 0; probe < 64; probe++) {
        uint64_t cur = *reinterpret_cast<volatile uint64_t*>(t.key + slot);
        if (cur == key) {
            if (c) atomicAdd(reinterpret_cast<unsigned long long*>(t.cnt + slot), (unsigned long long)c);
            return true;
        }
        if (cur == 0) return false;
        slot = (slot + 1) & t.cap_mask;
    }
    return false;
}

// Global table: `c` more occurrences of a key that is (normally) already there; first sighting -> insert with its bytes.
// c = 0 "announces" a key the CTA sees for the first time: the global table then knows the pretoken and its bytes, the
// CTA's count arrives with the fold at the end.
static __device__ __noinline__ void global_count(const FusedParams& P, uint64_t key, uint64_t abs_pos, uint32_t len, uint64_t c) {
    if (unique_bump_key(P.ut, key, c)) return;
    TokenKey k;
    k.key = key;
    unique_insert(P.ut, P.status, P.text, abs_pos, len, k, c);
}

// The second ring: tokens whose key was not in its home slot of the CTA's table.  32 at a time, with all lanes busy, each
// lane walks the probe sequence: found further on -> counted there; an empty slot while the table still admits keys ->
// the key moves in (and the global table is told about the pretoken and its bytes, count 0: the CTA's count arrives with
// the fold at the end); otherwise the occurrence goes to the global table directly.  Returns what is left in the ring.
static __device__ __noinline__ uint32_t flush_second(const FusedParams& P, FusedShared* shp, const uint4* ring2, uint64_t cta_base, uint32_t n2, bool all) {
    const uint32_t lane = threadIdx.x & 31;
    FusedShared& sh = *shp;
    while (n2 >= 32 || (all && n2)) {
        const uint32_t take = n2 >= 32 ? 32 : n2;
        const uint32_t base = n2 - take;
        if (lane < take) {
            const uint4 r = ring2[base + lane];
            const uint64_t key = r.x | ((uint64_t)r.y << 32);
            uint32_t slot = (uint32_t)(key >> 40) & (FC_SLOTS - 1);
            int outcome = 0;  // 0: no room -> global count; 1: counted in the CTA's table; 2: moved in and counted -> announce
#pragma unroll 1
            for (int probe = 0; probe < FC_PROBES; probe++) {
                unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(&sh.key[slot]);
                if (cur == 0) {
                    // an empty slot ends the search.  New keys are admitted only while the table is at most half full: a fuller
                    // table would make every lookup walk long probe chains.
                    if (*reinterpret_cast<volatile unsigned int*>(&sh.used) >= FC_ADMIT) break;
                    cur = atomicCAS(&sh.key[slot], 0ull, (unsigned long long)key);
                    if (cur == 0) {
                        atomicAdd(&sh.used, 1u);
                        outcome = 2;
                        cur = key;
                    }
                }
                if (cur == key) {
                    atomicAdd(&sh.cnt[slot], 1u);
                    if (!outcome) outcome = 1;
                    break;
                }
                slot = (slot + 1) & (FC_SLOTS - 1);
            }
            if (outcome != 1) global_count(P, key, cta_base + r.z, r.w, outcome ? 0 : 1);
        }
        n2 = base;
        __syncwarp();
    }
    return n2;
}

template <bool NLMODE, bool WRITE_MASK>
__global__ void __launch_bounds__(FC_THREADS, 1) fused_count_kernel(const __grid_constant__ FusedParams P, const __grid_constant__ Specials sp) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    FusedShared& sh = *reinterpret_cast<FusedShared*>(smem_raw);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpArea& wa = sh.wa[warp];
    for (uint32_t k = threadIdx.x; k < FC_SLOTS; k += FC_THREADS) { sh.key[k] = 0; sh.cnt[k] = 0; }
    if (threadIdx.x < BPE_MAX_SPECIALS) {  // keys of the special tokens, from their 8-byte words (kernel parameters)
        unsigned long long key = 0;
        if ((int)threadIdx.x < sp.n) {
            const uint32_t len = sp.off[threadIdx.x + 1] - sp.off[threadIdx.x];
            uint64_t h = key_seed(len);
            for (uint32_t k = 0; 8 * k < len; k++) h = mix64(h, sp.words[threadIdx.x][k]);  // zero padded: the last word is already masked
            key = key_finish(h);
        }
        sh.sp_key[threadIdx.x] = key;
    }
    if (threadIdx.x == 0) sh.used = 0;
    if (lane == 0) { mbar_init(&wa.bar[0], 1); mbar_init(&wa.bar[1], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();

    const SpecialFirst sf = special_first(sp);
    const uint64_t win_lo = (uint64_t)blockIdx.x * P.windows_per_cta;
    const uint64_t win_hi = win_lo + P.windows_per_cta < P.n_windows ? win_lo + P.windows_per_cta : P.n_windows;
    const uint64_t cta_base = win_lo * PT_WIN_BYTES;
    const bool core = lane >= PT_LG && lane < 32 - PT_RG;
    uint32_t n2 = 0;       // tokens in the second ring
    uint64_t tokens = 0;   // counted by this warp (the same number in every lane), plus per-lane extras
    uint32_t trips = 0;

    if (lane == 0 && win_lo + warp < win_hi) stage_window(P, win_lo + warp, wa.stage[0], &wa.bar[0]);
    for (uint64_t win = win_lo + warp; win < win_hi; win += FC_WARPS, trips++) {
        // the call is lost once the table or a pool is full (the host enlarges them and counts again): do not grind on
        // (looked at every 16th window: the status word lives in global memory)
        if ((trips & 15u) == 0 && __any_sync(PT_FULL, (*reinterpret_cast<volatile uint32_t*>(P.status) & (ST_TABLE_FULL | ST_POOL_FULL)) != 0)) break;
        const uint32_t buf = trips & 1u;
        const uint4* stage = wa.stage[buf];
        const int64_t chunk = (int64_t)win * PT_CORE - PT_LG + lane;
        const uint64_t c0 = chunk >= 0 ? (uint64_t)chunk * 32 : 0;
        const int64_t wb = (int64_t)win * PT_WIN_BYTES - 32 * PT_LG;  // text position of the window's first byte
        // the next window starts its way into the other buffer (whose readers finished a whole window ago), this one has landed
        if (lane == 0 && win + FC_WARPS < win_hi) stage_window(P, win + FC_WARPS, wa.stage[buf ^ 1u], &wa.bar[buf ^ 1u]);
        mbar_wait(&wa.bar[buf], (trips >> 1) & 1u);
        uint32_t w[8];
        {
            uint4 a = stage[2 * lane], b = stage[2 * lane + 1];
            if (!(chunk >= 0 && c0 < P.n)) a = b = make_uint4(0, 0, 0, 0);
            w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
        }
        uint32_t sps;
        uint32_t M = window_split<NLMODE>(P, sp, sf, chunk, w, lane, &sps);
        if (WRITE_MASK && core && chunk >= 0 && c0 < P.n) P.start[c0 >> 5] = M;
        if (chunk >= 0 && P.n >= c0 && P.n - c0 < 32) M |= 1u << (uint32_t)(P.n - c0);  // the end of the text ends the last token

        // ---- the token starts of the core lanes, in text order; the entry behind the last one = where the last token ends ---------
        const uint32_t mine = core ? (uint32_t)__popc(M) : 0u;
        uint32_t incl = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t y = __shfl_up_sync(PT_FULL, incl, d);
            if (lane >= (uint32_t)d) incl += y;
        }
        const uint32_t T = __shfl_sync(PT_FULL, incl, 31);
        if (T) {  // warp-uniform
            uint32_t at = incl - mine;
            for (uint32_t m = core ? M : 0u; m; m &= m - 1) wa.list[at++] = (unsigned short)(lane * 32 + (__ffs(m) - 1));
            // the last token ends at the right ghost's first start, if that bit is reliable; else the warp walks on
            const uint32_t m31 = __shfl_sync(PT_FULL, M, 31) & ((2u << PT_RELIABLE) - 1u);
            uint64_t end_abs = 0;  // != 0: the last token ends outside the window, here
            if (m31) {
                if (lane == 0) wa.list[T] = (unsigned short)(31 * 32 + (__ffs(m31) - 1));
            } else {
                end_abs = find_token_end<NLMODE>(P, sp, chunk - lane + 31);
                if (lane == 0) wa.list[T] = 0xFFFFu;
            }
            __syncwarp();
            // ---- 32 tokens at a time ----------------------------------------------------------------------------------------------
#pragma unroll 1
            for (uint32_t t0 = 0; t0 < T; t0 += 32) {
                const uint32_t t = t0 + lane;
                bool live = t < T;
                uint32_t s = 0, e = 0;
                if (live) { s = wa.list[t]; e = wa.list[t + 1]; }
                const uint64_t abs_s = (uint64_t)(wb + (int64_t)s);
                live = live && abs_s < P.n;  // the end of the text is in the list as a start (it ends the last token), not as a token
                uint64_t key = 0;
                uint32_t len = e - s;
                if (live) {
                    if (e <= 1024u) key = hash_key_staged(stage, s, len);
                    else {  // reaches beyond the window: its bytes come from global memory
                        const uint64_t full = end_abs - abs_s;
                        len = full < 0xFFFFFFFFull ? (uint32_t)full : 0xFFFFFFFFu;
                        key = hash_key(P.text, abs_s, full);
                    }
                }
                if (sp.n) {  // a special token is a token like any other up to here; training drops it (pretokenization.py:228-232).
                             // It is recognised by its key, like every pretoken on this path (what a 64-bit key collision would do).
#pragma unroll 1
                    for (int k = 0; k < sp.n; k++) live = live && key != sh.sp_key[k];
                }
                tokens += __popc(__ballot_sync(PT_FULL, live));
                const uint32_t slot = (uint32_t)(key >> 40) & (FC_SLOTS - 1);
                const bool hit = live && *reinterpret_cast<volatile unsigned long long*>(&sh.key[slot]) == key;
                if (hit) atomicAdd(&sh.cnt[slot], 1u);
                const bool push = live && !hit;
                const uint32_t mm = __ballot_sync(PT_FULL, push);
                if (mm) {
                    if (push) wa.second[n2 + __popc(mm & ((1u << lane) - 1u))] = make_uint4((uint32_t)key, (uint32_t)(key >> 32), (uint32_t)(abs_s - cta_base), len);
                    n2 += __popc(mm);
                    __syncwarp();
                    if (n2 >= 32) n2 = flush_second(P, &sh, wa.second, cta_base, n2, false);
                }
            }
        }
        __syncwarp();  // everyone is done with this window's buffer and list before the next trip reuses them
    }
    // what is left in the second ring
    __syncwarp();
    n2 = flush_second(P, &sh, wa.second, cta_base, n2, true);
    __syncthreads();
    // fold the CTA's table into the context's table (every key was announced there when it entered this one)
    if (!(*reinterpret_cast<volatile uint32_t*>(P.status) & (ST_TABLE_FULL | ST_POOL_FULL))) {
        for (uint32_t s = threadIdx.x; s < FC_SLOTS; s += FC_THREADS) {
            const unsigned long long key = sh.key[s];
            if (key && sh.cnt[s]) unique_bump_key(P.ut, key, sh.cnt[s]);
        }
    }
    if (lane == 0 && tokens) atomicAdd(reinterpret_cast<unsigned long long*>(P.ut.counters + 2), (unsigned long long)tokens);
}


}  // namespace

int launch_fused_count(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint32_t* d_cutmask, uint32_t flags, uint32_t* d_start, cudaStream_t s) {
    if (d_start) CUDA_TRY(ctx, cudaMemsetAsync(d_start + (n / 32), 0, 2 * sizeof(uint32_t), s));
    if (n == 0) return BPE_OK;
    FusedParams P{};
    P.text = d_text; P.n = n; P.cutmask = d_cutmask; P.flags = flags; P.status = ctx->d_status; P.cls = ctx->cls;
    P.start = d_start; P.ut = ctx->ut;
    P.n_windows = (n + PT_WIN_BYTES - 1) / PT_WIN_BYTES;
    uint64_t grid = (uint64_t)BPE_N_SM;  // one CTA per SM (the shared table and the per-warp areas take ~180 KB)
    uint64_t per = (P.n_windows + grid - 1) / grid;
    per = (per + FC_WARPS - 1) / FC_WARPS * FC_WARPS;  // every warp of a CTA gets the same number of windows
    grid = (P.n_windows + per - 1) / per;
    P.windows_per_cta = per;
    if (per * PT_WIN_BYTES >= 0xFFFF0000ull) return ctx_fail(ctx, BPE_ERR_ARGUMENT, "text too large for one call (more than 600 GB)");
    const bool nl = (flags & BPE_CUT_AFTER_NEWLINE) != 0;
    auto kern = nl ? (d_start ? fused_count_kernel<true, true> : fused_count_kernel<true, false>)
                   : (d_start ? fused_count_kernel<false, true> : fused_count_kernel<false, false>);
    const int variant = (nl ? 2 : 0) + (d_start ? 1 : 0);
    if (!(ctx->fused_attr_set & (1u << variant))) {  // per context = per device: function attributes do not carry over to another GPU
        CUDA_TRY(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FusedShared)));
        ctx->fused_attr_set |= 1u << variant;
    }
    KernelTimer t(ctx, "fused_count_kernel", s);
    kern<<<(unsigned)grid, FC_THREADS, sizeof(FusedShared), s>>>(P, ctx->sp);
    CUDA_TRY(ctx, cudaGetLastError());
    return BPE_OK;
}

cudaError_t set_key_mask_fused(unsigned long long mask) { return cudaMemcpyToSymbol(c_key_and, &mask, sizeof(mask)); }
