// fused.cu — one pass over the text that splits it into pretokens, hashes every pretoken and counts it: the reference's
// pretokenize_chunk + the fold over workers (bpe_transformer/tokenization/preprocessing/pretokenization.py:73-111, 171-252)
// without the token-start mask ever going to HBM and back.
//
// A warp takes windows of 928 bytes (pretok.cuh: 29 core lanes of 32 bytes + 3 ghost lanes).  Per window:
//   1. window_split: the token-start mask of every lane's 32 bytes, in registers.
//   2. Every lane runs Horner's rule (common.cuh: H = H * R + byte, 64 bits) over its own 32 bytes and restarts at
//      each token start; a token that ends inside the lane is written, as one 16-byte record {H, position, length}, to a
//      per-warp ring in shared memory.  The running values at the lane borders ("tails") give, with one shuffle, the
//      state in which a token enters the next lane; the token that was cut by a lane border is completed from that
//      state and the next lane's first record: H_left * R^n + H_right.
//   3. The ring is drained 32 records at a time with all lanes busy: key = finalise(H, length), look the key up in
//      the CTA's shared-memory table (the hot pretokens of Zipf-distributed text live there), one shared atomic per
//      token.  A key the CTA has not seen goes to the context's global table at once (which also copies the bytes of a
//      globally new pretoken into the pool); keys that find no room in the shared table are collected in a second
//      ring and bumped in the global table 32 at a time.
//   4. At the end the CTA adds its table to the global one.
// A token longer than the window's ghost lane lets its warp walk on (extend_token) until it ends.
// HBM traffic: N bytes of text read; everything else lives in shared memory / L2.
#include <cstdio>

#include "pretok.cuh"

#define FC_WARPS 16
#define FC_THREADS (FC_WARPS * 32)
#define FC_SLOTS 8192          // per-CTA table: 8 B key + 4 B count
#define FC_ADMIT (FC_SLOTS / 2)  // new keys are admitted while fewer than this many are in the table
#define FC_PROBES 8
#define FC_RING 272            // records: 31 left over + at most 29 lanes * 8 per slab, rounded up
#define FC_MISS 64
#define REC_VALID 0x10000u     // record flag: the token starts in a core lane of its window

namespace {

struct FusedParams : PretokParams {
    uint32_t* start;           // optional: the token-start mask is written too (tests, callers that want it)
    uint64_t n_windows;
    uint64_t windows_per_cta;  // a CTA owns a contiguous range: positions relative to its first byte fit 32 bits
    UniqueTable ut;
};

struct WarpRings {
    uint4 stage[64];     // the warp's next window of text (1 KiB), brought in by one bulk copy (TMA) while the current one is processed
    uint4 ring[FC_RING];
    uint4 head[64];      // [lane]: hash of the bytes in front of the lane's first token start; [32 ..]: scratch for the ghost lanes' records
    uint4 miss[FC_MISS];
    unsigned long long bar;  // mbarrier the bulk copy signals
    unsigned long long pad_;
};

struct FusedShared {
    unsigned long long key[FC_SLOTS];
    unsigned int cnt[FC_SLOTS];
    WarpRings wr[FC_WARPS];
    unsigned long long rpow[33];    // R^k
    unsigned long long r32pow[33];  // (R^32)^k
    unsigned long long sp_key[BPE_MAX_SPECIALS];
    unsigned int used;
};

__device__ __forceinline__ uint64_t mul_rpow(const FusedShared& sh, uint64_t h, uint64_t n) {
    if (n < 33) return h * sh.rpow[n];
    if (n < 33 * 32) return h * sh.r32pow[n >> 5] * sh.rpow[n & 31];
    return poly_shift(h, n);
}

// One lane's 32 bytes up to its first token start: hash and count of the bytes in front of it (all 32 when the mask is 0).
__device__ __forceinline__ void lane_head(const uint32_t (&w)[8], uint32_t mask, uint64_t* h_out, uint32_t* n_out) {
    uint64_t h = 0;
    bool open = true;
#pragma unroll
    for (int i = 0; i < 32; i++) {
        if ((mask >> i) & 1u) open = false;
        if (open) h = poly_step(h, __byte_perm(w[i >> 2], 0, 0x4440 + (i & 3)));
    }
    *h_out = h;
    *n_out = mask ? (uint32_t)__ffs(mask) - 1u : 32u;
}

// Inclusive segmented fold over the lanes: element = (hash, bytes, stop).  After the call lane L holds the hash / byte
// count of the run of bytes that ends with lane L's element and began at the last lane <= L whose `stop` is set
// (or at lane 0).  Exclusive variant = shuffle up by one.
__device__ __forceinline__ void warp_fold(const FusedShared& sh, uint32_t lane, uint64_t& h, uint64_t& n, bool& stop) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint64_t h2 = __shfl_up_sync(PT_FULL, h, d), n2 = __shfl_up_sync(PT_FULL, n, d);
        const bool s2 = __shfl_up_sync(PT_FULL, (int)stop, d) != 0;
        if (lane >= (uint32_t)d && !stop) {
            h = mul_rpow(sh, h2, n) + h;
            n += n2;
            stop = s2;
        }
    }
}

// Bytes behind the end of the text (the buffer is padded, BPE_TEXT_PAD) are loaded as they are: the classification masks
// them out and no token reaches them, because the end of the text counts as a token start.
__device__ __forceinline__ void load_chunk(const PretokParams& P, int64_t chunk, uint32_t (&w)[8]) {
    uint4 a = make_uint4(0, 0, 0, 0), b = a;
    if (chunk >= 0 && (uint64_t)chunk * 32 < P.n) {
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
// the lanes that own those chunks never look at their bytes.
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

// The token that is open when a window's core ends and does not end in the right ghost lane: walk on, window by
// window, until a token start shows up.  (h, n) = hash and bytes of the token up to `chunk` * 32.  Warp-uniform.
template <bool NLMODE>
__device__ __noinline__ void extend_token(const FusedParams& P, const Specials& sp, const FusedShared& sh, int64_t chunk, uint64_t& h, uint64_t& n) {
    const uint32_t lane = threadIdx.x & 31;
    const SpecialFirst sf = special_first(sp);
    for (;;) {
        // a window whose first core lane is `chunk`: lanes 2 .. 30 are exact
        const int64_t mine = chunk - PT_LG + lane;
        uint32_t w[8];
        load_chunk(P, mine, w);
        uint32_t sps;
        uint32_t mask = window_split<NLMODE>(P, sp, sf, mine, w, lane, &sps);
        const uint64_t c0 = mine >= 0 ? (uint64_t)mine * 32 : 0;
        if (mine >= 0 && P.n >= c0 && P.n - c0 < 32) mask |= 1u << (uint32_t)(P.n - c0);  // the end of the text ends the token
        uint64_t hh; uint32_t hn;
        lane_head(w, mask, &hh, &hn);
        const bool core = lane >= PT_LG && lane < 32 - PT_RG;
        const uint32_t stops = __ballot_sync(PT_FULL, core && mask != 0);
        const uint32_t last = stops ? (uint32_t)__ffs(stops) - 1u : 31 - PT_RG;  // fold the core lanes up to the first one with a start
        for (uint32_t l = PT_LG; l <= last; l++) {  // rare path: a plain sequential fold
            const uint64_t lh = __shfl_sync(PT_FULL, hh, l);
            const uint32_t ln = __shfl_sync(PT_FULL, hn, l);
            h = mul_rpow(sh, h, ln) + lh;
            n += ln;
        }
        if (stops) return;
        chunk += PT_CORE;
        if ((uint64_t)chunk * 32 >= P.n + 32) return;  // cannot happen (the end of the text stops the walk); keeps a bug from spinning
    }
}

// Per-warp state of the draining side.  Everything here stays in registers: the functions that take it are inlined; the
// out-of-line slow paths below get plain values (and references to the kernel's __grid_constant__ parameters).
struct DrainCtx {
    FusedShared& sh;
    WarpRings& wr;
    const FusedParams& P;
    const Specials& sp;
    uint64_t cta_base;
    uint32_t fb;  // first bytes of the special tokens (special_first_bytes)
    uint32_t n_miss;
    uint64_t tokens;
};

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

// The second ring: records whose key was not in its home slot of the CTA's table.  32 at a time, with all lanes busy, each
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
            const uint32_t len = r.w & 0xFFFFu;
            const uint64_t key = key_finalize(r.x | ((uint64_t)r.y << 32), len);
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
            if (outcome != 1) global_count(P, key, cta_base + r.z, len, outcome ? 0 : 1);
        }
        n2 = base;
        __syncwarp();
    }
    return n2;
}

// One round: lane takes record `idx` of the ring (active = idx is a record).  All 32 lanes call.
// Straight-line code for the common case — the key sits in its home slot of the CTA's table, which is where the hot
// pretokens of Zipf text are, because they arrived first: one 8-byte load, one compare, one shared atomic.  Everything else
// is handed, compacted, to the second ring.
__device__ __forceinline__ void drain_round(DrainCtx& C, uint32_t lane, uint32_t idx, bool active) {
    uint4 r = make_uint4(0, 0, 0, 0);
    if (active) r = C.wr.ring[idx];
    bool live = active && (r.w & REC_VALID);
    const uint32_t len = r.w & 0xFFFFu;
    const uint64_t key = key_finalize(r.x | ((uint64_t)r.y << 32), len);
    if (C.sp.n) {  // a special token is a token like any other up to here; training drops it (pretokenization.py:228-232)
        bool maybe = false;
#pragma unroll 1
        for (int s = 0; s < C.sp.n; s++) maybe |= key == C.sh.sp_key[s];
        if (live && maybe && token_is_special(C.sp, C.fb, C.P.text, C.cta_base + r.z, len)) live = false;
    }
    C.tokens += __popc(__ballot_sync(PT_FULL, live));  // the same number in every lane
    const uint32_t slot = (uint32_t)(key >> 40) & (FC_SLOTS - 1);
    const bool hit = live && *reinterpret_cast<volatile unsigned long long*>(&C.sh.key[slot]) == key;
    if (hit) atomicAdd(&C.sh.cnt[slot], 1u);
    const bool push = live && !hit;
    const uint32_t mm = __ballot_sync(PT_FULL, push);
    if (mm) {
        if (push) C.wr.miss[C.n_miss + __popc(mm & ((1u << lane) - 1u))] = r;
        C.n_miss += __popc(mm);
        __syncwarp();
        if (C.n_miss >= 32) C.n_miss = flush_second(C.P, &C.sh, C.wr.miss, C.cta_base, C.n_miss, false);
    }
}

// Drains full groups of 32 from the top of the ring; what is left (< 32 records) stays at the bottom.
__device__ __forceinline__ uint32_t drain_full(DrainCtx& C, uint32_t lane, uint32_t total) {
    __syncwarp();
    while (total >= 32) {
        total -= 32;
        drain_round(C, lane, total + lane, true);
    }
    return total;
}

// One byte of the per-lane sweep over a slab of 8 bytes (w0 / w1 = the slab's two words, Ms = its 8 mask bits, i0 = index of
// its first byte in the lane).  At a token start the token that ends here is written to ring slot `ri` — the lane's
// first one goes to its head slot, because it began in an earlier lane and is completed later — and Horner restarts.
// Branch-free: a predicated 16-byte store and selects, so the lanes of the warp stay together.
#define SWEEP_BYTE(i, word)                                                                                             \
    {                                                                                                                   \
        const uint32_t byte_ = __byte_perm(word, 0, 0x4440 + ((i) & 3));                                                \
        const bool st_ = (Ms & (1u << (i))) != 0;                                                                       \
        if (st_) slots[ri] = make_uint4(hlo, hhi, posv - tb, i0 + (uint32_t)(i) + tb);                                  \
        ri = st_ ? nx : ri;                                                                                             \
        nx += st_ ? 1u : 0u;                                                                                            \
        tb = st_ ? vfi - (uint32_t)(i) : tb;                                                                            \
        const uint32_t mul_ = st_ ? 0u : BPE_HASH_R;                                                                    \
        const uint64_t t_ = (uint64_t)hlo * mul_ + byte_;                                                               \
        hhi = hhi * mul_ + (uint32_t)(t_ >> 32);                                                                        \
        hlo = (uint32_t)t_;                                                                                             \
    }

template <bool NLMODE, bool WRITE_MASK>
__global__ void __launch_bounds__(FC_THREADS, 1) fused_count_kernel(const __grid_constant__ FusedParams P, const __grid_constant__ Specials sp) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    FusedShared& sh = *reinterpret_cast<FusedShared*>(smem_raw);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t k = threadIdx.x; k < FC_SLOTS; k += FC_THREADS) { sh.key[k] = 0; sh.cnt[k] = 0; }
    if (threadIdx.x < 33) {
        unsigned long long a = 1, b = 1, r32 = 1;
        for (int k = 0; k < 32; k++) r32 *= BPE_HASH_R;
        for (uint32_t k = 0; k < threadIdx.x; k++) { a *= BPE_HASH_R; b *= r32; }
        sh.rpow[threadIdx.x] = a;
        sh.r32pow[threadIdx.x] = b;
    }
    if (threadIdx.x < BPE_MAX_SPECIALS) {  // keys of the special tokens (their bytes are kernel parameters: plain byte reads)
        unsigned long long h = 0;
        uint32_t len = 0;
        if ((int)threadIdx.x < sp.n) {
            len = sp.off[threadIdx.x + 1] - sp.off[threadIdx.x];
            for (uint32_t k = 0; k < len; k++) h = poly_step(h, sp.blob[sp.off[threadIdx.x] + k]);
        }
        sh.sp_key[threadIdx.x] = len ? key_finalize(h, len) : 0ull;
    }
    if (threadIdx.x == 0) sh.used = 0;
    if (lane == 0) mbar_init(&sh.wr[warp].bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();

    const SpecialFirst sf = special_first(sp);
    WarpRings& wr = sh.wr[warp];
    uint4* const slots = wr.ring;  // head slots follow the ring: index FC_RING + lane
    const uint64_t win_lo = (uint64_t)blockIdx.x * P.windows_per_cta;
    const uint64_t win_hi = win_lo + P.windows_per_cta < P.n_windows ? win_lo + P.windows_per_cta : P.n_windows;
    DrainCtx C{sh, wr, P, sp, win_lo * PT_WIN_BYTES, special_first_bytes(sp), 0u, 0ull};
    const bool core = lane >= PT_LG && lane < 32 - PT_RG;
    const uint32_t vflag = core ? REC_VALID : 0u;
    uint32_t total = 0;  // records in the ring
    uint64_t long_tokens = 0;  // per lane: tokens that went straight to the global table

    uint32_t w[8];
    uint32_t parity = 0, trips = 0;
#ifndef FC_NO_STAGE
    if (lane == 0 && win_lo + warp < win_hi) stage_window(P, win_lo + warp, wr.stage, &wr.bar);
#endif
    for (uint64_t win = win_lo + warp; win < win_hi; win += FC_WARPS) {
        // the call is lost once the table or a pool is full (the host enlarges them and counts again): do not grind on
        // (looked at every 16th window: the status word lives in global memory)
        if ((trips++ & 15u) == 0 && __any_sync(PT_FULL, (*reinterpret_cast<volatile uint32_t*>(P.status) & (ST_TABLE_FULL | ST_POOL_FULL)) != 0)) break;
        const int64_t chunk = (int64_t)win * PT_CORE - PT_LG + lane;
        const uint64_t c0 = chunk >= 0 ? (uint64_t)chunk * 32 : 0;
#ifdef FC_NO_STAGE
        load_chunk(P, chunk, w);
#else
        {   // this window has landed in shared memory: 32 bytes per lane into registers, then the buffer is free for the next one
            mbar_wait(&wr.bar, parity);
            parity ^= 1u;
            uint4 a = wr.stage[2 * lane], b = wr.stage[2 * lane + 1];
            if (!(chunk >= 0 && c0 < P.n)) a = b = make_uint4(0, 0, 0, 0);
            w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
#ifdef FC_DEBUG_STAGE
            {
                uint32_t g[8];
                load_chunk(P, chunk, g);
                const uint64_t left = (chunk >= 0 && c0 < P.n) ? P.n - c0 : 0;
                for (int k = 0; k < 8; k++) {
                    uint32_t keep = left >= 4u * k + 4 ? 0xFFFFFFFFu : (left <= 4u * k ? 0u : (1u << (8 * (uint32_t)(left - 4u * k))) - 1u);
                    if ((g[k] ^ w[k]) & keep) { printf("stage mismatch win %llu lane %u word %d chunk %lld n %llu got %08x want %08x trip %u\n", (unsigned long long)win, lane, k, (long long)chunk, (unsigned long long)P.n, w[k], g[k], trips); break; }
                }
            }
#endif
            // The next copy overwrites the buffer: every lane's loads must have RETURNED before it is issued, not just have been
            // issued (a warp-level barrier alone does not wait for them).  A vote whose input depends on the loaded words does.
            const uint32_t landed = __ballot_sync(PT_FULL, (a.x ^ b.w) == 0x9E3779B9u);
            asm volatile("" ::"r"(landed));
            if (lane == 0 && win + FC_WARPS < win_hi) stage_window(P, win + FC_WARPS, wr.stage, &wr.bar);
        }
#endif
        uint32_t sps;
        uint32_t M = window_split<NLMODE>(P, sp, sf, chunk, w, lane, &sps);
        if (WRITE_MASK && core && chunk >= 0 && c0 < P.n) P.start[c0 >> 5] = M;
        if (chunk >= 0 && P.n >= c0 && P.n - c0 < 32) M |= 1u << (uint32_t)(P.n - c0);  // the end of the text ends the last token
        if (lane == 31 && (M & ((2u << PT_RELIABLE) - 1u)) == 0) M = 0;                  // right ghost: only its reliable bits count

        // ---- ring slots: per slab of 8 bytes, lane-major; the lane's first record goes to its head slot -----------
        const uint32_t first = M ? (uint32_t)__ffs(M) - 1u : 32u;
        uint32_t cnts = (uint32_t)__popc(M & 0xFFu) | ((uint32_t)__popc(M & 0xFF00u) << 8) | ((uint32_t)__popc(M & 0xFF0000u) << 16) |
                        ((uint32_t)__popc(M >> 24) << 24);
        if (first < 32) cnts -= 1u << (8 * (first >> 3));
        if (!core) cnts = 0;  // ghost lanes put nothing into the ring (the right ghost contributes its head record, later)
        uint32_t incl = cnts;      // four 8-bit prefix sums at once (a slab holds at most 29 * 8 = 232 records)
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t y = __shfl_up_sync(PT_FULL, incl, d);
            if (lane >= (uint32_t)d) incl += y;
        }
        const uint32_t excl = incl - cnts, tot4 = __shfl_sync(PT_FULL, incl, 31);

        uint32_t hlo = 0, hhi = 0;
        uint32_t tb = 0;                                                // flags - start of the open token (byte index in the lane)
        const uint32_t posv = (uint32_t)(c0 - C.cta_base) + vflag;     // record position = posv - tb (relative to the CTA's range)
        uint32_t Ms = M, exs = excl, tots = tot4, i0 = 0;
        __syncwarp();
        // Steps 0..3: one slab of 8 bytes each, then drain.  Step 4: the tokens cut by lane borders, then drain.  (One loop, so
        // that the draining code exists once: the kernel's hot loop has to fit the instruction cache.)
#pragma unroll 1
        for (int step = 0; step < 5; step++) {
            uint32_t added;
            if (step < 4) {
                const uint32_t base = total + (exs & 0xFFu);
                uint32_t ri, nx;
                if (first < i0) { ri = base; nx = base + 1; } else { ri = FC_RING + lane; nx = base; }
                if (!core) {  // ghost lanes: their records are not wanted (up to 8 per slab land in a scratch area); the right ghost keeps its head
                    const uint32_t scratch = FC_RING + 32 + 10 * (lane == 31 ? 2 : lane);
                    ri = (lane == 31 && first >= i0) ? FC_RING + 31 : scratch;
                    nx = scratch;
                }
                const uint32_t vfi = vflag - i0;
                SWEEP_BYTE(0, w[0]) SWEEP_BYTE(1, w[0]) SWEEP_BYTE(2, w[0]) SWEEP_BYTE(3, w[0])
                SWEEP_BYTE(4, w[1]) SWEEP_BYTE(5, w[1]) SWEEP_BYTE(6, w[1]) SWEEP_BYTE(7, w[1])
                added = tots & 0xFFu;
                w[0] = w[2]; w[1] = w[3]; w[2] = w[4]; w[3] = w[5]; w[4] = w[6]; w[5] = w[7];
                Ms >>= 8; exs >>= 8; tots >>= 8; i0 += 8;
            } else {
                // tail = the open token at the end of the lane: (hash, bytes since its start); a lane without any start passes
                // everything it has on.  in-state of lane L = fold of the tails in front of it.
                const bool has = M != 0;
                uint64_t th = hlo | ((uint64_t)hhi << 32), tn = has ? 32u - ((vflag - tb) & 0xFFFFu) : 32u;
                bool stop = has;
                if (__any_sync(PT_FULL, !has)) warp_fold(sh, lane, th, tn, stop);
                uint64_t ih = __shfl_up_sync(PT_FULL, th, 1), in = __shfl_up_sync(PT_FULL, tn, 1);
                const bool in_stop = __shfl_up_sync(PT_FULL, (int)stop, 1) != 0;
                if (lane == 0) { ih = 0; in = 0; }
                // the entering token began at c0 - in; it is this window's if that lies in the core (and a start was seen at all)
                const uint64_t core_lo = win * PT_WIN_BYTES;
                const bool mine = lane >= PT_LG && in_stop && c0 >= in && c0 - in >= core_lo && c0 - in < P.n;
                bool emit = has && mine;
                uint64_t H = 0, Ln = 0;
                if (emit) {
                    const uint4 hr = slots[FC_RING + lane];
                    H = mul_rpow(sh, ih, first) + (hr.x | ((uint64_t)hr.y << 32));
                    Ln = in + first;
                    if (Ln >= 0xFFFFu) {  // too long for a ring record: straight to the global table
                        if (!(sp.n && token_is_special(sp, C.fb, P.text, c0 - in, Ln))) { long_tokens++; global_count(P, key_finalize(H, Ln), c0 - in, (uint32_t)Ln, 1); }
                        emit = false;
                    }
                }
                const uint32_t em = __ballot_sync(PT_FULL, emit);
                if (emit) slots[total + __popc(em & ((1u << lane) - 1u))] = make_uint4((uint32_t)H, (uint32_t)(H >> 32), (uint32_t)(c0 - in - C.cta_base), (uint32_t)Ln | REC_VALID);
                added = __popc(em);
                // the token that is still open where the right ghost's reliable bits end (lane 31: no reliable start, and the
                // token entering it began in this window's core): walk on until it ends
                const bool open_end = __shfl_sync(PT_FULL, (int)(!has && mine), 31) != 0;
                if (open_end) {
                    uint64_t eh = __shfl_sync(PT_FULL, ih, 31), en = __shfl_sync(PT_FULL, in, 31);
                    const uint64_t begin = (uint64_t)(chunk - lane + 31) * 32 - en;  // all lanes: chunk - lane + 31 = lane 31's chunk
                    extend_token<NLMODE>(P, sp, sh, chunk - lane + 31, eh, en);
                    if (lane == 0 && !(sp.n && token_is_special(sp, C.fb, P.text, begin, en))) {
                        long_tokens++;
                        global_count(P, key_finalize(eh, en), begin, (uint32_t)en, 1);
                    }
                }
            }
            total = drain_full(C, lane, total + added);
        }
    }
    // what is left in the rings
    __syncwarp();
    drain_round(C, lane, lane, lane < total);
    C.n_miss = flush_second(P, &sh, wr.miss, C.cta_base, C.n_miss, true);
    __syncthreads();
    // fold the CTA's table into the context's table (every key was announced there when it entered this one)
    if (!(*reinterpret_cast<volatile uint32_t*>(P.status) & (ST_TABLE_FULL | ST_POOL_FULL))) {
        for (uint32_t s = threadIdx.x; s < FC_SLOTS; s += FC_THREADS) {
            const unsigned long long key = sh.key[s];
            if (key && sh.cnt[s]) unique_bump_key(P.ut, key, sh.cnt[s]);
        }
    }
    for (int o = 16; o; o >>= 1) long_tokens += __shfl_down_sync(PT_FULL, long_tokens, o);
    const uint64_t tokens = C.tokens + long_tokens;  // C.tokens is warp-uniform
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
    uint64_t grid = (uint64_t)BPE_N_SM;  // one CTA of 16 warps per SM (the shared table and the rings take ~190 KB)
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
