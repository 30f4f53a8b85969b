// common.cuh — shared device helpers and the context layout of libbpe_b200.so (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/bpe_b200.h"

#define BPE_N_SM 148  // B200: 2 dies x 74 SMs; persistent grids are sized in multiples of this

// ---- status word written by kernels (one per context, in device memory) --------------------------------
enum : uint32_t {
    ST_INVALID_UTF8 = 1u << 0,
    ST_TABLE_FULL = 1u << 1,
    ST_POOL_FULL = 1u << 2,
    ST_COLLISION = 1u << 3,
    ST_MISSING_ID = 1u << 4,
    ST_PAIR_TABLE_FULL = 1u << 5,
    ST_LIST_FULL = 1u << 6,
    ST_WATCHDOG = 1u << 7,
    ST_ENC_MISS = 1u << 8,  // encode: a pretoken of the text is not in the table yet (handled inside bpe_encode)
    ST_HAS_CR = 1u << 10,   // bpe_encode with BPE_FAIL_ON_CR: the text holds a '\r'
    ST_BLOB_FULL = 1u << 9,  // train: the pool of token byte strings is full (handled inside bpe_train: larger pool, run again)
};

// ---- code-point classes (include/bpe_unicode_tables.h) --------------------------------------------------
enum : uint32_t { CLS_O = 0, CLS_L = 1, CLS_N = 2, CLS_S = 3 };

struct ClassTables {
    const uint8_t* stage1;   // BPE_UC_N_STAGE1 block ids
    const uint32_t* stage2;  // 16 words (256 code points x 2 bits) per block
    uint32_t n_stage1;
};

__device__ __forceinline__ uint32_t cp_class(const ClassTables& t, uint32_t cp) {
    uint32_t b = cp >> 8;
    if (b >= t.n_stage1) return CLS_O;
    uint32_t w = __ldg(t.stage2 + (uint32_t)__ldg(t.stage1 + b) * 16u + ((cp & 255u) >> 4));
    return (w >> (2u * (cp & 15u))) & 3u;
}

// ---- special tokens (few, short): passed by value to kernels ---------------------------------------------
#define BPE_MAX_SPECIALS 16
#define BPE_MAX_SPECIAL_BYTES 256
struct Specials {
    int n;
    int max_len;
    uint32_t off[BPE_MAX_SPECIALS + 1];
    uint32_t first_byte_mask[8];  // 256-bit set of first bytes
    uint32_t newline_inside;      // bit s: token s holds a '\n' before its last byte (a cut after every newline splits it)
    uint8_t blob[BPE_MAX_SPECIAL_BYTES];
    unsigned long long words[BPE_MAX_SPECIALS][8];  // token s as little-endian 64-bit words, zero padded (at most 60 bytes)
};

// ---- the 64-bit key of a byte string ------------------------------------------------------------------------------
// Internal to the library (the oracle compares byte strings, never hashes).  Eight bytes per round; the same function
// wherever a pretoken is named by key (count, import, encode lookup), whether its bytes come from global memory, from
// a window of text staged in shared memory (fused.cu) or from a kernel parameter (the special tokens).
__device__ __forceinline__ uint64_t mix64(uint64_t h, uint64_t w) {
    h = (h ^ w) * 0x9FB21C651E98DF25ull;
    return h ^ (h >> 35);
}
__device__ __forceinline__ uint64_t key_seed(uint64_t len) { return 0x243F6A8885A308D3ull ^ (len * 0x9E3779B97F4A7C15ull); }
// Test hook (bpe_debug_key_mask): keys are ANDed with this, so that a test can make distinct pretokens share a key and see
// that the library reports it (BPE_ERR_HASH_COLLISION) instead of merging their counts.  All ones in normal operation.
// One copy per translation unit (no relocatable device code): every .cu that hashes has a setter, api.cu calls them all.
static __constant__ unsigned long long c_key_and = ~0ull;

__device__ __forceinline__ uint64_t key_finish(uint64_t h) {
    h ^= h >> 29;
    h &= c_key_and;
    return h ? h : 1ull;  // never 0 (0 marks an empty slot)
}

// Unaligned little-endian 64-bit read through two aligned reads (the buffer is padded: BPE_TEXT_PAD).
__device__ __forceinline__ uint64_t load_u64_unaligned(const uint8_t* __restrict__ base, uint64_t p) {
    const uint64_t* q = reinterpret_cast<const uint64_t*>(base) + (p >> 3);
    uint32_t sh = (uint32_t)(p & 7) * 8u;
    uint64_t lo = __ldg(q);
    if (sh == 0) return lo;
    uint64_t hi = __ldg(q + 1);
    return (lo >> sh) | (hi << (64u - sh));
}

struct TokenKey {
    uint64_t key;  // never 0
};

__device__ __forceinline__ uint64_t hash_key(const uint8_t* __restrict__ base, uint64_t p, uint64_t len) {
    uint64_t h = key_seed(len), o = 0;
#pragma unroll 1
    for (; o + 8 <= len; o += 8) h = mix64(h, load_u64_unaligned(base, p + o));
    if (o < len) h = mix64(h, load_u64_unaligned(base, p + o) & ((1ull << ((len - o) * 8)) - 1ull));
    return key_finish(h);
}

// hash_key for the common case in straight-line code that all lanes of a warp run together (the loop above leaves most lanes
// idle while the longest token of the round finishes): up to 16 bytes through three aligned 8-byte loads, funnel shifts to
// the token's first byte and a mask behind its last one; longer tokens take the loop.  Same value as hash_key().
__device__ __forceinline__ uint64_t hash_key_short(const uint8_t* __restrict__ base, uint64_t p, uint64_t len) {
    if (len > 16) return hash_key(base, p, len);
    const uint2* q = reinterpret_cast<const uint2*>(base + (p & ~7ull));  // reads at most 24 bytes from there: inside BPE_TEXT_PAD
    const uint2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
    const bool up = (p & 4u) != 0;  // the token starts in the upper half of the first 8 bytes
    const uint32_t d0 = up ? a.y : a.x, d1 = up ? b.x : a.y, d2 = up ? b.y : b.x, d3 = up ? c.x : b.y, d4 = up ? c.y : c.x;
    const uint32_t sh = ((uint32_t)p & 3u) * 8u;
    uint32_t w0 = __funnelshift_r(d0, d1, sh), w1 = __funnelshift_r(d1, d2, sh), w2 = __funnelshift_r(d2, d3, sh), w3 = __funnelshift_r(d3, d4, sh);
    const int L = (int)len;  // byte j (0..15) survives iff j < len: a clamped funnel shift of 0xFFFFFFFF by 32 - 8 * (bytes of the token in the word)
    w0 &= __funnelshift_rc(0xFFFFFFFFu, 0u, 32u - 8u * (uint32_t)min(L, 4));
    w1 &= __funnelshift_rc(0xFFFFFFFFu, 0u, 32u - 8u * (uint32_t)min(max(L - 4, 0), 4));
    w2 &= __funnelshift_rc(0xFFFFFFFFu, 0u, 32u - 8u * (uint32_t)min(max(L - 8, 0), 4));
    w3 &= __funnelshift_rc(0xFFFFFFFFu, 0u, 32u - 8u * (uint32_t)min(max(L - 12, 0), 4));
    uint64_t h = mix64(key_seed(len), w0 | ((uint64_t)w1 << 32));
    if (len > 8) h = mix64(h, w2 | ((uint64_t)w3 << 32));
    return key_finish(h);
}

__device__ __forceinline__ TokenKey hash_bytes(const uint8_t* __restrict__ base, uint64_t p, uint64_t len) {
    TokenKey k;
    k.key = hash_key(base, p, len);
    return k;
}

// text[p, p + len) == special token s ?  Eight bytes per compare (the text buffer is padded, so the last word may be read whole).
__device__ __forceinline__ bool special_equals(const Specials& sp, int s, const uint8_t* __restrict__ text, uint64_t p, uint32_t len) {
    bool ok = true;
#pragma unroll 1
    for (uint32_t k = 0; 8 * k < len && ok; k++) {
        uint64_t w = load_u64_unaligned(text, p + 8 * k);
        if (8 * k + 8 > len) w &= (1ull << (8 * (len - 8 * k))) - 1ull;
        ok = w == sp.words[s][k];
    }
    return ok;
}

// ---- the table of distinct pretokens (context-owned, global memory) ---------------------------------------
struct UniqueTable {
    uint64_t* key;       // [cap] 0 = empty
    uint64_t* cnt;       // [cap] occurrences
    uint64_t* pool_off;  // [cap] offset of the bytes in `pool`
    uint32_t* len;       // [cap]
    uint32_t* chk;       // [cap] 0 until the inserting thread has published length and bytes, then 1
    uint32_t* enc_off;   // [cap] offset of the cached encoding in the id pool (encode path)
    uint32_t* enc_len;   // [cap] 0xFFFFFFFF = not encoded yet
    uint4* enc;          // [cap] what the emit pass reads, one 16-byte load per probe: {key lo, key hi, w0, w1}, key 0 = empty.
                         //       w1 bit 31 clear: w1 ids at id-pool offset w0.  bit 31 set: (w1 >> 16) & 0x3FFF ids stored in place:
                         //       bit 30 set -> the single id w0; else 16-bit ids w0 & 0xFFFF, w0 >> 16, w1 & 0xFFFF.
    uint8_t* pool;
    uint64_t cap_mask;
    uint64_t pool_cap;
    uint64_t* counters;  // [0] pool cursor, [1] distinct, [2] total occurrences, [3] id-pool cursor
};

__device__ __forceinline__ void set_status(uint32_t* status, uint32_t bit) { atomicOr(status, bit); }

// Do text[p, p + len) and the pool copy of `slot` hold the same bytes?  The caller has seen chk[slot] != 0.  Eight bytes per
// compare; the pool is read through L2 (another SM may have written it during this kernel), the text through the read-only path.
__device__ __forceinline__ bool unique_same_bytes(const UniqueTable& t, uint64_t slot, const uint8_t* __restrict__ text, uint64_t p, uint64_t len) {
    if (*reinterpret_cast<volatile uint32_t*>(t.len + slot) != (uint32_t)len) return false;
    const uint64_t off = *reinterpret_cast<volatile uint64_t*>(t.pool_off + slot);
    const uint64_t* q = reinterpret_cast<const uint64_t*>(t.pool) + (off >> 3);
    const uint32_t sh = (uint32_t)(off & 7) * 8u;
    uint64_t lo = __ldcg(q);
#pragma unroll 1
    for (uint64_t o = 0; o < len; o += 8) {
        const uint64_t hi = __ldcg(q + (o >> 3) + 1);  // the pool is padded (BPE_TEXT_PAD)
        uint64_t a = sh ? (lo >> sh) | (hi << (64u - sh)) : lo;
        uint64_t b = load_u64_unaligned(text, p + o);
        if (len - o < 8) {
            const uint64_t keep = (1ull << ((len - o) * 8)) - 1ull;
            a &= keep;
            b &= keep;
        }
        if (a != b) return false;
        lo = hi;
    }
    return true;
}

// Adds `c` occurrences of text[p, p+len) to the table (copies the bytes into the pool on first sight).
// Identity is the BYTES: a key that is already present with other bytes raises ST_COLLISION (the reference's Counter
// is keyed by the bytes themselves, pretokenization.py:201-207).  Returns the slot, or ~0 when the table is full.
__device__ __forceinline__ uint64_t unique_insert(const UniqueTable& t, uint32_t* status, const uint8_t* __restrict__ text,
                                                  uint64_t p, uint32_t len, TokenKey k, uint64_t c) {
    uint64_t slot = (k.key * 0xD6E8FEB86659FD93ull >> 17) & t.cap_mask;
    // A probe sequence this long means the table is far too full for open addressing: report it instead of walking all slots
    // (the host enlarges the table and counts again).  Tables of up to 4096 slots may be walked completely.
    const uint64_t probe_limit = t.cap_mask < 4096 ? t.cap_mask + 1 : 4096;
    for (uint64_t probe = 0; probe < probe_limit; probe++) {
        uint64_t cur = *reinterpret_cast<volatile uint64_t*>(t.key + slot);
        if (cur == 0) {
            cur = atomicCAS(reinterpret_cast<unsigned long long*>(t.key + slot), 0ull, (unsigned long long)k.key);
            if (cur == 0) {  // ours: copy the bytes, publish
                uint64_t off = atomicAdd(reinterpret_cast<unsigned long long*>(t.counters + 0), (unsigned long long)len);
                if (off + len > t.pool_cap) {
                    set_status(status, ST_POOL_FULL);
                    off = 0;
                } else {
                    for (uint32_t i = 0; i < len; i++) t.pool[off + i] = text[p + i];
                }
                t.pool_off[slot] = off;
                t.len[slot] = len;
                t.enc_len[slot] = 0xFFFFFFFFu;
                __threadfence();
                *reinterpret_cast<volatile uint32_t*>(t.chk + slot) = 1u;
                const unsigned long long n_before = atomicAdd(reinterpret_cast<unsigned long long*>(t.counters + 1), 1ull);
                // beyond 7/8 full (large tables only) the probe sequences explode: stop early, the host enlarges the table
                if (t.cap_mask >= 4096 && n_before > (t.cap_mask + 1) / 8 * 7) set_status(status, ST_TABLE_FULL);
                if (c) atomicAdd(reinterpret_cast<unsigned long long*>(t.cnt + slot), (unsigned long long)c);
                return slot;
            }
        }
        if (cur == k.key) {
            // the inserting thread publishes within a few hundred cycles; a reader that arrives earlier waits for it
            uint32_t ready = 0;
            for (int spin = 0; spin < 4096 && !(ready = *reinterpret_cast<volatile uint32_t*>(t.chk + slot)); spin++) __nanosleep(32);
            if (ready) {
                __threadfence();
                if (!(*reinterpret_cast<volatile uint32_t*>(status) & ST_POOL_FULL) && !unique_same_bytes(t, slot, text, p, len)) set_status(status, ST_COLLISION);
            }
            if (c) atomicAdd(reinterpret_cast<unsigned long long*>(t.cnt + slot), (unsigned long long)c);
            return slot;
        }
        slot = (slot + 1) & t.cap_mask;
    }
    set_status(status, ST_TABLE_FULL);
    return ~0ull;
}

// Slot of an already-counted pretoken, ~0 if absent.
__device__ __forceinline__ uint64_t unique_find(const UniqueTable& t, TokenKey k) {
    uint64_t slot = (k.key * 0xD6E8FEB86659FD93ull >> 17) & t.cap_mask;
    for (uint64_t probe = 0; probe <= t.cap_mask; probe++) {
        uint64_t cur = __ldg(t.key + slot);
        if (cur == k.key) return slot;
        if (cur == 0) return ~0ull;
        slot = (slot + 1) & t.cap_mask;
    }
    return ~0ull;
}

// fb = up to four distinct first bytes of the special tokens packed in a word (0xFF = unused), so that the common
// "not a special token" answer costs a few register compares instead of an indexed read of the kernel parameters.
// Index of the special token equal to text[p, p + len), -1 if it is ordinary text.
__device__ __forceinline__ int special_index(const Specials& sp, uint32_t fb, const uint8_t* __restrict__ text, uint64_t p, uint64_t len) {
    const uint32_t b0 = text[p];
    if (fb != 0xFFFFFFFEu) {
        if (b0 != (fb & 0xFF) && b0 != ((fb >> 8) & 0xFF) && b0 != ((fb >> 16) & 0xFF) && b0 != (fb >> 24)) return -1;
    } else if (!((sp.first_byte_mask[b0 >> 5] >> (b0 & 31)) & 1u)) return -1;
    for (int s = 0; s < sp.n; s++) {
        if (sp.off[s + 1] - sp.off[s] != len) continue;
        if (special_equals(sp, s, text, p, (uint32_t)len)) return s;
    }
    return -1;
}

__device__ __forceinline__ bool token_is_special(const Specials& sp, uint32_t fb, const uint8_t* __restrict__ text, uint64_t p, uint64_t len) {
    return special_index(sp, fb, text, p, len) >= 0;
}

__device__ __forceinline__ uint32_t special_first_bytes(const Specials& sp) {
    uint32_t fb = 0xFFFFFFFFu;
    int used = 0;
    for (int s = 0; s < sp.n; s++) {
        uint32_t b = sp.blob[sp.off[s]];
        bool seen = false;
        for (int k = 0; k < used; k++) seen |= ((fb >> (8 * k)) & 0xFF) == b;
        if (seen) continue;
        if (used == 4 || b == 0xFF) return 0xFFFFFFFEu;  // too many / ambiguous: use the 256-bit set
        fb = (fb & ~(0xFFu << (8 * used))) | (b << (8 * used));
        used++;
    }
    return fb;
}

// ---- bit helpers -----------------------------------------------------------------------------------------
// End (exclusive) of the token that starts at bit `b` of start word `w` (word index wi), scanning forward.
__device__ __forceinline__ uint64_t token_end(const uint32_t* __restrict__ start, uint64_t n_words, uint64_t wi, uint32_t w,
                                              uint32_t b, uint64_t n) {
    uint32_t rest = (b == 31) ? 0u : (w >> (b + 1));
    if (rest) return wi * 32 + b + 1 + (__ffs(rest) - 1);
    for (uint64_t j = wi + 1; j < n_words; j++) {
        uint32_t x = __ldg(start + j);
        if (x) {
            uint64_t e = j * 32 + (__ffs(x) - 1);
            return e < n ? e : n;
        }
    }
    return n;
}

// ---- host side ---------------------------------------------------------------------------------------------
struct KernelTimer;

struct bpe_ctx {
    int device = 0;
    int n_sm = BPE_N_SM;  // multiprocessors of this device (cooperative grids are sized by it)
    std::string err;
    uint32_t* d_status = nullptr;  // status word
    ClassTables cls{};
    uint8_t* d_stage1 = nullptr;
    uint32_t* d_stage2 = nullptr;
    Specials sp{};
    // distinct pretokens
    int table_log2 = 22;
    uint64_t pool_bytes = 256ull << 20;
    UniqueTable ut{};
    bool ut_alloc = false;
    // encode model
    bool model_loaded = false;
    uint64_t* d_mkey = nullptr;   // merge table: (a << 32 | b) + 1
    uint32_t* d_mrank = nullptr;
    uint32_t* d_mid = nullptr;
    uint64_t m_mask = 0;
    uint32_t* d_byte_to_id = nullptr;  // 256
    uint32_t* d_special_id = nullptr;  // BPE_MAX_SPECIALS
    uint32_t* d_idpool = nullptr;      // cached encodings of distinct pretokens
    uint64_t idpool_cap = 0;
    uint64_t enc_done = 0;             // number of distinct pretokens already encoded
    bool enc_warm = false;             // the last bpe_encode found every pretoken in the table: try without counting first
    uint32_t fused_attr_set = 0;  // bit per variant of the fused kernels
    bool count_attr_set = false, emit_attr_set = false;  // dynamic shared memory limits raised for this device's kernels
    // scratch reused across calls
    void* scratch[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    size_t scratch_bytes[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    volatile uint32_t* h_read = nullptr;  // api.cu ctx_read_words: mapped host words small results are stored to
    uint32_t* d_read = nullptr;
    void* train_kern = nullptr;  // train.cu: the merge loop variant whose launch attributes are set, and its CTAs per SM
    int train_per_sm = 0;
    uint32_t* h_beat = nullptr;  // train.cu: pinned host words the merge loop writes its progress to
    // sharded heavy steps (bpe_peer_*): this GPU's exchange buffer, the peers' buffers (opened through CUDA IPC), sizes
    void* peer_buf = nullptr;
    void* peer_ptr[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    int peer_rank = 0, peer_world = 1;
    uint32_t peer_cap = 0;
    bool train_sharded = false;
    void* io_staging[2] = {nullptr, nullptr};  // io.cu: pinned staging buffers + copy streams; [0] loads, [1] stores (they may overlap)
    // timing
    bool profiling = false;
    std::vector<bpe_kernel_time> times;
    double train_ms[3] = {0, 0, 0};
};

int ctx_fail(bpe_ctx* ctx, int code, const char* what, cudaError_t e = cudaSuccess);
void* ctx_scratch(bpe_ctx* ctx, int which, size_t bytes);  // grows-only device scratch, nullptr on failure
int ctx_status_to_code(bpe_ctx* ctx, uint32_t st);
uint32_t ctx_read_status(bpe_ctx* ctx, cudaStream_t s, bool clear = true);
uint32_t ctx_read_words(bpe_ctx* ctx, cudaStream_t s, const void* d_src, int n_words, void* h_dst);
int launch_find_byte(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, uint8_t byte, uint32_t status_bit, cudaStream_t s);

#define CUDA_TRY(ctx, expr)                                                      \
    do {                                                                         \
        cudaError_t e__ = (expr);                                                \
        if (e__ != cudaSuccess) return ctx_fail((ctx), BPE_ERR_CUDA, #expr, e__); \
    } while (0)

// RAII launch record: always counts the launch under `name`; with profiling on it also brackets the launch
// with CUDA events on the launching stream and accumulates the device time.
struct KernelTimer {
    bpe_ctx* ctx;
    const char* name;
    cudaStream_t s;
    cudaEvent_t a = nullptr, b = nullptr;
    KernelTimer(bpe_ctx* c, const char* nm, cudaStream_t st) : ctx(c), name(nm), s(st) {
        if (ctx->profiling) {
            cudaEventCreate(&a);
            cudaEventCreate(&b);
            cudaEventRecord(a, s);
        }
    }
    ~KernelTimer() {
        float ms = 0;
        if (a) {
            cudaEventRecord(b, s);
            cudaEventSynchronize(b);
            cudaEventElapsedTime(&ms, a, b);
            cudaEventDestroy(a);
            cudaEventDestroy(b);
        }
        for (auto& t : ctx->times)
            if (t.name == name) { t.ms += ms; t.launches++; return; }
        ctx->times.push_back({name, (double)ms, 1});
    }
};

void io_staging_free(bpe_ctx* ctx);
cudaError_t set_key_mask_count(unsigned long long mask);   // count.cu
cudaError_t set_key_mask_fused(unsigned long long mask);   // fused.cu
cudaError_t set_key_mask_encode(unsigned long long mask);  // encode.cu

// kernels launched from api.cu, defined in the other translation units
int launch_scan(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint32_t* d_cutmask, uint32_t flags, uint32_t* d_start, cudaStream_t s);
int launch_cuts_to_mask(bpe_ctx* ctx, const uint64_t* d_cuts, uint64_t n_cuts, uint64_t n, uint32_t* d_cutmask, cudaStream_t s);
int launch_count(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint32_t* d_start, cudaStream_t s);
int launch_fused_count(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint32_t* d_cutmask, uint32_t flags, uint32_t* d_start, cudaStream_t s);
int launch_import(bpe_ctx* ctx, const uint8_t* d_blob, const uint64_t* d_off, const uint32_t* d_len, const uint64_t* d_cnt,
                  uint64_t n_unique, cudaStream_t s);
int launch_export(bpe_ctx* ctx, uint8_t* d_blob, uint64_t* d_off, uint32_t* d_len, uint64_t* d_cnt, cudaStream_t s);
int run_train(bpe_ctx* ctx, uint32_t vocab_size, uint32_t first_new_id, const uint8_t* h_init_blob, const uint32_t* h_init_off,
              uint32_t* h_merge_a, uint32_t* h_merge_b, uint32_t* n_merges, cudaStream_t s);
int run_encode(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint32_t* d_start, void* d_out, int out_elem_bytes,
               uint64_t cap, uint64_t* n_out, cudaStream_t s, bool pre_counted);
int run_decode(bpe_ctx* ctx, const void* d_ids, int id_elem_bytes, uint64_t n_ids, const uint8_t* h_vocab_blob,
               const uint64_t* h_vocab_off, uint32_t n_vocab, uint8_t* d_bytes, uint64_t cap, uint64_t* n_bytes, cudaStream_t s);
