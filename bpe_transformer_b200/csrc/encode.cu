// encode.cu — BPETokenizer.encode / encode_iterable / decode (bpe_transformer/tokenization/bpe_tokenizer.py:119-290).
//
// The reference encodes pretoken by pretoken and memoises the result per distinct pretoken (:57-60, :232-236).
// Here the memo is the context's table of distinct pretokens (count.cu):
//   1. scan.cu marks token starts; count.cu inserts every pretoken of the text into the table;
//   2. encode_new_kernel: one WARP per pretoken that has no cached encoding yet: byte ids, then repeatedly
//      join the leftmost adjacent pair of minimum merge rank ((a, b) -> (rank, id) hash, L2-resident) — the
//      literal rule of _find_pair_in_merges (:257-290), one occurrence per round;
//   3. emit: per 32-byte chunk look every token up (key -> slot -> cached ids), count the ids, prefix-sum the
//      counts (per tile in shared memory, across tiles with a small scan kernel), write the ids in text order
//      as uint16 or uint32.
// Special tokens map straight to their id (:172-176).  No communication between GPUs: shards are independent.
#include "common.cuh"

#define ENC_THREADS 256

namespace {

struct MergeTable {
    const uint64_t* key;  // (a << 32 | b) + 1, 0 = empty
    const uint32_t* rank;
    const uint32_t* id;
    uint64_t mask;
};

__device__ __forceinline__ bool merge_lookup(const MergeTable& M, uint32_t a, uint32_t b, uint32_t* rank, uint32_t* id) {
    uint64_t key = (((uint64_t)a << 32) | b) + 1ull;
    uint64_t slot = (key * 0x9E3779B97F4A7C15ull >> 20) & M.mask;
    for (;;) {
        uint64_t cur = __ldg(M.key + slot);
        if (cur == key) { *rank = __ldg(M.rank + slot); *id = __ldg(M.id + slot); return true; }
        if (cur == 0) return false;
        slot = (slot + 1) & M.mask;
    }
}

// ---- step 2a: list the table slots that have no encoding yet ---------------------------------------------------
__global__ void list_new_kernel(UniqueTable ut, uint32_t* __restrict__ list, unsigned long long* __restrict__ n_list) {
    uint64_t slot = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (slot > ut.cap_mask || ut.key[slot] == 0 || ut.enc_len[slot] != 0xFFFFFFFFu) return;
    unsigned long long at = atomicAdd(n_list, 1ull);
    list[at] = (uint32_t)slot;
}

// ---- step 2b: one warp per new pretoken -------------------------------------------------------------------------
__global__ void __launch_bounds__(ENC_THREADS) encode_new_kernel(UniqueTable ut, const uint32_t* __restrict__ list, uint64_t n_list,
                                                                 MergeTable M, const uint32_t* __restrict__ byte_to_id,
                                                                 uint32_t* __restrict__ idpool, uint64_t idpool_cap, uint32_t* status) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
    if (warp >= n_list) return;
    const uint32_t slot = list[warp];
    const uint32_t len = ut.len[slot];
    const uint8_t* src = ut.pool + ut.pool_off[slot];
    uint64_t base = 0;
    if (lane == 0) base = atomicAdd(reinterpret_cast<unsigned long long*>(ut.counters + 3), (unsigned long long)len);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (base + len > idpool_cap) { if (lane == 0) set_status(status, ST_POOL_FULL); return; }
    uint32_t* ids = idpool + base;
    for (uint32_t i = lane; i < len; i += 32) {
        uint32_t id = __ldg(byte_to_id + src[i]);
        if (id == 0xFFFFFFFFu) set_status(status, ST_MISSING_ID);
        ids[i] = id;
    }
    __syncwarp();
    uint32_t cur = len;
    while (cur >= 2) {
        // minimum rank, leftmost position: compare (rank, position) packed in 64 bits
        unsigned long long best = ~0ull;
        uint32_t best_id = 0;
        for (uint32_t i = lane; i + 1 < cur; i += 32) {
            uint32_t r, id;
            if (merge_lookup(M, ids[i], ids[i + 1], &r, &id)) {
                unsigned long long k = ((unsigned long long)r << 32) | i;
                if (k < best) { best = k; best_id = id; }
            }
        }
        for (int o = 16; o; o >>= 1) {
            unsigned long long ob = __shfl_xor_sync(0xFFFFFFFFu, best, o);
            uint32_t oid = __shfl_xor_sync(0xFFFFFFFFu, best_id, o);
            if (ob < best) { best = ob; best_id = oid; }
        }
        if (best == ~0ull) break;
        if (best_id == 0xFFFFFFFFu) { if (lane == 0) set_status(status, ST_MISSING_ID); break; }  // KeyError at bpe_tokenizer.py:290
        uint32_t at = (uint32_t)best;
        // ids[at] = merged; close the gap at at+1 (chunks of 32, each read before it is overwritten)
        if (lane == 0) ids[at] = best_id;
        for (uint32_t c0 = at + 1; c0 + 1 < cur; c0 += 32) {
            uint32_t i = c0 + lane;
            uint32_t v = (i + 1 < cur) ? ids[i + 1] : 0;
            __syncwarp();
            if (i + 1 < cur) ids[i] = v;
            __syncwarp();
        }
        cur--;
        __syncwarp();
    }
    if (lane == 0) { ut.enc_off[slot] = (uint32_t)base; ut.enc_len[slot] = cur; }
}

struct EmitParams {
    const uint8_t* text;
    uint64_t n;
    const uint32_t* start;
    uint64_t n_words;
    UniqueTable ut;
    const uint32_t* idpool;
    const uint32_t* special_id;
    uint32_t* status;
};

// Visits the tokens that start in mask word wi; calls f(ids pointer or null, count, special id).
template <class F>
__device__ __forceinline__ void for_each_token(const EmitParams& P, const Specials& sp, uint64_t wi, uint32_t w, F f) {
    for (uint32_t m = w; m; m &= m - 1) {
        uint32_t b = __ffs(m) - 1;
        uint64_t p = wi * 32 + b;
        uint64_t e = token_end(P.start, P.n_words, wi, w, b, P.n);
        uint64_t len = e - p;
        int s = -1;
        if (sp.n) {
            uint8_t b0 = P.text[p];
            if ((sp.first_byte_mask[b0 >> 5] >> (b0 & 31)) & 1u)
                for (int k = 0; k < sp.n && s < 0; k++) {
                    if (sp.off[k + 1] - sp.off[k] != len) continue;
                    bool ok = true;
                    for (uint32_t q = 0; q < len && ok; q++) ok = P.text[p + q] == sp.blob[sp.off[k] + q];
                    if (ok) s = k;
                }
        }
        if (s >= 0) {
            uint32_t sid = __ldg(P.special_id + s);
            if (sid == 0xFFFFFFFFu) set_status(P.status, ST_MISSING_ID);  // KeyError at bpe_tokenizer.py:175
            f((const uint32_t*)nullptr, 1u, sid);
            continue;
        }
        uint64_t slot = unique_find(P.ut, hash_bytes(P.text, p, len));
        if (slot == ~0ull) { set_status(P.status, ST_TABLE_FULL); continue; }
        f(P.idpool + P.ut.enc_off[slot], P.ut.enc_len[slot], 0u);
    }
}

// ---- step 3a: ids per tile --------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ENC_THREADS) emit_count_kernel(const EmitParams P, const __grid_constant__ Specials sp,
                                                                 uint32_t* __restrict__ tile_total, uint8_t* __restrict__ chunk_cnt) {
    __shared__ uint32_t warp_sum[ENC_THREADS / 32];
    const uint64_t wi = blockIdx.x * (uint64_t)ENC_THREADS + threadIdx.x;
    uint32_t mine = 0;
    if (wi < P.n_words) {
        uint32_t w = __ldg(P.start + wi);
        for_each_token(P, sp, wi, w, [&](const uint32_t*, uint32_t cnt, uint32_t) { mine += cnt; });
        // a chunk of 32 bytes holds at most 32 tokens but one token can have thousands of ids: saturate, recount on emit
        chunk_cnt[wi] = mine < 255 ? (uint8_t)mine : 255;
    }
    uint32_t v = mine;
    for (int o = 16; o; o >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, o);
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int k = 0; k < ENC_THREADS / 32; k++) t += warp_sum[k];
        tile_total[blockIdx.x] = t;
    }
}

// ---- step 3b: exclusive scan of the tile totals (one CTA; n_tiles is small: text bytes / 8192) -------------------
__global__ void __launch_bounds__(1024) scan_totals_kernel(const uint32_t* __restrict__ in, uint64_t n, uint64_t* __restrict__ out) {
    __shared__ uint64_t warp_tot[32];
    __shared__ uint64_t carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (uint64_t base = 0; base < n; base += 1024) {
        uint64_t i = base + threadIdx.x;
        uint64_t v = i < n ? in[i] : 0, x = v;
        for (int o = 1; o < 32; o <<= 1) {
            uint64_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if ((threadIdx.x & 31) >= (unsigned)o) x += y;
        }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint64_t t = warp_tot[threadIdx.x], z = t;
            for (int o = 1; o < 32; o <<= 1) {
                uint64_t y = __shfl_up_sync(0xFFFFFFFFu, z, o);
                if (threadIdx.x >= (unsigned)o) z += y;
            }
            warp_tot[threadIdx.x] = z - t;  // exclusive
        }
        __syncthreads();
        uint64_t carry = carry_s;
        uint64_t excl = carry + warp_tot[threadIdx.x >> 5] + x - v;
        if (i < n) out[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) out[n] = carry_s;
}

// ---- step 3c: write the ids ---------------------------------------------------------------------------------------
template <typename OutT>
__global__ void __launch_bounds__(ENC_THREADS) emit_write_kernel(const EmitParams P, const __grid_constant__ Specials sp,
                                                                 const uint64_t* __restrict__ tile_off, const uint8_t* __restrict__ chunk_cnt,
                                                                 OutT* __restrict__ out) {
    __shared__ uint32_t warp_sum[ENC_THREADS / 32];
    const uint64_t wi = blockIdx.x * (uint64_t)ENC_THREADS + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t w = 0, mine = 0;
    if (wi < P.n_words) {
        w = __ldg(P.start + wi);
        mine = chunk_cnt[wi];
        if (mine == 255) {  // saturated: recount
            mine = 0;
            for_each_token(P, sp, wi, w, [&](const uint32_t*, uint32_t cnt, uint32_t) { mine += cnt; });
        }
    }
    uint32_t x = mine;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
        if (lane >= (unsigned)o) x += y;
    }
    if (lane == 31) warp_sum[warp] = x;
    __syncthreads();
    uint32_t before = 0;
    for (uint32_t k = 0; k < warp; k++) before += warp_sum[k];
    uint64_t at = tile_off[blockIdx.x] + before + x - mine;
    if (wi < P.n_words && mine)
        for_each_token(P, sp, wi, w, [&](const uint32_t* ids, uint32_t cnt, uint32_t special) {
            if (!ids) { out[at++] = (OutT)special; return; }
            for (uint32_t k = 0; k < cnt; k++) out[at + k] = (OutT)ids[k];
            at += cnt;
        });
}

// ---- decode: ids -> bytes (bpe_tokenizer.py:119-137; U+FFFD for unknown ids) -------------------------------------------
template <typename IdT>
__global__ void __launch_bounds__(ENC_THREADS) decode_len_kernel(const IdT* __restrict__ ids, uint64_t n_ids, const uint64_t* __restrict__ voff,
                                                                 uint32_t n_vocab, uint32_t* __restrict__ tile_total) {
    __shared__ uint32_t warp_sum[ENC_THREADS / 32];
    uint64_t i = blockIdx.x * (uint64_t)ENC_THREADS + threadIdx.x;
    uint32_t v = 0;
    if (i < n_ids) {
        uint64_t id = (uint64_t)ids[i];
        uint32_t L = id < n_vocab ? (uint32_t)(voff[id + 1] - voff[id]) : 0;
        v = L ? L : 3;
    }
    for (int o = 16; o; o >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, o);
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int k = 0; k < ENC_THREADS / 32; k++) t += warp_sum[k];
        tile_total[blockIdx.x] = t;
    }
}

template <typename IdT>
__global__ void __launch_bounds__(ENC_THREADS) decode_write_kernel(const IdT* __restrict__ ids, uint64_t n_ids, const uint8_t* __restrict__ vblob,
                                                                   const uint64_t* __restrict__ voff, uint32_t n_vocab,
                                                                   const uint64_t* __restrict__ tile_off, uint8_t* __restrict__ out) {
    __shared__ uint32_t warp_sum[ENC_THREADS / 32];
    uint64_t i = blockIdx.x * (uint64_t)ENC_THREADS + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t L = 0; uint64_t src = 0; bool known = false;
    if (i < n_ids) {
        uint64_t id = (uint64_t)ids[i];
        uint32_t l = id < n_vocab ? (uint32_t)(voff[id + 1] - voff[id]) : 0;
        known = l != 0; L = known ? l : 3; src = known ? voff[id] : 0;
    }
    uint32_t x = L;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
        if (lane >= (unsigned)o) x += y;
    }
    if (lane == 31) warp_sum[warp] = x;
    __syncthreads();
    uint32_t before = 0;
    for (uint32_t k = 0; k < warp; k++) before += warp_sum[k];
    uint64_t at = tile_off[blockIdx.x] + before + x - L;
    if (i < n_ids) {
        if (known) for (uint32_t k = 0; k < L; k++) out[at + k] = vblob[src + k];
        else { out[at] = 0xEF; out[at + 1] = 0xBF; out[at + 2] = 0xBD; }
    }
}

}  // namespace

int run_encode(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint32_t* d_start, void* d_out, int out_elem_bytes,
               uint64_t cap, uint64_t* n_out, cudaStream_t s) {
    *n_out = 0;
    if (n == 0) return BPE_OK;
    // 1. every pretoken of the text into the table (the memo)
    int rc = launch_count(ctx, d_text, n, d_start, s);
    if (rc) return rc;
    // 2. encode the pretokens that are new
    uint64_t slots = ctx->ut.cap_mask + 1;
    uint64_t h_counters[4];
    CUDA_TRY(ctx, cudaMemcpyAsync(h_counters, ctx->ut.counters, sizeof(h_counters), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    uint32_t st = ctx_read_status(ctx, s);
    if (st) return ctx_status_to_code(ctx, st);
    const uint64_t n_unique = h_counters[1];
    MergeTable M{ctx->d_mkey, ctx->d_mrank, ctx->d_mid, ctx->m_mask};
    if (n_unique > ctx->enc_done) {
        uint64_t n_new = n_unique - ctx->enc_done;
        uint32_t* list = (uint32_t*)ctx_scratch(ctx, 2, n_new * 4 + 16);
        unsigned long long* n_list = (unsigned long long*)ctx_scratch(ctx, 0, 16);
        if (!list || !n_list) return BPE_ERR_CUDA;
        CUDA_TRY(ctx, cudaMemsetAsync(n_list, 0, 8, s));
        {
            KernelTimer t(ctx, "list_new_kernel", s);
            list_new_kernel<<<(unsigned)((slots + 255) / 256), 256, 0, s>>>(ctx->ut, list, n_list);
            CUDA_TRY(ctx, cudaGetLastError());
        }
        {
            KernelTimer t(ctx, "encode_new_kernel", s);
            uint64_t blocks = (n_new * 32 + ENC_THREADS - 1) / ENC_THREADS;
            encode_new_kernel<<<(unsigned)blocks, ENC_THREADS, 0, s>>>(ctx->ut, list, n_new, M, ctx->d_byte_to_id, ctx->d_idpool,
                                                                       ctx->idpool_cap, ctx->d_status);
            CUDA_TRY(ctx, cudaGetLastError());
        }
        ctx->enc_done = n_unique;
    }
    // 3. emit
    EmitParams P;
    P.text = d_text; P.n = n; P.start = d_start; P.n_words = (n + 31) / 32; P.ut = ctx->ut; P.idpool = ctx->d_idpool;
    P.special_id = ctx->d_special_id; P.status = ctx->d_status;
    const uint64_t n_tiles = (P.n_words + ENC_THREADS - 1) / ENC_THREADS;
    uint8_t* buf = (uint8_t*)ctx_scratch(ctx, 3, n_tiles * 4 + (n_tiles + 1) * 8 + P.n_words + 64);
    if (!buf) return BPE_ERR_CUDA;
    uint64_t* tile_off = (uint64_t*)buf;
    uint32_t* tile_total = (uint32_t*)(buf + (n_tiles + 1) * 8);
    uint8_t* chunk_cnt = buf + (n_tiles + 1) * 8 + n_tiles * 4;
    {
        KernelTimer t(ctx, "emit_count_kernel", s);
        emit_count_kernel<<<(unsigned)n_tiles, ENC_THREADS, 0, s>>>(P, ctx->sp, tile_total, chunk_cnt);
        CUDA_TRY(ctx, cudaGetLastError());
    }
    {
        KernelTimer t(ctx, "scan_totals_kernel", s);
        scan_totals_kernel<<<1, 1024, 0, s>>>(tile_total, n_tiles, tile_off);
        CUDA_TRY(ctx, cudaGetLastError());
    }
    uint64_t total = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&total, tile_off + n_tiles, 8, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    st = ctx_read_status(ctx, s);
    if (st) return ctx_status_to_code(ctx, st);
    *n_out = total;
    if (total > cap) return ctx_fail(ctx, BPE_ERR_CAPACITY, "output buffer too small");
    {
        KernelTimer t(ctx, "emit_write_kernel", s);
        if (out_elem_bytes == 2) emit_write_kernel<uint16_t><<<(unsigned)n_tiles, ENC_THREADS, 0, s>>>(P, ctx->sp, tile_off, chunk_cnt, (uint16_t*)d_out);
        else emit_write_kernel<uint32_t><<<(unsigned)n_tiles, ENC_THREADS, 0, s>>>(P, ctx->sp, tile_off, chunk_cnt, (uint32_t*)d_out);
        CUDA_TRY(ctx, cudaGetLastError());
    }
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    return ctx_status_to_code(ctx, ctx_read_status(ctx, s));
}

int run_decode(bpe_ctx* ctx, const void* d_ids, int id_elem_bytes, uint64_t n_ids, const uint8_t* h_vocab_blob,
               const uint64_t* h_vocab_off, uint32_t n_vocab, uint8_t* d_bytes, uint64_t cap, uint64_t* n_bytes, cudaStream_t s) {
    *n_bytes = 0;
    if (n_ids == 0) return BPE_OK;
    const uint64_t blob_bytes = h_vocab_off[n_vocab];
    const uint64_t n_tiles = (n_ids + ENC_THREADS - 1) / ENC_THREADS;
    uint8_t* vbuf = (uint8_t*)ctx_scratch(ctx, 4, (n_vocab + 1) * 8ull + blob_bytes + 64);
    uint8_t* buf = (uint8_t*)ctx_scratch(ctx, 3, n_tiles * 4 + (n_tiles + 1) * 8 + 64);
    if (!vbuf || !buf) return BPE_ERR_CUDA;
    uint64_t* voff = (uint64_t*)vbuf;
    uint8_t* vblob = vbuf + (n_vocab + 1) * 8ull;
    CUDA_TRY(ctx, cudaMemcpyAsync(voff, h_vocab_off, (n_vocab + 1) * 8ull, cudaMemcpyHostToDevice, s));
    if (blob_bytes) CUDA_TRY(ctx, cudaMemcpyAsync(vblob, h_vocab_blob, blob_bytes, cudaMemcpyHostToDevice, s));
    uint64_t* tile_off = (uint64_t*)buf;
    uint32_t* tile_total = (uint32_t*)(buf + (n_tiles + 1) * 8);
    {
        KernelTimer t(ctx, "decode_len_kernel", s);
        if (id_elem_bytes == 2) decode_len_kernel<uint16_t><<<(unsigned)n_tiles, ENC_THREADS, 0, s>>>((const uint16_t*)d_ids, n_ids, voff, n_vocab, tile_total);
        else decode_len_kernel<uint32_t><<<(unsigned)n_tiles, ENC_THREADS, 0, s>>>((const uint32_t*)d_ids, n_ids, voff, n_vocab, tile_total);
        CUDA_TRY(ctx, cudaGetLastError());
    }
    scan_totals_kernel<<<1, 1024, 0, s>>>(tile_total, n_tiles, tile_off);
    CUDA_TRY(ctx, cudaGetLastError());
    uint64_t total = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&total, tile_off + n_tiles, 8, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    *n_bytes = total;
    if (!d_bytes) return BPE_OK;  // size query
    if (total > cap) return ctx_fail(ctx, BPE_ERR_CAPACITY, "output buffer too small");
    {
        KernelTimer t(ctx, "decode_write_kernel", s);
        if (id_elem_bytes == 2) decode_write_kernel<uint16_t><<<(unsigned)n_tiles, ENC_THREADS, 0, s>>>((const uint16_t*)d_ids, n_ids, vblob, voff, n_vocab, tile_off, d_bytes);
        else decode_write_kernel<uint32_t><<<(unsigned)n_tiles, ENC_THREADS, 0, s>>>((const uint32_t*)d_ids, n_ids, vblob, voff, n_vocab, tile_off, d_bytes);
        CUDA_TRY(ctx, cudaGetLastError());
    }
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    return BPE_OK;
}

// ================================================================================================================
// bpe_decode_noalloc — ids -> bytes for callers that must not allocate device memory (include/bpe_b200.h).
//
// Same kernels as run_decode() above.  What differs is where the working memory comes from: the vocabulary is
// the caller's own device buffer, and the per-tile byte counts and offsets (the only scratch the three kernels
// need) live in a fixed array that is part of the library image.  The library reserves it on every GPU that
// loads it, for the life of the process: 768 KiB per GPU, enough for BPE_NOALLOC_MAX_IDS ids per call.
// ================================================================================================================
#include <cstdio>
#include <mutex>

#define NOALLOC_MAX_TILES ((uint64_t)BPE_NOALLOC_MAX_IDS / ENC_THREADS)
#define NOALLOC_OFF_BYTES ((NOALLOC_MAX_TILES + 1) * 8)                       /* uint64 tile offsets, one extra for the total */
#define NOALLOC_SCRATCH_BYTES ((NOALLOC_OFF_BYTES + NOALLOC_MAX_TILES * 4 + 255) / 256 * 256)

// In host code this name is only a handle for cudaGetSymbolAddress(); it is not a device address.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

#define NOALLOC_MAX_DEVICES 64
static void* g_noalloc_ptr[NOALLOC_MAX_DEVICES];  // device address of g_noalloc_scratch, per GPU; null until set up
static std::mutex g_noalloc_mutex;

static void noalloc_message(char* h_err, size_t h_err_cap, const char* what, cudaError_t e = cudaSuccess) {
    if (!h_err || !h_err_cap) return;
    if (e != cudaSuccess) snprintf(h_err, h_err_cap, "%s: %s", what, cudaGetErrorString(e));
    else snprintf(h_err, h_err_cap, "%s", what);
}

// Once per GPU, with that GPU current: ask the runtime where this GPU's copy of the array lives.
static cudaError_t noalloc_setup(int device, void** out) {
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return cudaSuccess;
}

extern "C" int bpe_decode_noalloc(const void* d_ids, int id_elem_bytes, uint64_t n_ids, const uint8_t* d_vocab_blob,
                                  const uint64_t* d_vocab_off, uint32_t n_vocab, uint8_t* d_bytes, uint64_t cap,
                                  uint64_t* n_bytes, void* stream, char* h_err, size_t h_err_cap) {
    if (h_err && h_err_cap) h_err[0] = 0;
    if (!n_bytes || (id_elem_bytes != 2 && id_elem_bytes != 4) || (n_ids && (!d_ids || !d_vocab_off))) {
        noalloc_message(h_err, h_err_cap, "bpe_decode_noalloc: bad arguments (ids must be 2 or 4 bytes wide, pointers non-null)");
        return BPE_ERR_ARGUMENT;
    }
    *n_bytes = 0;
    if (n_ids > (uint64_t)BPE_NOALLOC_MAX_IDS) {
        if (h_err && h_err_cap)
            snprintf(h_err, h_err_cap, "bpe_decode_noalloc: %llu ids in one call, the built-in scratch holds at most %llu; split the call or use bpe_decode",
                     (unsigned long long)n_ids, (unsigned long long)BPE_NOALLOC_MAX_IDS);
        return BPE_ERR_CAPACITY;
    }
    if (n_ids == 0) return BPE_OK;
    cudaStream_t s = (cudaStream_t)stream;
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess || device < 0 || device >= NOALLOC_MAX_DEVICES) {
        noalloc_message(h_err, h_err_cap, "bpe_decode_noalloc: cudaGetDevice", e);
        return BPE_ERR_CUDA;
    }
    void* base = nullptr;
    if ((e = noalloc_setup(device, &base)) != cudaSuccess) {
        noalloc_message(h_err, h_err_cap, "bpe_decode_noalloc: cudaGetSymbolAddress", e);
        return BPE_ERR_CUDA;
    }
    // Every word the kernels read below is written earlier in this same call (tile_total by decode_len_kernel,
    // tile_off by scan_totals_kernel), so what the previous call left behind is never seen: nothing needs clearing.
    uint64_t* tile_off = reinterpret_cast<uint64_t*>(base);
    uint32_t* tile_total = reinterpret_cast<uint32_t*>(static_cast<unsigned char*>(base) + NOALLOC_OFF_BYTES);
    const uint64_t n_tiles = (n_ids + ENC_THREADS - 1) / ENC_THREADS;
    if (id_elem_bytes == 2) decode_len_kernel<uint16_t><<<(unsigned)n_tiles, ENC_THREADS, 0, s>>>((const uint16_t*)d_ids, n_ids, d_vocab_off, n_vocab, tile_total);
    else decode_len_kernel<uint32_t><<<(unsigned)n_tiles, ENC_THREADS, 0, s>>>((const uint32_t*)d_ids, n_ids, d_vocab_off, n_vocab, tile_total);
    scan_totals_kernel<<<1, 1024, 0, s>>>(tile_total, n_tiles, tile_off);
    if ((e = cudaGetLastError()) != cudaSuccess) { noalloc_message(h_err, h_err_cap, "bpe_decode_noalloc: kernel launch", e); return BPE_ERR_CUDA; }
    uint64_t total = 0;
    if ((e = cudaMemcpyAsync(&total, tile_off + n_tiles, 8, cudaMemcpyDeviceToHost, s)) != cudaSuccess ||
        (e = cudaStreamSynchronize(s)) != cudaSuccess) {
        noalloc_message(h_err, h_err_cap, "bpe_decode_noalloc: reading the size", e);
        return BPE_ERR_CUDA;
    }
    *n_bytes = total;
    if (!d_bytes) return BPE_OK;  // size query
    if (total > cap) {
        noalloc_message(h_err, h_err_cap, "bpe_decode_noalloc: output buffer too small (*n_bytes holds the size needed)");
        return BPE_ERR_CAPACITY;
    }
    if (id_elem_bytes == 2) decode_write_kernel<uint16_t><<<(unsigned)n_tiles, ENC_THREADS, 0, s>>>((const uint16_t*)d_ids, n_ids, d_vocab_blob, d_vocab_off, n_vocab, tile_off, d_bytes);
    else decode_write_kernel<uint32_t><<<(unsigned)n_tiles, ENC_THREADS, 0, s>>>((const uint32_t*)d_ids, n_ids, d_vocab_blob, d_vocab_off, n_vocab, tile_off, d_bytes);
    if ((e = cudaGetLastError()) != cudaSuccess || (e = cudaStreamSynchronize(s)) != cudaSuccess) {
        noalloc_message(h_err, h_err_cap, "bpe_decode_noalloc: writing the bytes", e);
        return BPE_ERR_CUDA;
    }
    return BPE_OK;
}
