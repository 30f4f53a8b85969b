// count.cu — the Counter of distinct pretokens (pretokenize_chunk + the fold over workers in the reference:
// bpe_transformer/tokenization/preprocessing/pretokenization.py:73-111, 171-208).
//
// Every thread owns one 32-bit word of the start mask (32 bytes of text).  For each set bit it finds the
// token's end (next set bit), forms a 64-bit key + 32-bit check over the bytes (8 bytes per round, unaligned
// 64-bit reads), and counts the key:
//   1. in a per-CTA shared-memory table (4096 keys; Zipf text: the CTA's hot words live there), one
//      shared-memory atomic per token;
//   2. keys that do not fit go straight to the context's global table (L2-resident for ~10^5 keys);
//   3. at the end of the CTA's range the shared table is folded into the global one, which also copies the
//      bytes of a first-seen pretoken into the context's byte pool, so the text buffer is not needed later.
// A CTA owns a CONTIGUOUS range of the text so that a representative position fits 32 bits.
// HBM traffic: N bytes of text + N/8 of mask read; the tables stay in shared memory / L2.
#include "common.cuh"

#define COUNT_THREADS 1024
#define COUNT_SLOTS 8192  // per-CTA shared table: 8 B key + 4 B count + 4 B position = 16 B/slot = 128 KB
#define COUNT_PROBES 8

namespace {

struct CountParams {
    const uint8_t* text;
    uint64_t n;
    const uint32_t* start;
    uint64_t n_words;          // words of the start mask that cover [0, n)
    uint64_t words_per_block;  // contiguous range per CTA
    uint32_t* status;
    UniqueTable ut;
};

struct SharedTable {
    unsigned long long key[COUNT_SLOTS];
    uint32_t cnt[COUNT_SLOTS];
    uint32_t pos[COUNT_SLOTS];  // byte position of one occurrence, relative to the CTA's first byte (length: from the mask)
};

#define COUNT_SMEM (sizeof(SharedTable) + (COUNT_THREADS / 32) * 1056 * sizeof(unsigned short))

// Adds c to the count of an already-known key; false if the key is not in the table yet.
__device__ __forceinline__ bool unique_bump(const UniqueTable& t, uint64_t key, uint64_t c) {
    uint64_t slot = (key * 0xD6E8FEB86659FD93ull >> 17) & t.cap_mask;
    for (int probe = 0; probe < 64; probe++) {
        uint64_t cur = *reinterpret_cast<volatile uint64_t*>(t.key + slot);
        if (cur == key) {
            atomicAdd(reinterpret_cast<unsigned long long*>(t.cnt + slot), (unsigned long long)c);
            return true;
        }
        if (cur == 0) return false;
        slot = (slot + 1) & t.cap_mask;
    }
    return false;
}

// A warp takes 32 mask words (1 KiB of text) per round.  The set bits of the 32 words are spread evenly over the
// lanes — lane L takes tokens L, L + 32, ... of the 1 KiB — through a per-warp list of start positions in shared
// memory, so that every lane hashes and counts one token at a time instead of looping over "its" 32 bytes (token
// counts per 32 bytes vary a lot: that loop ran with ~6 of 32 lanes active).
__global__ void __launch_bounds__(COUNT_THREADS) count_kernel(const CountParams P, const __grid_constant__ Specials sp) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SharedTable& st = *reinterpret_cast<SharedTable*>(smem_raw);
    unsigned short* pos_all = reinterpret_cast<unsigned short*>(smem_raw + sizeof(SharedTable));
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned short* pos = pos_all + warp * 1056;  // 1024 starts + 1 end marker, padded
    __shared__ unsigned int sh_used_word;
    unsigned int* sh_used = &sh_used_word;
    for (uint32_t k = threadIdx.x; k < COUNT_SLOTS; k += COUNT_THREADS) { st.key[k] = 0; st.cnt[k] = 0; }
    if (threadIdx.x == 0) sh_used_word = 0;
    const uint32_t fb = special_first_bytes(sp);
    __syncthreads();

    const uint64_t w_lo = (uint64_t)blockIdx.x * P.words_per_block;
    const uint64_t w_hi = (w_lo + P.words_per_block < P.n_words) ? w_lo + P.words_per_block : P.n_words;
    const uint64_t base = w_lo * 32;  // first byte of this CTA's range
    uint64_t local_tokens = 0;

    for (uint64_t w0 = w_lo + warp * 32; w0 < w_hi; w0 += COUNT_THREADS) {  // COUNT_THREADS = 32 words x warps per CTA
        // the call is lost once the table or a pool is full (the host enlarges them and counts again): do not grind on
        // (one decision per warp: lanes that saw different values would leave the shuffles below a lane short)
        if (__any_sync(0xFFFFFFFFu, (*reinterpret_cast<volatile uint32_t*>(P.status) & (ST_TABLE_FULL | ST_POOL_FULL)) != 0)) break;
        const uint64_t wi = w0 + lane;
        const uint32_t w = wi < w_hi ? __ldg(P.start + wi) : 0u;
        // exclusive scan of the per-lane token counts
        const uint32_t mine = __popc(w);
        uint32_t incl = mine;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= (uint32_t)o) incl += y;
        }
        const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31);
        if (total == 0) continue;
        uint32_t at = incl - mine;
        for (uint32_t m = w; m; m &= m - 1) pos[at++] = (unsigned short)(lane * 32 + (__ffs(m) - 1));
        // end of the last token: the next set bit anywhere behind this KiB
        uint32_t last_lane = 31 - __clz(__ballot_sync(0xFFFFFFFFu, w != 0));
        uint64_t tail_end = 0;
        if (lane == last_lane) tail_end = token_end(P.start, P.n_words, wi, w, 31 - __clz(w), P.n);
        tail_end = __shfl_sync(0xFFFFFFFFu, tail_end, last_lane);
        __syncwarp();
        const uint64_t kib = w0 * 32;
        for (uint32_t t0 = 0; t0 < total; t0 += 32) {  // warp-uniform trip count: the lanes re-converge every round
            __syncwarp();
            const uint32_t t = t0 + lane;
            if (t >= total) continue;
            const uint64_t p = kib + pos[t];
            const uint64_t e = (t + 1 < total) ? kib + pos[t + 1] : tail_end;
            const uint64_t len = e - p;
            if (sp.n && token_is_special(sp, fb, P.text, p, len)) continue;
            local_tokens++;
            bool done = false;
            uint64_t key = 0;
            if (len < 0xFFFFFFFFull && p - base < 0xFFFFFFFFull) {
                key = hash_key(P.text, p, (uint32_t)len);
                uint32_t slot = (uint32_t)(key >> 40) & (COUNT_SLOTS - 1);
#pragma unroll 1
                for (int probe = 0; probe < COUNT_PROBES && !done; probe++) {
                    unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(&st.key[slot]);
                    if (cur == 0) {
                        // an empty slot ends the search.  New keys are admitted only while the table is at most half full:
                        // a fuller table would make every lookup walk long probe chains (and the lanes of a warp diverge).
                        if (*reinterpret_cast<volatile unsigned int*>(sh_used) >= COUNT_SLOTS / 2) break;
                        cur = atomicCAS(&st.key[slot], 0ull, (unsigned long long)key);
                        if (cur == 0) {
                            st.pos[slot] = (uint32_t)(p - base);
                            atomicAdd(sh_used, 1u);
                            cur = key;
                        }
                    }
                    if (cur == key) {
                        atomicAdd(&st.cnt[slot], 1u);
                        done = true;
                    }
                    slot = (slot + 1) & (COUNT_SLOTS - 1);
                }
            }
            // not in the shared table (full neighbourhood): bump the global entry; only a first sighting takes the slow path
            if (!done && !(key && unique_bump(P.ut, key, 1))) unique_insert(P.ut, P.status, P.text, p, (uint32_t)len, hash_bytes(P.text, p, len), 1);
        }
        __syncwarp();
    }
    __syncthreads();
    // fold the CTA's table into the context's table
    for (uint32_t s = threadIdx.x; s < COUNT_SLOTS; s += COUNT_THREADS) {
        if (st.key[s] == 0) continue;
        if (*reinterpret_cast<volatile uint32_t*>(P.status) & (ST_TABLE_FULL | ST_POOL_FULL)) break;  // the call is lost anyway
        uint64_t p = base + st.pos[s];
        uint32_t len = (uint32_t)(token_end(P.start, P.n_words, p >> 5, __ldg(P.start + (p >> 5)), (uint32_t)(p & 31), P.n) - p);
        TokenKey k = hash_bytes(P.text, p, len);  // key again, plus the check word (not kept in shared memory)
        unique_insert(P.ut, P.status, P.text, p, len, k, st.cnt[s]);
    }
    // total occurrences (block-reduced)
    for (int o = 16; o; o >>= 1) local_tokens += __shfl_down_sync(0xFFFFFFFFu, local_tokens, o);
    if ((threadIdx.x & 31) == 0 && local_tokens) atomicAdd(reinterpret_cast<unsigned long long*>(P.ut.counters + 2), (unsigned long long)local_tokens);
}

// One thread per imported pretoken: pretoken i = blob[off[i], off[i] + len[i]).
__global__ void import_kernel(const uint8_t* __restrict__ blob, const uint64_t* __restrict__ off, const uint32_t* __restrict__ len,
                              const uint64_t* __restrict__ cnt, uint64_t n_unique, UniqueTable ut, uint32_t* status) {
    uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (i >= n_unique) return;
    TokenKey k = hash_bytes(blob, off[i], len[i]);
    unique_insert(ut, status, blob, off[i], len[i], k, cnt[i]);
    atomicAdd(reinterpret_cast<unsigned long long*>(ut.counters + 2), (unsigned long long)cnt[i]);
}

// One thread per table slot: every occupied slot takes the next index and the next byte range.
__global__ void export_kernel(UniqueTable ut, uint8_t* __restrict__ blob, uint64_t* __restrict__ off, uint32_t* __restrict__ len,
                              uint64_t* __restrict__ cnt, unsigned long long* cursors /* [0] index, [1] bytes */) {
    uint64_t slot = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (slot > ut.cap_mask || ut.key[slot] == 0) return;
    uint32_t L = ut.len[slot];
    uint64_t idx = atomicAdd(&cursors[0], 1ull);
    uint64_t at = atomicAdd(&cursors[1], (unsigned long long)L);
    const uint8_t* src = ut.pool + ut.pool_off[slot];
    for (uint32_t k = 0; k < L; k++) blob[at + k] = src[k];
    off[idx] = at;
    len[idx] = L;
    cnt[idx] = ut.cnt[slot];
}

}  // namespace

int launch_count(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint32_t* d_start, cudaStream_t s) {
    if (n == 0) return BPE_OK;
    CountParams P;
    P.text = d_text; P.n = n; P.start = d_start; P.status = ctx->d_status; P.ut = ctx->ut;
    P.n_words = (n + 31) / 32;
    uint64_t grid = (uint64_t)BPE_N_SM;  // one CTA of 32 warps per SM (the shared table and the start lists take 144 KB)
    // keep ranges a multiple of the CTA width so that all threads stay busy; small inputs use fewer CTAs
    uint64_t per = (P.n_words + grid - 1) / grid;
    per = (per + COUNT_THREADS - 1) / COUNT_THREADS * COUNT_THREADS;
    grid = (P.n_words + per - 1) / per;
    P.words_per_block = per;
    if (!ctx->count_attr_set) {  // per context = per device: function attributes do not carry over to another GPU
        CUDA_TRY(ctx, cudaFuncSetAttribute(count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)COUNT_SMEM));
        ctx->count_attr_set = true;
    }
    KernelTimer t(ctx, "count_kernel", s);
    count_kernel<<<(unsigned)grid, COUNT_THREADS, COUNT_SMEM, s>>>(P, ctx->sp);
    CUDA_TRY(ctx, cudaGetLastError());
    return BPE_OK;
}

int launch_import(bpe_ctx* ctx, const uint8_t* d_blob, const uint64_t* d_off, const uint32_t* d_len, const uint64_t* d_cnt,
                  uint64_t n_unique, cudaStream_t s) {
    if (!n_unique) return BPE_OK;
    import_kernel<<<(unsigned)((n_unique + 255) / 256), 256, 0, s>>>(d_blob, d_off, d_len, d_cnt, n_unique, ctx->ut, ctx->d_status);
    CUDA_TRY(ctx, cudaGetLastError());
    return BPE_OK;
}

int launch_export(bpe_ctx* ctx, uint8_t* d_blob, uint64_t* d_off, uint32_t* d_len, uint64_t* d_cnt, cudaStream_t s) {
    unsigned long long* cursors = (unsigned long long*)ctx_scratch(ctx, 0, 2 * sizeof(unsigned long long));
    if (!cursors) return BPE_ERR_CUDA;
    CUDA_TRY(ctx, cudaMemsetAsync(cursors, 0, 2 * sizeof(unsigned long long), s));
    uint64_t slots = ctx->ut.cap_mask + 1;
    export_kernel<<<(unsigned)((slots + 255) / 256), 256, 0, s>>>(ctx->ut, d_blob, d_off, d_len, d_cnt, cursors);
    CUDA_TRY(ctx, cudaGetLastError());
    return BPE_OK;
}

cudaError_t set_key_mask_count(unsigned long long mask) { return cudaMemcpyToSymbol(c_key_and, &mask, sizeof(mask)); }
