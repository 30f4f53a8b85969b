// encode.cu — BPETokenizer.encode / encode_iterable / decode (bpe_transformer/tokenization/bpe_tokenizer.py:119-290).
//
// The reference encodes pretoken by pretoken and memoises the result per distinct pretoken (:57-60, :232-236).
// Here the memo is the context's table of distinct pretokens (count.cu):
//   1. scan.cu marks token starts; count.cu inserts every pretoken of the text into the table;
//   2. encode_new_kernel: one WARP per pretoken that has no cached encoding yet: byte ids, then repeatedly
//      join the leftmost adjacent pair of minimum merge rank ((a, b) -> (rank, id) hash, L2-resident) — the
//      literal rule of _find_pair_in_merges (:257-290), one occurrence per round;
//   3. emit_fused_kernel: one pass over the text; a warp per 1 KiB window looks every token up (key -> 16-byte table
//      entry holding up to three ids in place), the windows' id counts are prefix-summed with a chained scan
//      (decoupled look-back), and the ids are written in text order as uint16 or uint32.
// Once a text's pretokens are all cached (a second file of the same language, the next shard), step 1's count and step 2
// are skipped: the emit pass itself reports a pretoken it does not know and only then is the text counted.
// Special tokens map straight to their id (:172-176).  No communication between GPUs: shards are independent.
#include <algorithm>

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
__global__ void __launch_bounds__(ENC_THREADS) encode_new_kernel(UniqueTable ut, const uint32_t* __restrict__ list,
                                                                 const unsigned long long* __restrict__ n_list_dev, MergeTable M, const uint32_t* __restrict__ byte_to_id,
                                                                 uint32_t* __restrict__ idpool, uint64_t idpool_cap, uint32_t* status) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
    if (warp >= *n_list_dev) return;  // the grid is sized for the upper bound "distinct pretokens not seen by an earlier call"
    const uint32_t slot = list[warp];
    const uint32_t len = ut.len[slot];
    const uint8_t* src = ut.pool + ut.pool_off[slot];
    uint64_t base = 0;
    if (lane == 0) base = atomicAdd(reinterpret_cast<unsigned long long*>(ut.counters + 3), (unsigned long long)len);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (base + len > idpool_cap) { if (lane == 0) set_status(status, ST_POOL_FULL); return; }
    uint32_t* ids = idpool + base;
    // A pretoken that cannot be encoded (a byte or a merged pair without an id: KeyError at bpe_tokenizer.py:206 / :290)
    // gets a tombstone entry and nothing is raised here: the table may hold pretokens of earlier texts.  The emit pass
    // raises when the text being encoded really contains it, on every call.
    bool bad = false;
    for (uint32_t i = lane; i < len; i += 32) {
        uint32_t id = __ldg(byte_to_id + src[i]);
        if (id == 0xFFFFFFFFu) bad = true;
        ids[i] = id;
    }
    bad = __any_sync(0xFFFFFFFFu, bad);
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
        if (best_id == 0xFFFFFFFFu) { bad = true; break; }
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
    if (lane == 0 && bad) {
        // Tombstone: the key is in the table (probe chains stay whole, the slot is not listed again), the value says
        // "no ids": the emit pass raises ST_MISSING_ID whenever the text being encoded contains this pretoken.
        const uint64_t key = ut.key[slot];
        ut.enc_len[slot] = 0xFFFFFFFEu;
        ut.enc[slot] = make_uint4((uint32_t)key, (uint32_t)(key >> 32), 0u, 0xFFFFFFFFu);
    }
    if (lane == 0 && !bad) {
        ut.enc_off[slot] = (uint32_t)base;
        ut.enc_len[slot] = cur;
        const uint64_t key = ut.key[slot];
        uint32_t w0 = (uint32_t)base, w1 = cur;
        const uint32_t i0 = ids[0], i1 = cur > 1 ? ids[1] : 0u, i2 = cur > 2 ? ids[2] : 0u;
        if (cur <= 3 && (i0 | i1 | i2) <= 0xFFFFu) { w0 = i0 | (i1 << 16); w1 = 0x80000000u | (cur << 16) | i2; }
        else if (cur == 1) { w0 = i0; w1 = 0xC0010000u; }
        ut.enc[slot] = make_uint4((uint32_t)key, (uint32_t)(key >> 32), w0, w1);
    }
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

// ---- exclusive scan of per-tile totals (decode; one CTA, the number of tiles is small) --------------------------------
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

// ---- single-pass emit: look every token up once, place the ids with a chained scan over 1 KiB windows -----------------
// A warp takes a window of 32 mask words (tickets are handed out in increasing order), spreads the window's tokens over
// its lanes, looks each one up (key -> table slot -> cached ids), publishes the window's id count, obtains the count of
// all earlier windows by looking back at their published states (decoupled look-back), and writes its ids.
// HBM traffic: N bytes of text + N/8 of mask read, 2T (or 4T) bytes of ids written; the tables sit in L2.
#ifndef EMIT_THREADS
#define EMIT_THREADS 256
#endif
#define EMIT_CTAS_PER_SM (1536 / EMIT_THREADS)
#ifndef EMIT_SLEEP_NS
#define EMIT_SLEEP_NS 64  // pause between two looks at the state words of earlier tiles
#endif
#define EMIT_WARPS (EMIT_THREADS / 32)
#define EMIT_KEEP 288        // tokens of a window whose table entries are kept in shared memory between the two passes
#define ST_AGG (1ull << 62)  // state word: window's own count is available
#define ST_INC (2ull << 62)  // state word: count of this and all earlier windows is available

struct EmitWarp {
    uint2 ent[EMIT_KEEP];       // (w0, w1) of UniqueTable::enc per token; a window with more tokens looks the rest up twice
    unsigned short pos[1056];   // start positions inside the window
};

struct EmitLookup {
    const EmitParams& P;
    const Specials& sp;
    uint32_t fb;
    // (w0, w1) as in UniqueTable::enc; a miss gives zero ids and raises ST_ENC_MISS
    __device__ __forceinline__ uint2 operator()(uint64_t p, uint64_t len) const {
        if (sp.n) {
            const int s = special_index(sp, fb, P.text, p, len);
            if (s >= 0) {
                const uint32_t id = __ldg(P.special_id + s);
                if (id == 0xFFFFFFFFu) set_status(P.status, ST_MISSING_ID);  // KeyError at bpe_tokenizer.py:175
                return make_uint2(id, 0xC0010000u);
            }
        }
        if (len < 0xFFFFFFFFull) {
            const uint64_t key = hash_key_short(P.text, p, (uint32_t)len);
            uint64_t slot = (key * 0xD6E8FEB86659FD93ull >> 17) & P.ut.cap_mask;  // the slot sequence of unique_find
            for (uint64_t probe = 0; probe <= P.ut.cap_mask; probe++) {
                const uint4 e = __ldg(P.ut.enc + slot);
                if (e.x == (uint32_t)key && e.y == (uint32_t)(key >> 32)) {
                    if (e.w == 0xFFFFFFFFu) {  // tombstone of a pretoken that has no encoding (KeyError at bpe_tokenizer.py:206 / :290)
                        set_status(P.status, ST_MISSING_ID);
                        return make_uint2(0u, 0u);
                    }
                    return make_uint2(e.z, e.w);
                }
                if ((e.x | e.y) == 0) break;
                slot = (slot + 1) & P.ut.cap_mask;
            }
        }
        set_status(P.status, ST_ENC_MISS);  // not counted / not encoded yet: the host counts the text and runs the pass again
        return make_uint2(0u, 0u);
    }
};

__device__ __forceinline__ uint32_t ent_ids(uint2 e) { return (e.y >> 31) ? ((e.y >> 16) & 0x3FFFu) : e.y; }

template <typename OutT>
__global__ void __launch_bounds__(EMIT_THREADS, EMIT_CTAS_PER_SM) emit_fused_kernel(const EmitParams P, const __grid_constant__ Specials sp, uint64_t n_windows,
                                                                     unsigned long long* __restrict__ state, unsigned long long* __restrict__ ticket,
                                                                     OutT* __restrict__ out, uint64_t cap, unsigned long long* __restrict__ total_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    EmitWarp& W = reinterpret_cast<EmitWarp*>(smem_raw)[threadIdx.x >> 5];
    const uint32_t lane = threadIdx.x & 31;
    const EmitLookup lookup{P, sp, special_first_bytes(sp)};
    const uint32_t warp = threadIdx.x >> 5;
    __shared__ unsigned long long sh_tile, sh_before;
    __shared__ unsigned long long sh_ids[EMIT_WARPS];
    const uint64_t n_tiles = (n_windows + EMIT_WARPS - 1) / EMIT_WARPS;
    // A CTA takes a TILE of EMIT_WARPS consecutive windows per ticket, one window per warp; the chained scan below runs once
    // per tile, in warp 0, while the other warps wait at a barrier (which costs no issue slots: with a look-back per
    // window, a third of the kernel's instructions were polls of other windows' state words).
    for (;;) {
        if (threadIdx.x == 0) sh_tile = atomicAdd(ticket, 1ull);
        __syncthreads();  // also: everyone is done with the previous tile's sh_ids / sh_before
        const unsigned long long tile = sh_tile;
        if (tile >= n_tiles) return;
        const unsigned long long win = tile * EMIT_WARPS + warp;  // may lie behind the last window: an empty one
        const uint64_t w0 = win * 32, wi = w0 + lane, kib = w0 * 32;
        const uint32_t w = wi < P.n_words ? __ldg(P.start + wi) : 0u;
        const uint32_t mine = __popc(w);
        uint32_t incl = mine;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= (uint32_t)o) incl += y;
        }
        const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31);
        uint32_t at = incl - mine;
        for (uint32_t m = w; m; m &= m - 1) W.pos[at++] = (unsigned short)(lane * 32 + (__ffs(m) - 1));
        uint64_t tail_end = 0;
        if (total) {
            const uint32_t last_lane = 31 - __clz(__ballot_sync(0xFFFFFFFFu, w != 0));
            if (lane == last_lane) tail_end = token_end(P.start, P.n_words, wi, w, 31 - __clz(w), P.n);
            tail_end = __shfl_sync(0xFFFFFFFFu, tail_end, last_lane);
        }
        __syncwarp();
        // ---- look every token up -------------------------------------------------------------------------------------
        uint32_t my_ids = 0;
        for (uint32_t t0 = 0; t0 < total; t0 += 32) {
            __syncwarp();
            const uint32_t t = t0 + lane;
            if (t >= total) continue;
            const uint64_t p = kib + W.pos[t];
            const uint64_t e = (t + 1 < total) ? kib + W.pos[t + 1] : tail_end;
            const uint2 ent = lookup(p, e - p);
            if (t < EMIT_KEEP) W.ent[t] = ent;
            my_ids += ent_ids(ent);
        }
        __syncwarp();
        unsigned long long win_ids = my_ids;
        for (int o = 16; o; o >>= 1) win_ids += __shfl_xor_sync(0xFFFFFFFFu, win_ids, o);
        // ---- chained scan over the tiles -----------------------------------------------------------------------------------
        if (lane == 0) sh_ids[warp] = win_ids;
        __syncthreads();
        if (warp == 0) {
            unsigned long long tile_ids = lane < EMIT_WARPS ? sh_ids[lane] : 0ull;
            for (int o = 16; o; o >>= 1) tile_ids += __shfl_xor_sync(0xFFFFFFFFu, tile_ids, o);
            unsigned long long before = 0;
            if (tile == 0) {
                if (lane == 0) *reinterpret_cast<volatile unsigned long long*>(state) = ST_INC | tile_ids;
            } else {
                if (lane == 0) *reinterpret_cast<volatile unsigned long long*>(state + tile) = ST_AGG | tile_ids;
                long long j0 = (long long)tile - 1;  // lane L looks at tile j0 - L
                for (;;) {
                    const long long j = j0 - (long long)lane;
                    unsigned long long v = j >= 0 ? *reinterpret_cast<volatile unsigned long long*>(state + j) : ST_INC;  // in front of tile 0: prefix 0
                    const uint32_t inc = __ballot_sync(0xFFFFFFFFu, (v >> 62) == 2), none = __ballot_sync(0xFFFFFFFFu, (v >> 62) == 0);
                    const uint32_t first_inc = inc ? (uint32_t)__ffs(inc) - 1 : 32;  // nearest tile whose inclusive prefix is known
                    const uint32_t need = first_inc >= 31 ? 0xFFFFFFFFu : ((2u << first_inc) - 1u);
                    if (none & need) {  // a tile we depend on has not published yet: look again, a little later
                        __nanosleep(EMIT_SLEEP_NS);
                        continue;
                    }
                    unsigned long long part = (lane <= first_inc) ? (v & ~(3ull << 62)) : 0ull;
                    for (int o = 16; o; o >>= 1) part += __shfl_xor_sync(0xFFFFFFFFu, part, o);
                    before += part;
                    if (first_inc < 32) break;
                    j0 -= 32;
                }
                if (lane == 0) *reinterpret_cast<volatile unsigned long long*>(state + tile) = ST_INC | (before + tile_ids);
            }
            if (lane == 0) {
                sh_before = before;
                if (tile == n_tiles - 1) *total_out = before + tile_ids;
            }
        }
        __syncthreads();
        unsigned long long before = sh_before;
        for (uint32_t k = 0; k < warp; k++) before += sh_ids[k];
        // ---- write the ids -----------------------------------------------------------------------------------------------------
        unsigned long long running = before;
        for (uint32_t t0 = 0; t0 < total; t0 += 32) {
            const uint32_t t = t0 + lane;
            uint2 ent = make_uint2(0u, 0u);
            if (t < total) {
                if (t < EMIT_KEEP) ent = W.ent[t];
                else {
                    const uint64_t p = kib + W.pos[t];
                    ent = lookup(p, ((t + 1 < total) ? kib + W.pos[t + 1] : tail_end) - p);
                }
            }
            const uint32_t n_ids = ent_ids(ent);
            uint32_t x = n_ids;
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
                if (lane >= (uint32_t)o) x += y;
            }
            const unsigned long long dst = running + x - n_ids;
            if (n_ids) {
                if (!(ent.y >> 31)) {
                    const uint32_t* ids = P.idpool + ent.x;
                    for (uint32_t k = 0; k < n_ids; k++) if (dst + k < cap) out[dst + k] = (OutT)__ldg(ids + k);
                } else if (ent.y & 0x40000000u) {
                    if (dst < cap) out[dst] = (OutT)ent.x;
                } else {
                    if (dst < cap) out[dst] = (OutT)(ent.x & 0xFFFFu);
                    if (n_ids > 1 && dst + 1 < cap) out[dst + 1] = (OutT)(ent.x >> 16);
                    if (n_ids > 2 && dst + 2 < cap) out[dst + 2] = (OutT)(ent.y & 0xFFFFu);
                }
            }
            running += __shfl_sync(0xFFFFFFFFu, x, 31);
        }
        __syncwarp();
    }
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

static int encode_new_pretokens(bpe_ctx* ctx, cudaStream_t s) {
    uint64_t slots = ctx->ut.cap_mask + 1;
    uint64_t h_counters[4];
    uint32_t st = ctx_read_words(ctx, s, ctx->ut.counters, 8, h_counters);  // no copy engine: api.cu
    if (st) {
        if (st != 0x80000000u) cudaMemsetAsync(ctx->d_status, 0, 4, s);
        return ctx_status_to_code(ctx, st);
    }
    const uint64_t n_unique = h_counters[1];
    if (n_unique <= ctx->enc_done) return BPE_OK;
    MergeTable M{ctx->d_mkey, ctx->d_mrank, ctx->d_mid, ctx->m_mask};
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
        encode_new_kernel<<<(unsigned)blocks, ENC_THREADS, 0, s>>>(ctx->ut, list, n_list, M, ctx->d_byte_to_id, ctx->d_idpool, ctx->idpool_cap,
                                                                   ctx->d_status);
        CUDA_TRY(ctx, cudaGetLastError());
    }
    ctx->enc_done = n_unique;
    return BPE_OK;
}

int run_encode(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint32_t* d_start, void* d_out, int out_elem_bytes,
               uint64_t cap, uint64_t* n_out, cudaStream_t s, bool pre_counted) {
    *n_out = 0;
    if (n == 0) return BPE_OK;
    EmitParams P;
    P.text = d_text; P.n = n; P.start = d_start; P.n_words = (n + 31) / 32; P.ut = ctx->ut; P.idpool = ctx->d_idpool;
    P.special_id = ctx->d_special_id; P.status = ctx->d_status;
    const uint64_t n_windows = (P.n_words + 31) / 32;
    const uint64_t n_tiles = (n_windows + EMIT_WARPS - 1) / EMIT_WARPS;  // one state word of the chained scan per tile of windows
    unsigned long long* buf = (unsigned long long*)ctx_scratch(ctx, 3, (n_tiles + 4) * 8);
    if (!buf) return BPE_ERR_CUDA;
    unsigned long long* ticket = buf, *total_out = buf + 1, *state = buf + 2;
    if (!ctx->emit_attr_set) {  // per context = per device
        CUDA_TRY(ctx, cudaFuncSetAttribute(emit_fused_kernel<uint16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(EMIT_WARPS * sizeof(EmitWarp))));
        CUDA_TRY(ctx, cudaFuncSetAttribute(emit_fused_kernel<uint32_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(EMIT_WARPS * sizeof(EmitWarp))));
        ctx->emit_attr_set = true;
    }
    // With a warm table (every pretoken of earlier texts is cached with its ids) the text goes straight to the emit pass;
    // if it meets a pretoken it does not know, the text is counted, the new pretokens are encoded, and the emit runs again.
    // pre_counted: the caller already counted this text into the table (one pass with the split, api.cu).
    bool counted = false;
    for (int attempt = 0; attempt < 2; attempt++) {
        if (!ctx->enc_warm || attempt == 1) {
            int rc = pre_counted ? BPE_OK : launch_count(ctx, d_text, n, d_start, s);
            if (rc) return rc;
            rc = encode_new_pretokens(ctx, s);
            if (rc) return rc;
            counted = true;
        }
        CUDA_TRY(ctx, cudaMemsetAsync(buf, 0, (n_tiles + 4) * 8, s));
        {
            KernelTimer t(ctx, "emit_fused_kernel", s);
            const uint64_t blocks = std::min<uint64_t>(n_tiles, (uint64_t)BPE_N_SM * EMIT_CTAS_PER_SM);
            const size_t smem = EMIT_WARPS * sizeof(EmitWarp);
            if (out_elem_bytes == 2) emit_fused_kernel<uint16_t><<<(unsigned)blocks, EMIT_THREADS, smem, s>>>(P, ctx->sp, n_windows, state, ticket, (uint16_t*)d_out, cap, total_out);
            else emit_fused_kernel<uint32_t><<<(unsigned)blocks, EMIT_THREADS, smem, s>>>(P, ctx->sp, n_windows, state, ticket, (uint32_t*)d_out, cap, total_out);
            CUDA_TRY(ctx, cudaGetLastError());
        }
        unsigned long long total = 0;
        uint32_t st = ctx_read_words(ctx, s, total_out, 2, &total);  // total and status in one round trip, no copy engine
        if (st && st != 0x80000000u) cudaMemsetAsync(ctx->d_status, 0, 4, s);
        if ((st & ST_ENC_MISS) && !counted) { ctx->enc_warm = false; continue; }  // cold table: count, then emit again
        if (st & ~ST_ENC_MISS) { ctx->enc_warm = false; return ctx_status_to_code(ctx, st & ~ST_ENC_MISS); }
        if (st & ST_ENC_MISS) {  // counted, yet without an encoding: encode_new_kernel could not encode it
            ctx->enc_warm = false;
            return ctx_status_to_code(ctx, ST_MISSING_ID);
        }
        ctx->enc_warm = true;
        *n_out = total;
        if (total > cap) return ctx_fail(ctx, BPE_ERR_CAPACITY, "output buffer too small");
        return BPE_OK;
    }
    return ctx_fail(ctx, BPE_ERR_STATE, "encode did not converge");
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

cudaError_t set_key_mask_encode(unsigned long long mask) { return cudaMemcpyToSymbol(c_key_and, &mask, sizeof(mask)); }
