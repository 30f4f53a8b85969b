// train.cu — the greedy merge loop of BPETrainer (bpe_transformer/tokenization/bpe_trainer.py:169-426).
//
// State, all in HBM / L2 (a 2 GB TinyStories-shaped corpus has ~10^5 distinct pretokens, ~10^6 symbols):
//   words     distinct pretokens as arrays of token ids (sym[off .. off+len)), their counts
//   pairs     open-addressing table (a, b) -> signed 64-bit count, plus a list of occupied slots
//   tokens    byte strings, a lexicographic rank per token (ties in count are broken on bytes, :36-67)
//   postings  for every learned token X the words that were rewritten when X was created.  A pair that
//             contains X can only occur in those words, so a merge step visits wl[X] instead of all words
//             (the reference's pair -> pretokens inverted index, :186-199, :291-301, keyed by token;
//             entries go stale and are re-checked on visit like :260-262).
// One persistent cooperative kernel runs the whole loop; a step is: argmax over live pairs -> grid sync ->
// rewrite the posting list's words + count deltas + rank maintenance -> grid sync.  No host round trip.
//
// The count deltas reproduce the reference exactly, including its double counting:
// removed neighbours are read from the ORIGINAL word (:414-417), added neighbours are counted once per new
// token position on each side (:349-357).  Counts can therefore go negative; pairs with count > 0 but no
// occurrence are retired without producing a merge, as the reference's heap pop does (:251-262, :309).
#include <cooperative_groups.h>

#include <algorithm>

#include "common.cuh"

namespace cg = cooperative_groups;

#define TRAIN_THREADS 512
#define PAIR_PROBE_LIMIT 8192

namespace {

struct Best {
    long long cnt;  // 0 = none
    uint32_t ra, rb;
    unsigned long long key;
    uint32_t slot;
};

__device__ __forceinline__ bool better(const Best& x, const Best& y) {
    if (x.cnt != y.cnt) return x.cnt > y.cnt;
    if (x.ra != y.ra) return x.ra > y.ra;
    if (x.rb != y.rb) return x.rb > y.rb;
    return x.key > y.key;  // equal (count, bytes, bytes) on different id pairs: unpinned in the reference
}

struct Ctrl {
    unsigned int n_live;      // occupied pair slots (length of `live`)
    unsigned int wl_cursor;   // posting pool cursor
    unsigned int n_words;
    unsigned int sym_cursor;
    unsigned int rmax[2];     // 1 + highest rank among tokens that sort below the new token
    unsigned int req[2];      // 1 + rank of a token with the same bytes as the new token
    unsigned int n_merges;
    unsigned int n_iters;
    unsigned int blob_cursor;
    unsigned int pad;
};

struct TrainState {
    // words
    uint32_t* w_off; uint32_t* w_len; long long* w_cnt; uint32_t* sym;
    // pairs
    unsigned long long* ph_key; long long* ph_cnt; uint8_t* ph_dead; uint32_t* live; uint64_t ph_mask;
    // tokens
    uint32_t* tk_off; uint8_t* tk_blob; uint32_t blob_cap; uint32_t* rank; uint32_t* wl_off; uint32_t* wl_len;
    uint32_t* wl_pool; uint32_t wl_cap;
    uint32_t* m_a; uint32_t* m_b;
    Best* block_best;
    Ctrl* ctrl;
    uint32_t* status;
    uint32_t first_new_id, vocab_size;
};

__device__ __forceinline__ void pair_add(const TrainState& S, uint32_t a, uint32_t b, long long delta) {
    unsigned long long key = (((unsigned long long)a << 32) | b) + 1ull;
    uint64_t slot = (key * 0x9E3779B97F4A7C15ull >> 20) & S.ph_mask;
    for (int probe = 0; probe < PAIR_PROBE_LIMIT; probe++) {
        unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(S.ph_key + slot);
        if (cur == 0) {
            cur = atomicCAS(S.ph_key + slot, 0ull, key);
            if (cur == 0) {
                unsigned int idx = atomicAdd(&S.ctrl->n_live, 1u);
                S.live[idx] = (uint32_t)slot;  // n_live <= number of slots
                cur = key;
            }
        }
        if (cur == key) {
            atomicAdd(reinterpret_cast<unsigned long long*>(S.ph_cnt + slot), (unsigned long long)delta);
            return;
        }
        slot = (slot + 1) & S.ph_mask;
    }
    set_status(S.status, ST_PAIR_TABLE_FULL);
}

// ---- build the word table from the table of distinct pretokens (bpe_trainer.py:188-190: len 1 is skipped) ----
__global__ void words_build_kernel(UniqueTable ut, TrainState S) {
    uint64_t slot = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (slot > ut.cap_mask || ut.key[slot] == 0) return;
    uint32_t len = ut.len[slot];
    if (len < 2) return;
    unsigned int w = atomicAdd(&S.ctrl->n_words, 1u);
    unsigned int off = atomicAdd(&S.ctrl->sym_cursor, len);
    S.w_off[w] = off; S.w_len[w] = len; S.w_cnt[w] = (long long)ut.cnt[slot];
    const uint8_t* src = ut.pool + ut.pool_off[slot];
    for (uint32_t k = 0; k < len; k++) S.sym[off + k] = src[k];
}

// ---- pair histogram (bpe_trainer.py:188-199) ------------------------------------------------------------------
__global__ void pairs_init_kernel(TrainState S, uint32_t n_words) {
    uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    const uint32_t* s = S.sym + S.w_off[w];
    uint32_t len = S.w_len[w];
    long long c = S.w_cnt[w];
    uint32_t prev = s[0];
    for (uint32_t k = 1; k < len; k++) {
        uint32_t cur = s[k];
        pair_add(S, prev, cur, c);
        prev = cur;
    }
}

// Lexicographic comparison of token t's bytes with the concatenation bytes(a) + bytes(b).
__device__ int cmp_with_concat(const TrainState& S, uint32_t t, const uint8_t* A, uint32_t la, const uint8_t* B, uint32_t lb) {
    uint32_t o = __ldcg(S.tk_off + t), lt = __ldcg(S.tk_off + t + 1) - o;
    const uint8_t* T = S.tk_blob + o;
    uint32_t lx = la + lb, n = lt < lx ? lt : lx;
    for (uint32_t i = 0; i < n; i++) {
        uint32_t x = i < la ? A[i] : B[i - la];
        uint32_t y = T[i];
        if (y != x) return y < x ? -1 : 1;
    }
    return (lt > lx) - (lt < lx);
}

// One word: replace (a, b) by x left to right without overlap, apply the reference's count deltas.
__device__ void rewrite_word(const TrainState& S, uint32_t w, uint32_t a, uint32_t b, uint32_t x) {
    uint32_t* s = S.sym + S.w_off[w];
    uint32_t len = __ldcg(S.w_len + w);
    if (len < 2) return;
    long long c = S.w_cnt[w];
    uint32_t m = 0, j = 0, hits = 0;
    uint32_t cur = __ldcg(s), prev = 0;
    while (j < len) {
        uint32_t nxt = (j + 1 < len) ? __ldcg(s + j + 1) : 0xFFFFFFFFu;
        if (j + 1 < len && cur == a && nxt == b) {
            if (j >= 1) pair_add(S, prev, cur, -c);                       // (s[j-1], s[j]) from the original word
            uint32_t nn = (j + 2 < len) ? __ldcg(s + j + 2) : 0xFFFFFFFFu;
            if (j + 2 < len) pair_add(S, nxt, nn, -c);                    // (s[j+1], s[j+2]) from the original word
            s[m++] = x;
            hits++;
            prev = nxt; cur = nn; j += 2;
        } else {
            s[m++] = cur;
            prev = cur; cur = nxt; j += 1;
        }
    }
    if (!hits) return;
    for (uint32_t k = 0; k < m; k++) {
        if (s[k] != x) continue;
        if (k >= 1) pair_add(S, s[k - 1], x, c);
        if (k + 1 < m) pair_add(S, x, s[k + 1], c);
    }
    S.w_len[w] = m;
    unsigned int at = atomicAdd(&S.ctrl->wl_cursor, 1u);
    if (at < S.wl_cap) S.wl_pool[at] = w; else set_status(S.status, ST_LIST_FULL);
}

__global__ void __launch_bounds__(TRAIN_THREADS) merge_loop_kernel(TrainState S) {
    cg::grid_group grid = cg::this_grid();
    __shared__ Best sh_best[TRAIN_THREADS / 32];
    __shared__ Best sh_win;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t gtid = blockIdx.x * blockDim.x + tid, gsize = gridDim.x * blockDim.x;
    const bool leader = blockIdx.x == 0 && tid == 0;
    const uint32_t n_words = S.ctrl->n_words;
    const uint32_t NONE = 0xFFFFFFFFu;

    uint32_t next_id = S.first_new_id;
    // Results of the previous step.  They reach memory at the start of the next step (leader thread), so
    // phase 1 of that step reads them from these registers; from phase 2 on memory is current.
    uint32_t prev_slot = NONE, prev_x = NONE, prev_rank = 0, prev_matched = 0;
    uint32_t iter = 0;

    while (next_id < S.vocab_size) {
        const uint32_t par = iter & 1;
        // ---- phase 1: argmax over live pairs of (count, rank[a], rank[b]) --------------------------------------
        if (leader) {
            if (prev_slot != NONE) S.ph_dead[prev_slot] = 1;       // the winner never occurs again (bpe_trainer.py:243)
            if (prev_x != NONE) { S.rank[prev_x] = prev_rank; S.wl_len[prev_x] = prev_matched; }
            S.ctrl->rmax[par] = 0; S.ctrl->req[par] = 0;
        }
        const uint32_t n_live = __ldcg(&S.ctrl->n_live);
        const uint32_t wl_start = __ldcg(&S.ctrl->wl_cursor);
        Best best{0, 0, 0, 0, 0};
        for (uint32_t i = gtid; i < n_live; i += gsize) {
            uint32_t slot = __ldcg(S.live + i);
            long long cnt = __ldcg(S.ph_cnt + slot);
            if (cnt <= 0 || slot == prev_slot || __ldcg(S.ph_dead + slot)) continue;
            unsigned long long key = __ldcg(S.ph_key + slot);
            uint32_t a = (uint32_t)((key - 1) >> 32), b = (uint32_t)(key - 1);
            Best c{cnt, a == prev_x ? prev_rank : __ldcg(S.rank + a), b == prev_x ? prev_rank : __ldcg(S.rank + b), key, slot};
            if (better(c, best)) best = c;
        }
        for (int o = 16; o; o >>= 1) {
            Best other;
            other.cnt = __shfl_down_sync(0xFFFFFFFFu, best.cnt, o);
            other.ra = __shfl_down_sync(0xFFFFFFFFu, best.ra, o);
            other.rb = __shfl_down_sync(0xFFFFFFFFu, best.rb, o);
            other.key = __shfl_down_sync(0xFFFFFFFFu, best.key, o);
            other.slot = __shfl_down_sync(0xFFFFFFFFu, best.slot, o);
            if (better(other, best)) best = other;
        }
        if (lane == 0) sh_best[warp] = best;
        __syncthreads();
        if (tid == 0) {
            Best bb = sh_best[0];
            for (uint32_t k = 1; k < TRAIN_THREADS / 32; k++) if (better(sh_best[k], bb)) bb = sh_best[k];
            S.block_best[blockIdx.x] = bb;
        }
        grid.sync();
        if (warp == 0) {
            Best bb{0, 0, 0, 0, 0};
            for (uint32_t k = lane; k < gridDim.x; k += 32) {
                Best c;
                c.cnt = __ldcg(&S.block_best[k].cnt); c.ra = __ldcg(&S.block_best[k].ra); c.rb = __ldcg(&S.block_best[k].rb);
                c.key = __ldcg(&S.block_best[k].key); c.slot = __ldcg(&S.block_best[k].slot);
                if (better(c, bb)) bb = c;
            }
            for (int o = 16; o; o >>= 1) {
                Best other;
                other.cnt = __shfl_xor_sync(0xFFFFFFFFu, bb.cnt, o);
                other.ra = __shfl_xor_sync(0xFFFFFFFFu, bb.ra, o);
                other.rb = __shfl_xor_sync(0xFFFFFFFFu, bb.rb, o);
                other.key = __shfl_xor_sync(0xFFFFFFFFu, bb.key, o);
                other.slot = __shfl_xor_sync(0xFFFFFFFFu, bb.slot, o);
                if (better(other, bb)) bb = other;
            }
            if (lane == 0) sh_win = bb;
        }
        __syncthreads();
        const Best win = sh_win;
        if (win.cnt <= 0) break;  // nothing qualifies (bpe_trainer.py:242: heap exhausted)
        const uint32_t a = (uint32_t)((win.key - 1) >> 32), b = (uint32_t)(win.key - 1), x = next_id;

        // ---- phase 2: rewrite the words that can contain (a, b); keep the ranks ready for the new token ----------
        const uint32_t oa = __ldcg(S.tk_off + a), ob = __ldcg(S.tk_off + b);
        const uint32_t la_tok = __ldcg(S.tk_off + a + 1) - oa, lb_tok = __ldcg(S.tk_off + b + 1) - ob;
        const uint8_t* A = S.tk_blob + oa;
        const uint8_t* B = S.tk_blob + ob;
        uint32_t list_tok = NONE, list_len = n_words;  // byte pairs: every word
        if (a >= S.first_new_id) { list_tok = a; list_len = __ldcg(S.wl_len + a); }
        if (b >= S.first_new_id) {
            uint32_t lb = __ldcg(S.wl_len + b);
            if (list_tok == NONE || lb < list_len) { list_tok = b; list_len = lb; }
        }
        const uint32_t* list = (list_tok == NONE) ? nullptr : S.wl_pool + __ldcg(S.wl_off + list_tok);
        for (uint32_t i = gtid; i < list_len; i += gsize) rewrite_word(S, list ? __ldcg(list + i) : i, a, b, x);

        // Rank maintenance for x = bytes(a) + bytes(b): tokens that sort above it move up by one, its own rank
        // becomes (highest rank below it) + 1, or the rank of a token with identical bytes.  Done speculatively:
        // if the step merges nothing the order of the existing tokens is still intact.
        unsigned int my_rmax = 0;
        for (uint32_t t = gtid; t < x; t += gsize) {
            int c = cmp_with_concat(S, t, A, la_tok, B, lb_tok);
            uint32_t r = __ldcg(S.rank + t);
            if (c > 0) S.rank[t] = r + 1;
            else if (c < 0) my_rmax = max(my_rmax, r + 1);
            else S.ctrl->req[par] = r + 1;
        }
        my_rmax = __reduce_max_sync(0xFFFFFFFFu, my_rmax);
        if (lane == 0 && my_rmax) atomicMax(&S.ctrl->rmax[par], my_rmax);
        if (leader) {  // speculative too: overwritten by the next attempt if this step merges nothing
            uint32_t o = __ldcg(S.tk_off + x);
            if (o + la_tok + lb_tok <= S.blob_cap) {
                for (uint32_t i = 0; i < la_tok; i++) S.tk_blob[o + i] = A[i];
                for (uint32_t i = 0; i < lb_tok; i++) S.tk_blob[o + la_tok + i] = B[i];
            } else set_status(S.status, ST_LIST_FULL);
            S.tk_off[x + 1] = o + la_tok + lb_tok;
            S.m_a[x - S.first_new_id] = a; S.m_b[x - S.first_new_id] = b;
            S.wl_off[x] = wl_start;
        }
        grid.sync();
        const uint32_t matched = __ldcg(&S.ctrl->wl_cursor) - wl_start;
        const uint32_t req = __ldcg(&S.ctrl->req[par]), rmax = __ldcg(&S.ctrl->rmax[par]);
        prev_slot = win.slot;
        if (matched) {  // bpe_trainer.py:309-316
            prev_x = x; prev_rank = req ? req - 1 : rmax; prev_matched = matched;
            next_id++;
        } else {        // count > 0 but no occurrence: retired, no merge recorded
            prev_x = NONE;
        }
        iter++;
    }
    if (leader) { S.ctrl->n_iters = iter; S.ctrl->n_merges = next_id - S.first_new_id; }
}

size_t next_pow2(size_t v) { size_t p = 1; while (p < v) p <<= 1; return p; }

}  // namespace

int run_train(bpe_ctx* ctx, uint32_t vocab_size, uint32_t first_new_id, const uint8_t* h_init_blob, const uint32_t* h_init_off,
              uint32_t* h_merge_a, uint32_t* h_merge_b, uint32_t* n_merges, cudaStream_t s) {
    *n_merges = 0;
    if (!ctx->ut_alloc) return ctx_fail(ctx, BPE_ERR_NO_PAIRS, "no pretokens counted");
    uint64_t h_counters[4];
    CUDA_TRY(ctx, cudaMemcpyAsync(h_counters, ctx->ut.counters, sizeof(h_counters), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    const uint64_t n_unique = h_counters[1], n_sym = h_counters[0];
    if (n_unique == 0 || n_sym >= 0xFFFFFFF0ull) return ctx_fail(ctx, n_unique ? BPE_ERR_CAPACITY : BPE_ERR_NO_PAIRS, "word table size");
    if (vocab_size < first_new_id) vocab_size = first_new_id;  // nothing to learn, but "no pairs" is still an error

    cudaEvent_t ev[3];
    for (auto& e : ev) cudaEventCreate(&e);
    cudaEventRecord(ev[0], s);

    TrainState S{};
    S.first_new_id = first_new_id; S.vocab_size = vocab_size; S.status = ctx->d_status;
    const size_t ph_cap = next_pow2(std::max<size_t>(1u << 16, 2 * n_sym));
    S.ph_mask = ph_cap - 1;
    const uint32_t init_bytes = h_init_off[first_new_id];
    S.blob_cap = init_bytes + vocab_size * 256u;
    S.wl_cap = (uint32_t)n_sym + 16;
    // one allocation, carved up (all sizes rounded to 256 B)
    struct Part { void** p; size_t bytes; };
    Part parts[] = {
        {(void**)&S.w_off, n_unique * 4}, {(void**)&S.w_len, n_unique * 4}, {(void**)&S.w_cnt, n_unique * 8}, {(void**)&S.sym, (n_sym + 4) * 4},
        {(void**)&S.ph_key, ph_cap * 8}, {(void**)&S.ph_cnt, ph_cap * 8}, {(void**)&S.ph_dead, ph_cap}, {(void**)&S.live, ph_cap * 4},
        {(void**)&S.tk_off, (vocab_size + 2) * 4ull}, {(void**)&S.tk_blob, S.blob_cap}, {(void**)&S.rank, (vocab_size + 1) * 4ull},
        {(void**)&S.wl_off, (vocab_size + 1) * 4ull}, {(void**)&S.wl_len, (vocab_size + 1) * 4ull}, {(void**)&S.wl_pool, (size_t)S.wl_cap * 4},
        {(void**)&S.m_a, (vocab_size + 1) * 4ull}, {(void**)&S.m_b, (vocab_size + 1) * 4ull},
        {(void**)&S.block_best, sizeof(Best) * 4096}, {(void**)&S.ctrl, sizeof(Ctrl)},
    };
    size_t total = 0;
    for (auto& p : parts) total += (p.bytes + 255) / 256 * 256;
    uint8_t* base = (uint8_t*)ctx_scratch(ctx, 1, total);
    if (!base) return BPE_ERR_CUDA;
    size_t at = 0;
    for (auto& p : parts) { *p.p = base + at; at += (p.bytes + 255) / 256 * 256; }
    // zero the pair table, the control block and the posting lengths
    CUDA_TRY(ctx, cudaMemsetAsync(S.ph_key, 0, (uint8_t*)S.tk_off - (uint8_t*)S.ph_key, s));
    CUDA_TRY(ctx, cudaMemsetAsync(S.wl_off, 0, (uint8_t*)S.wl_pool - (uint8_t*)S.wl_off, s));
    CUDA_TRY(ctx, cudaMemsetAsync(S.ctrl, 0, sizeof(Ctrl), s));

    // initial tokens: bytes, offsets, lexicographic ranks (equal byte strings share a rank)
    std::vector<uint32_t> order(first_new_id), rank(first_new_id);
    for (uint32_t i = 0; i < first_new_id; i++) order[i] = i;
    auto tok = [&](uint32_t t) { return std::string((const char*)h_init_blob + h_init_off[t], h_init_off[t + 1] - h_init_off[t]); };
    std::sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) { return tok(x) < tok(y); });
    uint32_t r = 0;
    for (uint32_t i = 0; i < first_new_id; i++) {
        if (i && tok(order[i]) != tok(order[i - 1])) r++;
        rank[order[i]] = r;
    }
    CUDA_TRY(ctx, cudaMemcpyAsync(S.tk_blob, h_init_blob, init_bytes, cudaMemcpyHostToDevice, s));
    CUDA_TRY(ctx, cudaMemcpyAsync(S.tk_off, h_init_off, (first_new_id + 1) * 4ull, cudaMemcpyHostToDevice, s));
    CUDA_TRY(ctx, cudaMemcpyAsync(S.rank, rank.data(), first_new_id * 4ull, cudaMemcpyHostToDevice, s));

    {
        KernelTimer t(ctx, "words_build_kernel", s);
        uint64_t slots = ctx->ut.cap_mask + 1;
        words_build_kernel<<<(unsigned)((slots + 255) / 256), 256, 0, s>>>(ctx->ut, S);
        CUDA_TRY(ctx, cudaGetLastError());
    }
    Ctrl h_ctrl;
    CUDA_TRY(ctx, cudaMemcpyAsync(&h_ctrl, S.ctrl, sizeof(Ctrl), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    const uint32_t n_words = h_ctrl.n_words;
    if (n_words) {
        KernelTimer t(ctx, "pairs_init_kernel", s);
        pairs_init_kernel<<<(n_words + 255) / 256, 256, 0, s>>>(S, n_words);
        CUDA_TRY(ctx, cudaGetLastError());
    }
    CUDA_TRY(ctx, cudaMemcpyAsync(&h_ctrl, S.ctrl, sizeof(Ctrl), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    cudaEventRecord(ev[1], s);
    if (h_ctrl.n_live == 0) {
        for (auto& e : ev) cudaEventDestroy(e);
        return ctx_fail(ctx, BPE_ERR_NO_PAIRS, "no byte pairs to merge");
    }
    if (vocab_size == first_new_id) {
        for (auto& e : ev) cudaEventDestroy(e);
        return BPE_OK;
    }

    int per_sm = 0;
    CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, merge_loop_kernel, TRAIN_THREADS, 0));
    cudaDeviceProp prop;
    CUDA_TRY(ctx, cudaGetDeviceProperties(&prop, ctx->device));
    int grid = prop.multiProcessorCount * (per_sm > 2 ? 2 : per_sm);
    if (grid > 4096) grid = 4096;
    // small problems: fewer CTAs make the grid barrier cheaper
    uint64_t work = std::max<uint64_t>(h_ctrl.n_live, n_words);
    int want = (int)std::min<uint64_t>((work + TRAIN_THREADS - 1) / TRAIN_THREADS, (uint64_t)grid);
    if (want < 1) want = 1;
    grid = want;
    {
        KernelTimer t(ctx, "merge_loop_kernel", s);
        void* args[] = {&S};
        CUDA_TRY(ctx, cudaLaunchCooperativeKernel((void*)merge_loop_kernel, dim3(grid), dim3(TRAIN_THREADS), args, 0, s));
    }
    CUDA_TRY(ctx, cudaMemcpyAsync(&h_ctrl, S.ctrl, sizeof(Ctrl), cudaMemcpyDeviceToHost, s));
    cudaEventRecord(ev[2], s);
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    const uint32_t learned = h_ctrl.n_merges;
    CUDA_TRY(ctx, cudaMemcpyAsync(h_merge_a, S.m_a, learned * 4ull, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(ctx, cudaMemcpyAsync(h_merge_b, S.m_b, learned * 4ull, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    *n_merges = learned;
    float ms0 = 0, ms1 = 0;
    cudaEventElapsedTime(&ms0, ev[0], ev[1]);
    cudaEventElapsedTime(&ms1, ev[1], ev[2]);
    ctx->train_ms[0] = ms0; ctx->train_ms[1] = ms1; ctx->train_ms[2] = h_ctrl.n_iters;
    for (auto& e : ev) cudaEventDestroy(e);
    return ctx_status_to_code(ctx, ctx_read_status(ctx, s));
}
