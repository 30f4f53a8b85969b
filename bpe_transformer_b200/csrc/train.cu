// train.cu — the greedy merge loop of BPETrainer (bpe_transformer/tokenization/bpe_trainer.py:169-426).
//
// State, all in HBM / L2 (a 2 GB TinyStories-shaped corpus has ~2*10^5 distinct pretokens, ~2*10^6 symbols):
//   words     distinct pretokens as arrays of token ids (sym[off .. off+len), 16-byte aligned), their counts
//   pairs     open-addressing table (a, b) -> signed 64-bit count, a list of occupied slots, and a HOT list: the
//             pairs whose count is at least a threshold T — the argmax looks only at those
//   tokens    byte strings plus an 8-byte big-endian prefix key (ties in count are broken on bytes, :36-67)
//   postings  for every learned token X the words that were rewritten when X was created.  A pair that contains
//             X can only occur in those words, so a merge step visits wl[X] instead of all words (the reference's
//             pair -> pretokens inverted index, :186-199, :291-301, keyed by token; entries go stale and are
//             re-checked on visit like :260-262).
//
// The loop is strictly sequential (~10^4 dependent steps) and each step touches little data, so it is LATENCY bound: what
// counts is the number of dependent round trips to L2 (~0.8 us each) per step.  One persistent kernel runs it, launched
// as a cooperative grid with one CTA per SM (BPE_B200_TRAIN_MODE=cluster: one 16-CTA cluster instead):
//   * CTA 0 drives.  Pairs with a count >= T live in its shared memory (HotCache), so the argmax never leaves the SM.
//     A "light" step (posting list of <= 2048 words: ~88 % of the steps) is done by CTA 0 alone with __syncthreads only:
//     argmax -> filter the posting list (four postings in flight per thread) -> rewrite the hits (one WARP per word).
//     Its count updates do not wait for global memory either: additions go to a per-step shared table (NewPairs),
//     everything that must reach the global table is queued (Deferred) and applied behind the next step's loads.
//   * A "heavy" step (a pair of bytes, whose candidates are ALL words, or a long posting list) is broadcast to the
//     other CTAs, which poll a flag in global memory in between (cluster mode: the hardware cluster barrier).
//     (Four candidates per thread in the heavy filter too was tried: -7 % on a corpus with 6*10^6 words, +4 % on the
//     2 GB TinyStories-shaped one through register pressure in this shared kernel; not kept.)
// No host round trip inside the loop; a pinned-host heartbeat and a device watchdog turn a bug into an error, not a hang.
//
// The count deltas reproduce the reference exactly, including its double counting:
// removed neighbours are read from the ORIGINAL word (:414-417), added neighbours are counted once per new
// token position on each side (:349-357).  Counts can therefore go negative; pairs with count > 0 but no
// occurrence are retired without producing a merge, as the reference's heap pop does (:251-262, :309).
#include <cooperative_groups.h>

#include <algorithm>
#include <chrono>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

#include "common.cuh"

namespace cg = cooperative_groups;

#define PAIR_PROBE_LIMIT 8192
#define RETIRED (-(1ll << 62))  // count of a pair that was merged (or found to have no occurrence): never positive again
#ifndef LIGHT_MAX
#define LIGHT_MAX 2048u         // posting lists up to this length are handled by the driver CTA alone (measured: 1024 / 1536 / 2048 / 3072 / 4096 -> 90.9 / 88.9 / 87.6 / 91.4 / 93.3 ms)
#endif
#ifndef QUEUE_CAP
#define QUEUE_CAP 2048u
#endif
#ifndef HEAVY_MIN_THREADS
#define HEAVY_MIN_THREADS 32  // heavy step: a CTA that gets candidates gets at least this many
#endif
#ifndef HELPER_SLEEP_NS
#define HELPER_SLEEP_NS 32   // helper CTAs look at the driver's flag this often while the driver works alone (measured: 16 / 32 / 64 / 256 / 1024 ns -> 86.9 / 86.8 / 87.3 / 87.6 / 88.4 ms)
#endif
#define NONE 0xFFFFFFFFu
#define MAX_TIES 1024u            // pairs that tie at the top count and are ordered by their bytes
#define W_MASK 0x1FFFFFFFu        // a posting's word index; the three bits above it name the GPU that filters the word in a sharded heavy step
#define W_OWNER_SHIFT 29
#define MAX_PEERS 8

namespace {

struct Best {
    long long cnt;  // 0 = none
    unsigned long long key;
    uint32_t slot;
};

struct Command {  // driver CTA -> helper CTAs, through global memory, ordered by the cluster / grid barrier
    uint32_t op;  // 0 = exit, 1 = rewrite (filter + rewrite; sharded: this GPU's share), 2 = sharded: rewrite the peers' hits
    uint32_t a, b, x, list_tok, list_len;
    long long threshold;
    uint32_t heavy_id;  // number of the heavy step (its parity picks the exchange buffers)
    uint32_t pad;
};

struct Ctrl {
    unsigned int n_live;      // occupied pair slots (length of `live`)
    unsigned int wl_cursor;   // posting pool cursor
    unsigned int n_words;
    unsigned int sym_cursor;
    unsigned int n_merges;
    unsigned int n_iters;
    unsigned int n_hot;       // length of `hot`
    unsigned int n_rebuilds;
    unsigned int n_heavy;
    unsigned int out_n;       // sharded heavy step: hits this GPU has sent to its peers
    unsigned int in_n[MAX_PEERS];  // ... and the hits each peer has sent here
    unsigned int go;          // grid variant: heavy steps announced by the driver CTA (helpers poll it)
    unsigned int done;        // grid variant: helper CTAs that finished their share, summed over all heavy steps
    unsigned int pad;
    Command cmd;
    unsigned int dbg[8];      // flags bit 1: [0] stale cache copies, [1] first iteration with one, [2] hot pairs missing from the cache, [3] first iteration
    unsigned long long phase_clk[8];  // developer trace: SM cycles of the driver's thread 0 per phase, summed over iterations
    unsigned long long sub_clk[8];    // developer trace: inside the argmax (scan, barrier, ties, barrier, order + barrier)
    unsigned long long heavy_clk[6];  // developer trace, heavy steps: drain, wake, own share, join, reload, rest
    unsigned long long t_begin, t_end; // globaltimer when the driver CTA entered / left the kernel
};

struct TrainState {
    // words
    uint32_t* w_off; uint32_t* w_len; long long* w_cnt; uint32_t* sym;
    // pairs
    unsigned long long* ph_key; long long* ph_cnt; uint32_t* live; ulonglong2* hot; uint64_t ph_mask;
    // tokens
    uint32_t* tk_off; uint8_t* tk_blob; uint32_t blob_cap; unsigned long long* tk_pre; uint32_t* wl_off; uint32_t* wl_len;
    uint2* wl_pool; uint32_t wl_cap;  // postings: (word, offset of its symbols): one load less on the way to the symbols
    uint32_t* m_a; uint32_t* m_b;
    Ctrl* ctrl;
    uint32_t* trace;  // optional (BPE_B200_TRAIN_TRACE): per iteration {ns, list length, 1 = all words, matched}
    uint32_t* status;
    uint32_t first_new_id, vocab_size;
    uint32_t flags;  // bit 0: never use the shared-memory cache (experiments)
    uint32_t watchdog_s;
    uint32_t cache_fill;  // entries a rebuild of the hot-pair cache aims for (the rest of its HC entries is room for pairs that rise past T)
    int driver_sm;  // -1: block 0 drives
    volatile uint32_t* beat;  // pinned host memory: [2 * cta] = iteration, [2 * cta + 1] = phase (what the host prints if the loop stalls)
    // ---- heavy steps shared between the GPUs of one box (csrc/train.cu header) ----
    uint32_t rank, world;              // world == 1: everything below is unused
    uint8_t* w_owner;                  // [n_words] the GPU that filters word w: a function of the pretoken's key, the same on every GPU
    unsigned long long* w_key;         // [n_words] the pretoken's key: how a hit is named on the wire (word indices differ between GPUs)
    uint32_t* slot_word;               // [table slots] word index of the pretoken in that slot of the context's table
    const unsigned long long* ut_key;  // the context's table: key per slot
    unsigned long long ut_mask;
    unsigned long long* inbox[MAX_PEERS];  // inbox[p]: GPU p's exchange buffer (peer memory for p != rank).  Layout: see peer_entry / peer_flag
    uint32_t inbox_cap;                // hits one sender may put into one receiver's buffer per heavy step
};

// Exchange buffer of one GPU: for each parity (heavy steps alternate) and each sender a flag word — (heavy step << 32) | hits —
// followed by the senders' hit areas of inbox_cap keys each.
__device__ __forceinline__ unsigned long long* peer_flag(const TrainState& S, uint32_t receiver, uint32_t parity, uint32_t sender) {
    return S.inbox[receiver] + parity * MAX_PEERS + sender;
}
__device__ __forceinline__ unsigned long long* peer_entry(const TrainState& S, uint32_t receiver, uint32_t parity, uint32_t sender) {
    return S.inbox[receiver] + 2 * MAX_PEERS + ((size_t)parity * MAX_PEERS + sender) * S.inbox_cap;
}

// Slot of pair `key` in the global table, inserted on first sight (compare-and-swap first: one round trip for a new pair).
__device__ __forceinline__ uint64_t pair_slot(const TrainState& S, unsigned long long key) {
    uint64_t slot = (key * 0x9E3779B97F4A7C15ull >> 20) & S.ph_mask;
    for (int probe = 0; probe < PAIR_PROBE_LIMIT; probe++) {
        unsigned long long cur = atomicCAS(S.ph_key + slot, 0ull, key);
        if (cur == 0) {
            unsigned int idx = atomicAdd(&S.ctrl->n_live, 1u);
            S.live[idx] = (uint32_t)slot;  // n_live <= number of slots
            return slot;
        }
        if (cur == key) return slot;
        slot = (slot + 1) & S.ph_mask;
    }
    set_status(S.status, ST_PAIR_TABLE_FULL);
    return 0;
}

// The same search when the first compare-and-swap has been issued earlier (its latency hidden behind other loads).
__device__ __forceinline__ uint64_t pair_slot_resume(const TrainState& S, unsigned long long key, uint64_t slot, unsigned long long cur) {
    for (int probe = 0; probe < PAIR_PROBE_LIMIT; probe++) {
        if (cur == 0) {
            unsigned int idx = atomicAdd(&S.ctrl->n_live, 1u);
            S.live[idx] = (uint32_t)slot;
            return slot;
        }
        if (cur == key) return slot;
        slot = (slot + 1) & S.ph_mask;
        cur = atomicCAS(S.ph_key + slot, 0ull, key);
    }
    set_status(S.status, ST_PAIR_TABLE_FULL);
    return 0;
}

// Count updates straight to the global table.  A pair whose count rises to or above the threshold is appended to the
// global "pending hot" list, which the driver drains (duplicates are harmless there).
struct GlobalAdder {
    const TrainState& S;
    long long threshold;
    __device__ __forceinline__ void operator()(uint32_t a, uint32_t b, long long delta) const {
        unsigned long long key = (((unsigned long long)a << 32) | b) + 1ull;
        uint64_t slot = pair_slot(S, key);
        if (delta < 0) {  // cannot enter the hot list: no need to wait for the old value
            atomicAdd(reinterpret_cast<unsigned long long*>(S.ph_cnt + slot), (unsigned long long)delta);
            return;
        }
        long long before = (long long)atomicAdd(reinterpret_cast<unsigned long long*>(S.ph_cnt + slot), (unsigned long long)delta);
        if (before < threshold && before + delta >= threshold) {
            unsigned int h = atomicAdd(&S.ctrl->n_hot, 1u);
            if (h <= S.ph_mask) S.hot[h] = make_ulonglong2(key, slot);
        }
    }
};

// ---- the driver CTA's shared-memory cache of the hot pairs -------------------------------------------------------------
// key / global slot / a COPY of the count for every pair with count >= T (plus leftovers), and a small hash index over
// the keys.  The global table stays exact (every update goes there too, fire-and-forget); the copy lets the argmax of a
// light step run without touching global memory.  Entries appended during a step get their count from global memory
// when the step closes; after a heavy step (other CTAs updated counts) all copies are reloaded.
#define HC 4096u      // cache entries
#define PRE_UNKNOWN (~0ull)  // HotCache::pre of an entry whose prefix key has not been fetched (yet): ties go to global memory
#define SUB_BINS 1024u
#define HIDX 8192u    // index slots (power of two)
struct HotCache {
    unsigned long long key[HC];  // 0 = hole (duplicate that lost the race)
    long long cnt[HC];
    uint32_t slot[HC];
    unsigned long long pre[HC];  // tk_pre of the pair's FIRST token: most ties in count are decided without leaving shared memory
    unsigned short index[HIDX];  // entry + 1, 0 = empty
    unsigned int n;              // entries in use (may exceed HC: overflow, forces a rebuild)
};

__device__ __forceinline__ uint32_t hc_hash(unsigned long long key) { return (uint32_t)((key * 0xD6E8FEB86659FD93ull) >> 50); }

__device__ __forceinline__ int hc_find(const HotCache& H, unsigned long long key) {
    uint32_t h = hc_hash(key) & (HIDX - 1);
    for (uint32_t probe = 0; probe < HIDX; probe++) {
        unsigned int v = *reinterpret_cast<const volatile unsigned short*>(&H.index[h]);
        if (v == 0) return -1;
        if (H.key[v - 1] == key) return (int)v - 1;
        h = (h + 1) & (HIDX - 1);
    }
    return -1;
}

// Appends (key, slot) with count `cnt`; a concurrent duplicate loses and becomes a hole.
__device__ __forceinline__ int hc_append(HotCache& H, unsigned long long key, uint32_t slot, long long cnt, unsigned long long pre = PRE_UNKNOWN) {
    unsigned int idx = atomicAdd(&H.n, 1u);
    if (idx >= HC) return -1;
    H.key[idx] = key; H.slot[idx] = slot; H.cnt[idx] = cnt; H.pre[idx] = pre;
    __threadfence_block();
    uint32_t h = hc_hash(key) & (HIDX - 1);
    for (uint32_t probe = 0; probe < HIDX; probe++) {
        unsigned short v = *reinterpret_cast<volatile unsigned short*>(&H.index[h]);
        if (v == 0) {
            v = atomicCAS(&H.index[h], (unsigned short)0, (unsigned short)(idx + 1));
            if (v == 0) return (int)idx;
        }
        if (H.key[v - 1] == key) { H.key[idx] = 0; H.cnt[idx] = RETIRED; return -1; }
        h = (h + 1) & (HIDX - 1);
    }
    return -1;
}

struct CacheAdder {
    const TrainState& S;
    HotCache& H;
    long long threshold;
    __device__ __forceinline__ void operator()(uint32_t a, uint32_t b, long long delta) const {
        unsigned long long key = (((unsigned long long)a << 32) | b) + 1ull;
        int e = hc_find(H, key);
        if (e >= 0) {  // hot already: update the copy, post the update to the global table without waiting
            atomicAdd(reinterpret_cast<unsigned long long*>(&H.cnt[e]), (unsigned long long)delta);
            atomicAdd(reinterpret_cast<unsigned long long*>(S.ph_cnt + H.slot[e]), (unsigned long long)delta);
            return;
        }
        uint64_t slot = pair_slot(S, key);
        if (delta < 0) {
            atomicAdd(reinterpret_cast<unsigned long long*>(S.ph_cnt + slot), (unsigned long long)delta);
            return;
        }
        long long before = (long long)atomicAdd(reinterpret_cast<unsigned long long*>(S.ph_cnt + slot), (unsigned long long)delta);
        if (before < threshold && before + delta >= threshold) hc_append(H, key, (uint32_t)slot, 0);  // count: loaded when the step closes
    }
};

// ---- light steps without waiting for global memory ------------------------------------------------------------------------
// Two facts about the reference's deltas (rewrite_word_warp below): every count that a step ADDS belongs to a pair that
// contains the step's new token x, i.e. to a pair that did not exist before the step; every pair that existed before
// only ever loses count.  So, in a light step,
//   * additions are summed in a small shared-memory table keyed by (other token, side).  When the step closes the table
//     holds the exact count of every new pair: those >= T enter the cache with that count (nothing to read back), and
//     all of them are handed to the deferred queue as one (key, +count) item each;
//   * subtractions from cached pairs update the cached copy at once and post a fire-and-forget update to the global
//     table; subtractions from uncached pairs (which can never become hot again) go to the deferred queue;
//   * the deferred queue is applied to the global table (slot lookup + add) at the start of the NEXT step, by threads
//     that are waiting for that step's posting lists anyway, and always before anything reads global counts (cache
//     rebuild, heavy step).  A new cached pair learns its global slot there, before any later step can subtract from it.
#define NPS 1024u     // new-pair slots per step (power of two); a fuller step falls back to CacheAdder for the rest
#define DQ_CAP 1024u  // deferred updates; a fuller queue applies the update on the spot
struct NewPairs {
    uint32_t key[NPS];  // ((other << 1) | side) + 1, side 1 = (x, other), 0 = (other, x); 0 = empty
    long long cnt[NPS];
    unsigned int overflow;  // some addition of this step went through CacheAdder
};
struct Deferred {
    unsigned long long key[DQ_CAP];
    long long delta[DQ_CAP];
    unsigned short entry[DQ_CAP];  // cache entry that waits for its global slot, 0xFFFF = none
    unsigned int n;
};

__device__ __forceinline__ void global_add(const TrainState& S, uint64_t slot, long long delta) {
    atomicAdd(reinterpret_cast<unsigned long long*>(S.ph_cnt + slot), (unsigned long long)delta);
}

// One deferred item: find (or claim) the pair's slot, post the update, tell the cache entry its slot.
__device__ __forceinline__ void deferred_apply(const TrainState& S, HotCache& H, unsigned long long key, long long delta, unsigned short entry) {
    const uint64_t slot = pair_slot(S, key);
    global_add(S, slot, delta);
    if (entry != 0xFFFFu) {
        H.slot[entry] = (uint32_t)slot;
        if (H.pre[entry] == PRE_UNKNOWN) H.pre[entry] = __ldcg(S.tk_pre + (uint32_t)((key - 1) >> 32));
    }
}

__device__ __forceinline__ void deferred_push(const TrainState& S, HotCache& H, Deferred& DQ, unsigned long long key, long long delta, unsigned short entry) {
    const unsigned int q = atomicAdd(&DQ.n, 1u);
    if (q < DQ_CAP) { DQ.key[q] = key; DQ.delta[q] = delta; DQ.entry[q] = entry; }
    else deferred_apply(S, H, key, delta, entry);
}

// Applies everything that is queued (all threads of the CTA call this; ends with a barrier, queue empty, updates landed).
template <int NT>
__device__ __forceinline__ void deferred_drain(const TrainState& S, HotCache& H, Deferred& DQ) {
    __syncthreads();
    const unsigned int n = min(DQ.n, DQ_CAP);
    for (unsigned int i = threadIdx.x; i < n; i += NT) deferred_apply(S, H, DQ.key[i], DQ.delta[i], DQ.entry[i]);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) DQ.n = 0;
    __syncthreads();
}

struct LightAdder {
    const TrainState& S;
    HotCache& H;
    NewPairs& NP;
    Deferred& DQ;
    long long threshold;
    uint32_t x;
    __device__ __forceinline__ void operator()(uint32_t a, uint32_t b, long long delta) const {
        if (delta > 0) {  // (a, x) or (x, b) or (x, x): a pair born in this step
            const uint32_t code = (a == x ? ((b << 1) | 1u) : (a << 1)) + 1u;
            uint32_t h = (code * 0x9E3779B1u) >> 22;  // 10 bits
            for (uint32_t probe = 0; probe < NPS; probe++) {
                uint32_t cur = *reinterpret_cast<volatile uint32_t*>(&NP.key[h]);
                if (cur == 0) cur = atomicCAS(&NP.key[h], 0u, code);
                if (cur == 0 || cur == code) {
                    atomicAdd(reinterpret_cast<unsigned long long*>(&NP.cnt[h]), (unsigned long long)delta);
                    return;
                }
                h = (h + 1) & (NPS - 1);
            }
            NP.overflow = 1;
            CacheAdder{S, H, threshold}(a, b, delta);
            return;
        }
        const unsigned long long key = (((unsigned long long)a << 32) | b) + 1ull;
        const int e = hc_find(H, key);
        if (e >= 0) {
            atomicAdd(reinterpret_cast<unsigned long long*>(&H.cnt[e]), (unsigned long long)delta);
            global_add(S, H.slot[e], delta);
            return;
        }
        deferred_push(S, H, DQ, key, delta, 0xFFFFu);
    }
};

struct DriverShared {  // the driver CTA's dynamic shared memory
    HotCache H;
    NewPairs NP;
    Deferred DQ;
    uint32_t tie[MAX_TIES];      // cache entries (or global slots) that share the top count
    unsigned int sub[SUB_BINS];  // rebuild: histogram inside the threshold's power-of-two bin
};

// ---- build the word table from the table of distinct pretokens (bpe_trainer.py:188-190: len 1 is skipped) ----
__global__ void words_build_kernel(UniqueTable ut, TrainState S) {
    uint64_t slot = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (slot > ut.cap_mask || ut.key[slot] == 0) return;
    uint32_t len = ut.len[slot];
    if (len < 2) return;
    unsigned int w = atomicAdd(&S.ctrl->n_words, 1u);
    unsigned int off = atomicAdd(&S.ctrl->sym_cursor, (len + 3u) & ~3u);  // 16-byte aligned: scanned with uint4 loads
    S.w_off[w] = off; S.w_len[w] = len; S.w_cnt[w] = (long long)ut.cnt[slot];
    if (S.world > 1) {
        const unsigned long long key = ut.key[slot];
        S.w_key[w] = key;
        S.w_owner[w] = (uint8_t)((key >> 40) % S.world);
        S.slot_word[slot] = w;
    }
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
        GlobalAdder{S, LLONG_MAX}(prev, cur, c);  // hot pairs are collected by the merge loop's first rebuild
        prev = cur;
    }
}

// ---- ordering of candidates: count, then (bytes(a), bytes(b)) as a tuple, then the id pair --------------------------
// 8-byte big-endian zero-padded prefixes order two byte strings whenever they differ; equal prefixes fall back to the
// bytes (a shorter string that is a prefix of a longer one pads with 0x00 <= any byte, so "prefixes differ" is decisive).
__device__ int token_cmp(const TrainState& S, uint32_t t, uint32_t u) {
    if (S.flags & 4u) atomicAdd(&S.ctrl->dbg[6], 1u);  // experiments: how often ties in count reach the byte comparison
    if (t == u) return 0;
    unsigned long long pt = __ldcg(S.tk_pre + t), pu = __ldcg(S.tk_pre + u);
    if (pt != pu) return pt < pu ? -1 : 1;
    uint32_t ot = __ldcg(S.tk_off + t), lt = __ldcg(S.tk_off + t + 1) - ot, ou = __ldcg(S.tk_off + u), lu = __ldcg(S.tk_off + u + 1) - ou;
    uint32_t n = lt < lu ? lt : lu;
    for (uint32_t i = 0; i < n; i++) {
        uint32_t x = __ldcg(S.tk_blob + ot + i), y = __ldcg(S.tk_blob + ou + i);
        if (x != y) return x < y ? -1 : 1;
    }
    return (lt > lu) - (lt < lu);
}

__device__ __forceinline__ bool better(const TrainState& S, const Best& x, const Best& y) {
    if (x.cnt != y.cnt) return x.cnt > y.cnt;
    if (x.cnt <= 0 || x.key == y.key) return false;
    int c = token_cmp(S, (uint32_t)((x.key - 1) >> 32), (uint32_t)((y.key - 1) >> 32));  // bpe_trainer.py:59-64
    if (!c) c = token_cmp(S, (uint32_t)(x.key - 1), (uint32_t)(y.key - 1));
    if (c) return c > 0;
    return x.key > y.key;  // equal (count, bytes, bytes) on different id pairs: unpinned in the reference
}

__device__ __forceinline__ Best shfl_best(const Best& b, int o, bool down) {
    Best r;
    r.cnt = down ? __shfl_down_sync(0xFFFFFFFFu, b.cnt, o) : __shfl_xor_sync(0xFFFFFFFFu, b.cnt, o);
    r.key = down ? __shfl_down_sync(0xFFFFFFFFu, b.key, o) : __shfl_xor_sync(0xFFFFFFFFu, b.key, o);
    r.slot = down ? __shfl_down_sync(0xFFFFFFFFu, b.slot, o) : __shfl_xor_sync(0xFFFFFFFFu, b.slot, o);
    return r;
}

// ---- word rewriting -----------------------------------------------------------------------------------------------------
// Does the word contain (a, b)?  One thread, 16-byte loads (independent of one another, so they overlap).
__device__ __forceinline__ bool word_has_pair_loaded(const TrainState& S, uint32_t off, uint4 v, uint32_t len, uint32_t a, uint32_t b) {
    if (len < 2) return false;
    const uint4* s4 = reinterpret_cast<const uint4*>(S.sym + off);
    bool hit = (v.x == a && v.y == b) | (2 < len && v.y == a && v.z == b) | (3 < len && v.z == a && v.w == b);
    uint32_t prev = v.w;
    for (uint32_t k = 4; k < len; k += 4) {
        v = __ldcg(s4 + (k >> 2));
        hit |= (prev == a && v.x == b) | (k + 1 < len && v.x == a && v.y == b) | (k + 2 < len && v.y == a && v.z == b) |
               (k + 3 < len && v.z == a && v.w == b);
        prev = v.w;
    }
    return hit;
}

// One word, one thread: replace (a, b) by x left to right without overlap, apply the reference's count deltas.
// Returns whether anything matched.  (Words longer than a warp, and queue overflow.)
template <class Adder>
__device__ bool rewrite_word_serial(const TrainState& S, uint32_t w, uint32_t off, uint32_t len, uint32_t a, uint32_t b, uint32_t x, const Adder& add) {
    w &= W_MASK;
    uint32_t* s = S.sym + off;
    if (len < 2) return false;
    long long c = S.w_cnt[w];
    uint32_t m = 0, j = 0, hits = 0;
    uint32_t cur = __ldcg(s), prev = 0;
    while (j < len) {
        uint32_t nxt = (j + 1 < len) ? __ldcg(s + j + 1) : NONE;
        if (j + 1 < len && cur == a && nxt == b) {
            if (j >= 1) add(prev, cur, -c);                               // (s[j-1], s[j]) from the original word
            uint32_t nn = (j + 2 < len) ? __ldcg(s + j + 2) : NONE;
            if (j + 2 < len) add(nxt, nn, -c);                            // (s[j+1], s[j+2]) from the original word
            s[m++] = x;
            hits++;
            prev = nxt; cur = nn; j += 2;
        } else {
            s[m++] = cur;
            prev = cur; cur = nxt; j += 1;
        }
    }
    if (!hits) return false;
    for (uint32_t k = 0; k < m; k++) {
        if (s[k] != x) continue;
        if (k >= 1) add(s[k - 1], x, c);
        if (k + 1 < m) add(x, s[k + 1], c);
    }
    S.w_len[w] = m;
    return true;
}

// One word per WARP (up to 32 symbols): every lane holds one symbol, the matches come from a ballot, the four count
// deltas of a match are issued by two lanes at once, the new word is written back compacted.  Same deltas as the
// serial version: for a match at j, removed (s[j-1], s[j]) and (s[j+1], s[j+2]) of the original word; added
// (left neighbour in the new word, x) and (x, right neighbour in the new word), a neighbour being x itself when the
// adjacent original positions were a match too.  Returns (warp-uniform) whether anything matched.
template <class Adder>
__device__ bool rewrite_word_warp(const TrainState& S, uint32_t w, uint32_t off, uint32_t len, uint32_t a, uint32_t b, uint32_t x, const Adder& add) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t w_tagged = w;
    w &= W_MASK;
    if (len == 0xFFFFu) len = __ldcg(S.w_len + w);  // the queue keeps 16 bits
    if (len < 2) return false;
    if (len > 32) {
        bool hit = lane == 0 ? rewrite_word_serial(S, w_tagged, off, len, a, b, x, add) : false;
        return __shfl_sync(0xFFFFFFFFu, hit, 0);
    }
    uint32_t* s = S.sym + off;
    const uint32_t me = lane < len ? __ldcg(s + lane) : NONE;
    const long long c = S.w_cnt[w];
    const uint32_t next1 = __shfl_down_sync(0xFFFFFFFFu, me, 1), prev1 = __shfl_up_sync(0xFFFFFFFFu, me, 1);
    const bool cand = lane + 1 < len && me == a && next1 == b;
    const uint32_t C = __ballot_sync(0xFFFFFFFFu, cand);
    if (!C) return false;
    bool match = cand;
    if (a == b) {  // a run of candidates: left to right without overlap = every other one from the run's start
        uint32_t starts = C & ~(C << 1);
        uint32_t my_start = 31 - __clz(starts & (0xFFFFFFFFu >> (31 - lane)));
        match = cand && !((lane - my_start) & 1u);
    }
    const uint32_t M = __ballot_sync(0xFFFFFFFFu, match);
    const bool second = lane >= 1 && ((M >> (lane - 1)) & 1u);  // this lane holds the right half of a match
    if (match && lane >= 1) {
        add(prev1, me, -c);
        add((lane >= 2 && ((M >> (lane - 2)) & 1u)) ? x : prev1, x, c);
    }
    if (second && lane + 1 < len) {
        add(me, next1, -c);
        add(x, ((M >> (lane + 1)) & 1u) ? x : next1, c);
    }
    __syncwarp();
    if (lane < len && !second) s[lane - __popc(M & ((1u << lane) - 1u))] = match ? x : me;
    if (lane == 0) S.w_len[w] = len - __popc(M);
    return true;
}

// The candidates of a heavy step, spread over all CTAs: every thread filters one candidate word per round, the CTA's
// warps rewrite the hits.  Rewritten words are appended to the posting pool through the global cursor.
// Hits of the filter stage, handed from threads to warps through shared memory.
struct Queue {
    uint32_t w[QUEUE_CAP], off[QUEUE_CAP];
    unsigned short len[QUEUE_CAP];  // 0xFFFF = long word: the rewriter reads the length again
    unsigned int n;
};

// Sharded heavy step: one more hit of this GPU's share goes to every peer, named by the pretoken's key.
__device__ __forceinline__ void send_hit(const TrainState& S, uint32_t parity, uint32_t w) {
    const unsigned long long key = S.w_key[w & W_MASK];
    const unsigned int at = atomicAdd(&S.ctrl->out_n, 1u);
    if (at >= S.inbox_cap) { set_status(S.status, ST_LIST_FULL); return; }
    if (S.flags & 32u) return;  // experiments: no stores into peer memory
    for (uint32_t p = 0; p < S.world; p++)
        if (p != S.rank) peer_entry(S, p, parity, S.rank)[at] = key;  // a store into peer memory: over NVLink, no reply awaited
}

template <int NT, bool SHARED_STEPS>
__device__ void heavy_rewrite(const TrainState& S, const Command& cmd, uint32_t n_words, Queue& Q) {
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint2* list = cmd.list_tok == NONE ? nullptr : S.wl_pool + __ldcg(S.wl_off + cmd.list_tok);
    const uint32_t n_cand = list ? cmd.list_len : n_words;
    const GlobalAdder add{S, cmd.threshold};
    const bool sharded = SHARED_STEPS && S.world > 1;  // every GPU filters the words it owns and tells the others what it found
    const uint32_t parity = cmd.heavy_id & 1u;
    // Candidates are dealt out evenly: a list of 5,000 postings gives every CTA of the grid a warp or two of them (and at most one
    // round of rewriting), instead of ten CTAs 512 each (and three rounds) while the others only report back.  A thread takes
    // up to HEAVY_R candidates per pass with their loads in flight together (postings, then symbols + lengths): an all-words step
    // over 2*10^7 words is 126 passes instead of 252.
    constexpr uint32_t HEAVY_R = 2;  // measured 2 / 3 / 4: 81.4 / 81.8 / 84.8 ms at 2.1 GB, 2.12 / - / 2.02 s at 11 GB; beyond 2 the kernel spills
    const uint32_t R = n_cand > gridDim.x * NT ? HEAVY_R : 1u;  // only when one candidate per thread of the grid would take several passes
    const uint32_t per = min((uint32_t)NT, max((uint32_t)HEAVY_MIN_THREADS, ((n_cand + gridDim.x * R - 1) / (gridDim.x * R) + 31u) & ~31u));  // threads in use
    const uint32_t chunk = per * R;  // <= QUEUE_CAP
    for (uint32_t base = blockIdx.x * chunk; base < n_cand; base += gridDim.x * chunk) {
        if (tid == 0) Q.n = 0;
        __syncthreads();
        uint32_t cw[HEAVY_R], coff[HEAVY_R], clen[HEAVY_R];
        uint4 cv[HEAVY_R];
        bool ok[HEAVY_R];
#pragma unroll
        for (uint32_t r = 0; r < HEAVY_R; r++) {
            const uint32_t i = base + r * per + tid;
            ok[r] = r < R && tid < per && i < n_cand;
            if (ok[r]) {
                if (list) { const uint2 e = __ldcg(list + i); cw[r] = e.x; coff[r] = e.y; }
                else { cw[r] = sharded ? (i | ((uint32_t)S.w_owner[i] << W_OWNER_SHIFT)) : i; coff[r] = S.w_off[i]; }
            }
        }
#pragma unroll
        for (uint32_t r = 0; r < HEAVY_R; r++) {
            ok[r] = ok[r] && (!sharded || (cw[r] >> W_OWNER_SHIFT) == S.rank);
            if (ok[r]) {
                cv[r] = __ldcg(reinterpret_cast<const uint4*>(S.sym + coff[r]));  // every word owns at least 4 symbol slots
                clen[r] = __ldcg(S.w_len + (cw[r] & W_MASK));
            }
        }
#pragma unroll
        for (uint32_t r = 0; r < HEAVY_R; r++) {
            if (ok[r] && word_has_pair_loaded(S, coff[r], cv[r], clen[r], cmd.a, cmd.b)) {
                const uint32_t w = cw[r], off = coff[r], len = clen[r];
                unsigned int q = atomicAdd(&Q.n, 1u);
                if (q < QUEUE_CAP) { Q.w[q] = w; Q.off[q] = off; Q.len[q] = (unsigned short)min(len, 0xFFFFu); }
                else if (rewrite_word_serial(S, w, off, len, cmd.a, cmd.b, cmd.x, add)) {
                    unsigned int at = atomicAdd(&S.ctrl->wl_cursor, 1u);
                    if (at < S.wl_cap) S.wl_pool[at] = make_uint2(w, off); else set_status(S.status, ST_LIST_FULL);
                    if (sharded) send_hit(S, parity, w);
                }
            }
        }
        __syncthreads();
        const uint32_t qn = min(Q.n, QUEUE_CAP);
        if (sharded)
            for (uint32_t k = tid; k < qn; k += NT) send_hit(S, parity, Q.w[k]);
        for (uint32_t k = warp; k < qn; k += NT / 32) {
            if (rewrite_word_warp(S, Q.w[k], Q.off[k], Q.len[k], cmd.a, cmd.b, cmd.x, add) && lane == 0) {
                unsigned int at = atomicAdd(&S.ctrl->wl_cursor, 1u);
                if (at < S.wl_cap) S.wl_pool[at] = make_uint2(Q.w[k], Q.off[k]); else set_status(S.status, ST_LIST_FULL);
            }
        }
        __syncthreads();
    }
}

// Second half of a sharded heavy step: the words the peers found in their shares, rewritten here as well (every GPU keeps
// all words).  A hit arrives as the pretoken's key: key -> slot of the context's table -> this GPU's word index.
template <int NT>
__device__ void apply_remote_hits(const TrainState& S, const Command& cmd) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t gwarp = blockIdx.x * (NT / 32) + (threadIdx.x >> 5), n_warps = gridDim.x * (NT / 32);
    const GlobalAdder add{S, cmd.threshold};
    const uint32_t parity = cmd.heavy_id & 1u;
    uint32_t base = 0;  // hits of the peers in front of p
    for (uint32_t p = 0; p < S.world; p++) {
        if (p == S.rank) continue;
        const uint32_t n_p = __ldcg(&S.ctrl->in_n[p]);
        const unsigned long long* hits = peer_entry(S, S.rank, parity, p);
        // warps take hits round-robin over the concatenation of all peers' lists
        for (uint32_t j = (gwarp + n_warps - base % n_warps) % n_warps; j < n_p; j += n_warps) {
            const unsigned long long key = __ldcg(hits + j);
            uint64_t slot = (key * 0xD6E8FEB86659FD93ull >> 17) & S.ut_mask;  // the slot sequence of unique_insert
            uint32_t w = NONE;
            for (uint32_t probe = 0; probe < 4096; probe++) {
                const unsigned long long cur = __ldg(S.ut_key + slot);
                if (cur == key) { w = __ldg(S.slot_word + slot); break; }
                if (cur == 0) break;
                slot = (slot + 1) & S.ut_mask;
            }
            if (w == NONE) { if (lane == 0) set_status(S.status, ST_LIST_FULL); continue; }  // cannot happen: every GPU holds the same pretokens
            const uint32_t off = __ldcg(S.w_off + w);
            const uint32_t tagged = w | (p << W_OWNER_SHIFT);
            if (rewrite_word_warp(S, tagged, off, 0xFFFFu, cmd.a, cmd.b, cmd.x, add) && lane == 0) {
                unsigned int at = atomicAdd(&S.ctrl->wl_cursor, 1u);
                if (at < S.wl_cap) S.wl_pool[at] = make_uint2(tagged, off); else set_status(S.status, ST_LIST_FULL);
            }
        }
        base += n_p;
    }
}

// SHARED_STEPS: the variant that shares heavy steps with the other GPUs of the box (kept out of the single-GPU kernel's code).
template <bool CLUSTER, int NT, bool SHARED_STEPS = false>
__global__ void __launch_bounds__(NT, 1) merge_loop_kernel(TrainState S) {
    // Driver <-> helper hand-over around a heavy step.  CLUSTER: the hardware cluster barrier on both sides.  Otherwise
    // (a cooperative grid, one CTA per SM): helpers poll a counter in global memory, so the driver's light steps involve
    // no barrier at all and a heavy step has all 148 SMs.
    uint32_t seen = 0;  // helper: heavy steps seen so far; driver: heavy steps issued so far
    auto helper_wait = [&] {
        if constexpr (CLUSTER) cg::this_cluster().sync();
        else {
            if (threadIdx.x == 0) {
                while (*reinterpret_cast<volatile unsigned int*>(&S.ctrl->go) == seen) __nanosleep(HELPER_SLEEP_NS);
                __threadfence();
            }
            seen++;
            __syncthreads();
        }
    };
    auto helper_done = [&] {
        if constexpr (CLUSTER) cg::this_cluster().sync();
        else {
            __syncthreads();
            if (threadIdx.x == 0) {
                if (SHARED_STEPS && S.world > 1) __threadfence_system(); else __threadfence();  // shared steps: the CTA's stores into peer memory come first, for everyone
                atomicAdd(&S.ctrl->done, 1u);
            }
        }
    };
    auto driver_wake = [&] {  // the command has been written by thread 0
        if constexpr (CLUSTER) cg::this_cluster().sync();
        else {
            seen++;
            if (threadIdx.x == 0) { __threadfence(); atomicAdd(&S.ctrl->go, 1u); }
        }
    };
    auto driver_join = [&] {
        if constexpr (CLUSTER) cg::this_cluster().sync();
        else {
            __syncthreads();
            if (threadIdx.x == 0) {
                const unsigned int want = seen * (gridDim.x - 1);
                unsigned int spins_ = 0;
                while (*reinterpret_cast<volatile unsigned int*>(&S.ctrl->done) < want) {
                    if ((S.flags & 16u) && (++spins_ & 0xFFFFu) == 0) { S.beat[8000] = *reinterpret_cast<volatile unsigned int*>(&S.ctrl->done); S.beat[8001] = want; S.beat[8002] = seen; }
                }
                __threadfence();
            }
            __syncthreads();
        }
    };
    extern __shared__ __align__(16) unsigned char smem_raw[];
    DriverShared& DS = *reinterpret_cast<DriverShared*>(smem_raw);  // used by the driver CTA only
    HotCache& H = DS.H;
    NewPairs& NP = DS.NP;
    Deferred& DQ = DS.DQ;
    uint32_t* sh_tie = DS.tie;
    unsigned int* sh_sub = DS.sub;
    __shared__ long long sh_wmax[NT / 32];
    __shared__ unsigned int sh_ntie;
    __shared__ Best sh_win;
    __shared__ Best sh_bestw[NT / 32];
    __shared__ Queue Q;
    __shared__ unsigned int sh_hist[64];
    __shared__ long long sh_T;
    __shared__ int sh_bit;
    __shared__ unsigned int sh_above;
    __shared__ Command sh_cmd;
    __shared__ bool sh_stop;
    __shared__ unsigned long long sh_px;  // prefix key of the step's new token
    __shared__ bool sh_blob_full;         // the token byte pool is full: leave the loop, the host enlarges it and trains again
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t n_words = S.ctrl->n_words;

    // Which CTA drives: block 0, or (experiments: BPE_B200_TRAIN_DRIVER_SM) the CTA that runs on a given SM — the loop is pure
    // latency, and the distance from the driver's SM to the L2 slices differs between SMs of a two-die part.
    bool is_driver = blockIdx.x == 0;
    if (!CLUSTER && S.driver_sm >= 0 && gridDim.x > (unsigned)S.driver_sm) {
        unsigned int smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        is_driver = (int)smid == S.driver_sm;
    }
    if (!is_driver) {  // helper CTAs: asleep on the barrier until the driver broadcasts a heavy step
        for (uint32_t round = 0;; round++) {
            if (tid == 0 && blockIdx.x < 16 && blockIdx.x > 0) { S.beat[2 * blockIdx.x] = round; S.beat[2 * blockIdx.x + 1] = 20; }
            if (tid == 0 && (S.flags & 16u)) { S.beat[64 + 2 * blockIdx.x] = round; S.beat[64 + 2 * blockIdx.x + 1] = 20; }
            helper_wait();
            if (tid == 0) {
                if (blockIdx.x < 16 && blockIdx.x > 0) S.beat[2 * blockIdx.x + 1] = 21;
                const volatile Command* c = &S.ctrl->cmd;
                sh_cmd.op = c->op; sh_cmd.a = c->a; sh_cmd.b = c->b; sh_cmd.x = c->x; sh_cmd.list_tok = c->list_tok;
                sh_cmd.list_len = c->list_len; sh_cmd.threshold = c->threshold; sh_cmd.heavy_id = c->heavy_id;
            }
            __syncthreads();
            const Command cmd = sh_cmd;
            if (cmd.op == 0) return;
            if (SHARED_STEPS && cmd.op == 2) apply_remote_hits<NT>(S, cmd);
            else heavy_rewrite<NT, SHARED_STEPS>(S, cmd, n_words, Q);
            if (tid == 0 && blockIdx.x < 16 && blockIdx.x > 0) S.beat[2 * blockIdx.x + 1] = 22;
            if (tid == 0 && (S.flags & 16u)) S.beat[64 + 2 * blockIdx.x + 1] = 22;
            helper_done();
        }
    }

    // ---- driver CTA -------------------------------------------------------------------------------------------------
    const bool leader = tid == NT - 32;  // a warp of its own: its bookkeeping runs beside the other warps' work
    const bool tracer = S.trace && tid == 0;
    uint32_t next_id = S.first_new_id, iter = 0, wl_cur = 0, n_heavy = 0;
    bool need_rebuild = true;
    bool use_cache = !(S.flags & 1u);  // false: counts too flat for the cache (see the rebuild): argmax over the global table
    long long T = LLONG_MAX, failed_T = 0;
    uint32_t spins = 0;           // loop passes that did not finish a step (rebuilds): bounded, a bug must not hang the GPU
    unsigned long long t_prev = 0;
    if (S.trace && leader) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_prev));

    unsigned long long t_start;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_start));
    uint32_t passes = 0;
    const bool legacy_adders = (S.flags & 8u) != 0;  // experiments: every light step through CacheAdder, as before
    const bool beats = (S.flags & 16u) != 0;         // debugging: heartbeat (iteration, phase) in pinned host memory on every step;
                                                     // otherwise only with the watchdog check — a store over PCIe per phase is not free
    constexpr int PEND = (int)((NPS + NT - 1) / NT);  // new-pair slots per thread when a light step closes
    int pend_e[PEND];                   // cache entries whose prefix key this thread has asked global memory for
    unsigned long long pend_v[PEND];
#pragma unroll
    for (int r = 0; r < PEND; r++) { pend_e[r] = -1; pend_v[r] = 0; }
    for (uint32_t k = tid; k < NPS; k += NT) { NP.key[k] = 0; NP.cnt[k] = 0; }
    if (tid == 0) { NP.overflow = 0; DQ.n = 0; sh_blob_full = false; }
    __syncthreads();
    while (next_id < S.vocab_size) {
        if ((++passes & 255u) == 0) {  // watchdog: a bug must not hang the GPU (default 120 s, far beyond any real training)
            if (tid == 0) {            // one thread decides, so that the whole CTA leaves together
                unsigned long long now;
                asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
                sh_stop = now - t_start > (unsigned long long)S.watchdog_s * 1000000000ull && now - t_start < (1ull << 62);
                if (sh_stop) { set_status(S.status, ST_WATCHDOG); S.ctrl->dbg[4] = iter; S.ctrl->dbg[5] = passes; }
                S.beat[0] = iter; S.beat[1] = 3;
            }
            __syncthreads();
            if (sh_stop) break;
        }
        // Cache invariant: every pair with count >= T has an entry (pairs are appended when their count rises past T;
        // leftovers below T and holes are allowed).  If the best entry has count >= T it is the global argmax;
        // otherwise T is lowered and the cache rebuilt from a log2 histogram of all counts.
        if (beats && tid == 0) { S.beat[0] = iter; S.beat[1] = need_rebuild ? 1 : 2; }
        if (need_rebuild && use_cache) {
            if (++spins > 64) { if (tid == 0) set_status(S.status, ST_PAIR_TABLE_FULL); break; }
#pragma unroll
            for (int r = 0; r < PEND; r++) pend_e[r] = -1;  // the entries they belong to are about to be wiped
            deferred_drain<NT>(S, H, DQ);  // the rebuild reads global counts: queued updates first (ends with a barrier:
                                           // every thread has read H.n / the winner that sent it here before the cache is wiped)
            if (tid < 64) sh_hist[tid] = 0;
            for (uint32_t k = tid; k < HIDX; k += NT) H.index[k] = 0;
            if (tid == 0) H.n = 0;
            __syncthreads();
            const uint32_t n_all = __ldcg(&S.ctrl->n_live);
            for (uint32_t i = tid; i < n_all; i += NT) {
                long long cnt = __ldcg(S.ph_cnt + __ldcg(S.live + i));
                if (cnt > 0) atomicAdd(&sh_hist[63 - __clzll(cnt)], 1u);
            }
            __syncthreads();
            if (tid == 0) {  // the bin (power of two) in which the cache_fill-th largest count lies
                unsigned int above = sh_hist[63];
                int bit = 63;
                for (; bit > 0; bit--) {
                    if (above + sh_hist[bit - 1] > S.cache_fill) break;
                    above += sh_hist[bit - 1];
                }
                sh_bit = bit;      // bins >= bit fit entirely (`above` pairs); bin bit - 1 does not (unless bit == 0)
                sh_above = above;
                sh_T = 1ll << bit;
                S.ctrl->n_rebuilds++;
            }
            for (uint32_t k = tid; k < SUB_BINS; k += NT) sh_sub[k] = 0;
            __syncthreads();
            if (sh_bit > 0) {  // refine inside bin bit - 1 = [lo, 2 lo): SUB_BINS equal slices, take the upper ones that still fit
                const long long lo = 1ll << (sh_bit - 1), width = lo >= SUB_BINS ? lo / SUB_BINS : 1;
                for (uint32_t i = tid; i < n_all; i += NT) {
                    long long cnt = __ldcg(S.ph_cnt + __ldcg(S.live + i));
                    if (cnt >= lo && cnt < 2 * lo) atomicAdd(&sh_sub[(cnt - lo) / width], 1u);
                }
                __syncthreads();
                if (tid == 0) {
                    unsigned int above = sh_above;
                    long long slice = (lo + width - 1) / width;  // slices in use
                    while (slice > 0 && above + sh_sub[slice - 1] <= S.cache_fill) above += sh_sub[--slice];
                    sh_T = lo + slice * width;
                }
                __syncthreads();
            }
            T = sh_T;
            // The previous threshold found nothing and the next lower one does not fit (many pairs with equal small
            // counts): leave the cache and take the argmax over the whole table from here on.
            if (T == failed_T) use_cache = false;
            for (uint32_t i = tid; use_cache && i < n_all; i += NT) {
                uint32_t slot = __ldcg(S.live + i);
                long long cnt = __ldcg(S.ph_cnt + slot);
                if (cnt >= T) {
                    const unsigned long long key = __ldcg(S.ph_key + slot);
                    hc_append(H, key, slot, cnt, __ldcg(S.tk_pre + (uint32_t)((key - 1) >> 32)));
                }
            }
            __syncthreads();
            need_rebuild = false;
        }
        if (!use_cache) T = LLONG_MAX;  // GlobalAdder never reports crossings

        if ((S.flags & 2u) && use_cache) {  // self-check (experiments): the cache against the global table
            deferred_drain<NT>(S, H, DQ);
            const uint32_t n_c = min(H.n, HC);
            for (uint32_t i = tid; i < n_c; i += NT)
                if (H.key[i] && H.cnt[i] != __ldcg(S.ph_cnt + H.slot[i])) {
                    if (atomicAdd(&S.ctrl->dbg[0], 1u) == 0) S.ctrl->dbg[1] = iter;
                }
            const uint32_t n_all = __ldcg(&S.ctrl->n_live);
            for (uint32_t i = tid; i < n_all; i += NT) {
                uint32_t slot = __ldcg(S.live + i);
                if (H.n <= HC && __ldcg(S.ph_cnt + slot) >= T && hc_find(H, __ldcg(S.ph_key + slot)) < 0) {
                    if (atomicAdd(&S.ctrl->dbg[2], 1u) == 0) S.ctrl->dbg[3] = iter;
                }
            }
            __syncthreads();
        }
        // ---- argmax over the cached pairs of (count, bytes(a), bytes(b)): shared memory only -------------------------------
        // Two steps, so that byte strings are compared only for pairs that really tie at the top count (comparing
        // them inside the reduction would chase token bytes through global memory for every tie among also-rans):
        // (1) the maximum count, (2) the entries that have it — almost always exactly one.
        long long c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0;
        if (tracer) c0 = clock64();
        const uint32_t n_cached = use_cache ? H.n : 0;
        if (n_cached > HC) { need_rebuild = true; continue; }  // overflowed: compact (holes, leftovers) or raise T
        const uint32_t n_scan = use_cache ? n_cached : __ldcg(&S.ctrl->n_live);
        long long my_max = 0;
        if (use_cache) {  // holes and retired entries carry a negative count: no need to look at the keys
            for (uint32_t i = tid; i < n_scan; i += NT) my_max = max(my_max, H.cnt[i]);
        } else {
            for (uint32_t i = tid; i < n_scan; i += NT) my_max = max(my_max, __ldcg(S.ph_cnt + __ldcg(S.live + i)));
        }
        const long long my_own = my_max;
        long long a1 = 0, a2 = 0, a3 = 0, a4 = 0;
        if (tracer) a1 = clock64();
        for (int o = 16; o; o >>= 1) my_max = max(my_max, __shfl_xor_sync(0xFFFFFFFFu, my_max, o));
        if (lane == 0) sh_wmax[warp] = my_max;
        if (tid == 0) sh_ntie = 0;
        // Prefix keys requested when the previous step closed have had the scan to arrive.
#pragma unroll
        for (int r = 0; r < PEND; r++)
            if (pend_e[r] >= 0) { H.pre[pend_e[r]] = pend_v[r]; pend_e[r] = -1; }
        __syncthreads();
        if (tracer) a2 = clock64();
        long long top = lane < NT / 32 ? sh_wmax[lane] : 0;  // every warp folds the per-warp maxima itself: no second barrier
        for (int o = 16; o; o >>= 1) top = max(top, __shfl_xor_sync(0xFFFFFFFFu, top, o));
        if (top > 0 && my_own == top) {  // only a thread that saw the top count looks again, for the entries that have it
            if (use_cache) {
                for (uint32_t i = tid; i < n_scan; i += NT)
                    if (H.cnt[i] == top) { uint32_t k = atomicAdd(&sh_ntie, 1u); if (k < MAX_TIES) sh_tie[k] = i; }
            } else {
                for (uint32_t i = tid; i < n_scan; i += NT) {
                    const uint32_t slot = __ldcg(S.live + i);
                    if (__ldcg(S.ph_cnt + slot) == top) { uint32_t k = atomicAdd(&sh_ntie, 1u); if (k < MAX_TIES) sh_tie[k] = slot; }
                }
            }
        }
        if (tracer) a3 = clock64();
        __syncthreads();
        if (tracer) a4 = clock64();
        Best win{0, 0, 0};
        if (top > 0 && sh_ntie > MAX_TIES) {
            // very many pairs tie at the top (late in a small corpus): the general reduction, restricted to the tied entries
            Best best{0, 0, 0};
            for (uint32_t i = tid; i < n_scan; i += NT) {
                const uint32_t e = use_cache ? i : __ldcg(S.live + i);
                const long long cnt = use_cache ? (H.key[i] ? H.cnt[i] : 0) : __ldcg(S.ph_cnt + e);
                if (cnt != top) continue;
                Best c{top, use_cache ? H.key[i] : __ldcg(S.ph_key + e), e};
                if (best.cnt == 0 || better(S, c, best)) best = c;
            }
            for (int o = 16; o; o >>= 1) {
                Best other = shfl_best(best, o, false);
                if (other.cnt > 0 && (best.cnt == 0 || better(S, other, best))) best = other;
            }
            if (lane == 0) sh_bestw[warp] = best;
            __syncthreads();
            if (warp == 0) {
                Best bb = lane < NT / 32 ? sh_bestw[lane] : Best{0, 0, 0};
                for (int o = 16; o; o >>= 1) {
                    Best other = shfl_best(bb, o, false);
                    if (other.cnt > 0 && (bb.cnt == 0 || better(S, other, bb))) bb = other;
                }
                if (lane == 0) sh_win = bb;
            }
            __syncthreads();
            win = sh_win;
        } else {  // the tied entries (usually one): order by bytes.  Every warp does it for itself: no hand-over, no barrier.
            const uint32_t n_tie = top > 0 ? sh_ntie : 0;
            Best bb{0, 0, 0};
            // equal counts: order by (bytes(a), bytes(b)).  With the cache, different first tokens whose prefix keys
            // differ are ordered from shared memory; anything else falls back to the token bytes in global memory.
            auto wins = [&](const Best& c, const Best& d) -> bool {  // c before d ?
                if (use_cache && (c.key >> 32) != (d.key >> 32)) {  // key - 1 cannot carry into the high word for b < 2^32 - 1
                    const unsigned long long pc = H.pre[c.slot], pd = H.pre[d.slot];
                    if (pc != PRE_UNKNOWN && pd != PRE_UNKNOWN && pc != pd) return pc > pd;
                }
                return better(S, c, d);
            };
            if (n_tie == 1) {
                const uint32_t e = sh_tie[0];
                bb = Best{top, use_cache ? H.key[e] : __ldcg(S.ph_key + e), e};
            } else if (n_tie > 1) {
                for (uint32_t k = lane; k < n_tie; k += 32) {
                    const uint32_t e = sh_tie[k];
                    Best c{top, use_cache ? H.key[e] : __ldcg(S.ph_key + e), e};
                    if (bb.cnt == 0 || wins(c, bb)) bb = c;
                }
                for (int o = 16; o; o >>= 1) {
                    Best other = shfl_best(bb, o, false);
                    if (other.cnt > 0 && (bb.cnt == 0 || wins(other, bb))) bb = other;
                }
                bb.cnt = __shfl_sync(0xFFFFFFFFu, bb.cnt, 0); bb.key = __shfl_sync(0xFFFFFFFFu, bb.key, 0); bb.slot = __shfl_sync(0xFFFFFFFFu, bb.slot, 0);
            }
            win = bb;
            if (tracer) S.ctrl->sub_clk[6] += n_tie;
        }
        if (tracer) {
            c1 = clock64();
            S.ctrl->sub_clk[0] += a1 - c0; S.ctrl->sub_clk[1] += a2 - a1; S.ctrl->sub_clk[2] += a3 - a2; S.ctrl->sub_clk[3] += a4 - a3;
            S.ctrl->sub_clk[4] += c1 - a4; S.ctrl->sub_clk[5] += n_scan;
        }
        // win.slot = index of the cache entry (global slot without the cache)
        if (!use_cache) {
            if (win.cnt <= 0) break;  // nothing qualifies (bpe_trainer.py:242: heap exhausted)
        } else if (win.cnt < T) {     // the best cached pair is below the threshold: pairs outside the cache may beat it
            if (T <= 1) break;        // the cache held every pair with a positive count: nothing qualifies
            failed_T = T;
            need_rebuild = true;
            continue;
        }
        spins = 0;
        const uint32_t a = (uint32_t)((win.key - 1) >> 32), b = (uint32_t)(win.key - 1), x = next_id;

        // ---- the words that can contain (a, b): the shorter posting list of the two tokens, all words for two bytes ------
        uint32_t list_tok = NONE, list_len = n_words, list_off = 0;
        {   // both tokens' posting heads in one round trip
            const bool ma = a >= S.first_new_id, mb = b >= S.first_new_id;
            const uint32_t la = ma ? __ldcg(S.wl_len + a) : 0, oa = ma ? __ldcg(S.wl_off + a) : 0;
            const uint32_t lb = mb ? __ldcg(S.wl_len + b) : 0, ob = mb ? __ldcg(S.wl_off + b) : 0;
            if (ma) { list_tok = a; list_len = la; list_off = oa; }
            if (mb && (!ma || lb < la)) { list_tok = b; list_len = lb; list_off = ob; }
        }
        const uint32_t wl_start = wl_cur;
        const uint32_t first_new_entry = n_cached;
        uint32_t matched = 0;
        if (tid == 0) Q.n = 0;
        if (warp == NT / 32 - 1) {  // bookkeeping for x beside the rewrite (speculative: overwritten if the step merges nothing)
            // The last warp copies the bytes of a and b behind one another, a lane per byte: all loads are in flight together
            // (a one-thread loop would pay a round trip to L2 per byte, and the whole CTA waits for it at the next barrier).
            const uint32_t oa = __ldcg(S.tk_off + a), ob = __ldcg(S.tk_off + b), o = __ldcg(S.tk_off + x);
            const uint32_t la = __ldcg(S.tk_off + a + 1) - oa, lb = __ldcg(S.tk_off + b + 1) - ob;
            uint32_t px_hi = 0, px_lo = 0;
            if (o + la + lb <= S.blob_cap) {
                for (uint32_t i0 = 0; i0 < la + lb; i0 += 32) {
                    const uint32_t i = i0 + lane;
                    if (i < la + lb) {
                        const uint32_t v = i < la ? __ldcg(S.tk_blob + oa + i) : __ldcg(S.tk_blob + ob + i - la);
                        S.tk_blob[o + i] = (uint8_t)v;
                        if (i < 4) px_hi = v << (24 - 8 * i);
                        else if (i < 8) px_lo = v << (56 - 8 * i);
                    }
                }
            } else if (lane == 0) { set_status(S.status, ST_BLOB_FULL); sh_blob_full = true; }
            px_hi = __reduce_or_sync(0xFFFFFFFFu, px_hi);
            px_lo = __reduce_or_sync(0xFFFFFFFFu, px_lo);
            if (lane == 0) {
                S.tk_off[x + 1] = o + la + lb;
                S.tk_pre[x] = ((unsigned long long)px_hi << 32) | px_lo;
                sh_px = ((unsigned long long)px_hi << 32) | px_lo;
                S.m_a[x - S.first_new_id] = a; S.m_b[x - S.first_new_id] = b;
                S.wl_off[x] = wl_start;
            }
        }
        if (tracer) c2 = clock64();
        if (list_tok != NONE && list_len <= LIGHT_MAX) {
            // light step, this CTA alone: filter (one thread per posting), then one warp per hit
            const uint2* list = S.wl_pool + list_off;
            const bool light_adders = use_cache && !legacy_adders;
            // The updates deferred by the previous step: the first compare-and-swap of each is issued here, in front of
            // the posting loads, and looked at behind the filter — its round trip disappears behind theirs.
            const unsigned int dq_n = min(DQ.n, DQ_CAP);
            unsigned long long dq_key = 0, dq_cur = 0;
            uint64_t dq_slot = 0;
            if (tid < dq_n) {
                dq_key = DQ.key[tid];
                dq_slot = (dq_key * 0x9E3779B97F4A7C15ull >> 20) & S.ph_mask;
                dq_cur = atomicCAS(S.ph_key + dq_slot, 0ull, dq_key);
            }
            __syncthreads();  // Q.n = 0; every thread has read DQ.n
            if (tid == 0) DQ.n = 0;  // nothing is queued between the two barriers around the filter; the items stay readable
            // Four postings per thread and pass: the posting loads of a pass are in flight together, then the symbol
            // loads — a list of up to 4 * NT words costs two round trips, not eight.
            for (uint32_t i0 = tid; i0 < list_len; i0 += 4 * NT) {
                uint2 e[4];
                uint4 v[4];
                uint32_t len[4];
#pragma unroll
                for (int r = 0; r < 4; r++)
                    if (i0 + r * NT < list_len) e[r] = __ldcg(list + i0 + r * NT);
#pragma unroll
                for (int r = 0; r < 4; r++)
                    if (i0 + r * NT < list_len) {
                        v[r] = __ldcg(reinterpret_cast<const uint4*>(S.sym + e[r].y));
                        len[r] = __ldcg(S.w_len + (e[r].x & W_MASK));
                    }
#pragma unroll
                for (int r = 0; r < 4; r++)
                    if (i0 + r * NT < list_len && word_has_pair_loaded(S, e[r].y, v[r], len[r], a, b)) {  // list_len <= QUEUE_CAP
                        const unsigned int q = atomicAdd(&Q.n, 1u);
                        Q.w[q] = e[r].x; Q.off[q] = e[r].y; Q.len[q] = (unsigned short)min(len[r], 0xFFFFu);
                    }
            }
            if (tid < dq_n) {
                const unsigned short entry = DQ.entry[tid];
                unsigned long long pre = PRE_UNKNOWN;
                const bool want_pre = entry != 0xFFFFu && H.pre[entry] == PRE_UNKNOWN;
                if (want_pre) pre = __ldcg(S.tk_pre + (uint32_t)((dq_key - 1) >> 32));
                const uint64_t slot = pair_slot_resume(S, dq_key, dq_slot, dq_cur);
                global_add(S, slot, DQ.delta[tid]);
                if (entry != 0xFFFFu) { H.slot[entry] = (uint32_t)slot; if (want_pre) H.pre[entry] = pre; }
            }
            for (unsigned int i = tid + NT; i < dq_n; i += NT) deferred_apply(S, H, DQ.key[i], DQ.delta[i], DQ.entry[i]);
            __syncthreads();
            matched = Q.n;
            if (beats && tid == 0) S.beat[1] = 5;
            if (tracer) c3 = clock64();
            if (light_adders) {
                const LightAdder add{S, H, NP, DQ, T, x};
                for (uint32_t k = warp; k < matched; k += NT / 32) rewrite_word_warp(S, Q.w[k], Q.off[k], Q.len[k], a, b, x, add);
            } else if (use_cache) {
                const CacheAdder add{S, H, T};
                for (uint32_t k = warp; k < matched; k += NT / 32) rewrite_word_warp(S, Q.w[k], Q.off[k], Q.len[k], a, b, x, add);
            } else {
                const GlobalAdder add{S, LLONG_MAX};
                for (uint32_t k = warp; k < matched; k += NT / 32) rewrite_word_warp(S, Q.w[k], Q.off[k], Q.len[k], a, b, x, add);
            }
            for (uint32_t k = tid; k < matched; k += NT) {  // the hits ARE x's posting list
                if (wl_start + k < S.wl_cap) S.wl_pool[wl_start + k] = make_uint2(Q.w[k], Q.off[k]); else set_status(S.status, ST_LIST_FULL);
            }
            if (beats && tid == 0) S.beat[1] = 6;
            if (light_adders) {
                __syncthreads();
                // Close the step's new pairs: exact counts, straight from shared memory.
                const uint32_t n_after = min(H.n, HC);  // beyond first_new_entry: appended by the CacheAdder fallback, if any
                const bool fell_back = NP.overflow != 0;
                __syncthreads();
#pragma unroll
                for (int r = 0; r < PEND; r++) {
                    const uint32_t k = tid + r * NT;
                    pend_e[r] = -1;
                    if (k >= NPS) continue;
                    const uint32_t code = NP.key[k];
                    if (!code) continue;
                    const long long c = NP.cnt[k];
                    NP.key[k] = 0; NP.cnt[k] = 0;
                    const uint32_t other = (code - 1) >> 1;
                    const uint32_t pa = ((code - 1) & 1u) ? x : other, pb = ((code - 1) & 1u) ? other : x;
                    const unsigned long long key = (((unsigned long long)pa << 32) | pb) + 1ull;
                    int e = -1;
                    if (c >= T) {
                        e = hc_append(H, key, 0u, c, pa == x ? sh_px : PRE_UNKNOWN);  // the global slot arrives with the deferred update
                        if (e >= 0 && pa != x) {  // the prefix key of `other`: asked for now, stored on the way into the next tie-break
                            pend_e[r] = e;
                            pend_v[r] = __ldcg(S.tk_pre + other);
                        }
                    }
                    deferred_push(S, H, DQ, key, c, e >= 0 ? (unsigned short)e : (unsigned short)0xFFFFu);
                }
                if (fell_back) {  // rare: more distinct new pairs than NewPairs holds; those went the CacheAdder way
                    __threadfence();
                    __syncthreads();
                    for (uint32_t i = first_new_entry + tid; i < n_after; i += NT)
                        if (H.key[i]) H.cnt[i] = __ldcg(S.ph_cnt + H.slot[i]);
                    if (tid == 0) NP.overflow = 0;
                }
            } else {
                __threadfence();  // the count updates were posted without waiting: they must have landed before the reload below
                __syncthreads();
                // entries appended during this step: their count comes from the (exact) global table
                const uint32_t n_now = use_cache ? min(H.n, HC) : 0;
                for (uint32_t i = first_new_entry + tid; i < n_now; i += NT)
                    if (H.key[i]) H.cnt[i] = __ldcg(S.ph_cnt + H.slot[i]);
            }
        } else {
            // heavy step: wake the helper CTAs; counts change behind the cache's back, so reload it afterwards
            long long h0 = 0, h1 = 0, h2 = 0, h3 = 0, h4 = 0;
            if (tracer) h0 = clock64();
            deferred_drain<NT>(S, H, DQ);  // queued updates land (and new cache entries learn their slots) before anyone reads global counts
            if (tracer) h1 = clock64();
            if (tid == 0) {
                S.ctrl->wl_cursor = wl_start;
                S.ctrl->n_hot = 0;
                S.ctrl->out_n = 0;
                S.ctrl->cmd = Command{1u, a, b, x, list_tok, list_len, T, n_heavy, 0u};
                __threadfence();
                if (beats) S.beat[1] = 7;
            }
            driver_wake();
            if (beats && tid == 0) S.beat[1] = 8;
            if (tracer) { c3 = clock64(); h2 = c3; }
            const Command cmd{1u, a, b, x, list_tok, list_len, T, n_heavy, 0u};
            heavy_rewrite<NT, SHARED_STEPS>(S, cmd, n_words, Q);
            if (tracer) h3 = clock64();
            if (beats && tid == 0) S.beat[1] = 9;
            driver_join();
            if (tracer) h4 = clock64();
            if (SHARED_STEPS && S.world > 1) {
                // ---- the exchange: every GPU has filtered and rewritten its share and put the keys of its hits into the peers'
                // buffers; now the counts go out (behind the hits: system-wide fence), the peers' counts come in, and the
                // peers' hits are rewritten here too.  Same step on every GPU: they all run the same loop on the same data.
                const uint32_t parity = n_heavy & 1u;
                if (tid == 0) {
                    __threadfence_system();
                    const unsigned long long flag = ((unsigned long long)(n_heavy + 1u) << 32) | __ldcg(&S.ctrl->out_n);
                    for (uint32_t p = 0; p < S.world; p++)
                        if (p != S.rank) *reinterpret_cast<volatile unsigned long long*>(peer_flag(S, p, parity, S.rank)) = flag;
                    __threadfence_system();
                    if (beats) { S.beat[0] = iter; S.beat[1] = 30; }
                    unsigned long long t_wait0;
                    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_wait0));
                    for (uint32_t p = 0; p < S.world; p++) {
                        if (p == S.rank) { S.ctrl->in_n[p] = 0; continue; }
                        unsigned long long v;
                        uint32_t polls = 0;
                        for (;;) {
                            v = *reinterpret_cast<volatile unsigned long long*>(peer_flag(S, S.rank, parity, p));
                            if ((uint32_t)(v >> 32) == n_heavy + 1u) break;
                            if ((++polls & 1023u) == 0) {  // a peer that never arrives must not hang this GPU for ever
                                unsigned long long now;
                                asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
                                if (now - t_wait0 > (unsigned long long)S.watchdog_s * 1000000000ull) {
                                    set_status(S.status, ST_WATCHDOG);
                                    S.ctrl->dbg[4] = iter; S.ctrl->dbg[5] = 0x1000u + p; S.ctrl->dbg[0] = (uint32_t)(v >> 32); S.ctrl->dbg[1] = (uint32_t)v;
                                    v = 0;
                                    break;
                                }
                            }
                        }
                        S.ctrl->in_n[p] = (uint32_t)v;
                    }
                    __threadfence_system();
                    if (beats) S.beat[1] = 31;
                    S.ctrl->cmd = Command{2u, a, b, x, list_tok, list_len, T, n_heavy, 0u};
                    __threadfence();
                }
                __syncthreads();  // the rest of this CTA reads the peers' counts too
                driver_wake();
                const Command cmd2{2u, a, b, x, list_tok, list_len, T, n_heavy, 0u};
                apply_remote_hits<NT>(S, cmd2);
                driver_join();
            }
            if (beats && tid == 0) S.beat[1] = 10;
            matched = __ldcg(&S.ctrl->wl_cursor) - wl_start;
            for (uint32_t i = tid; i < n_cached; i += NT)
                if (H.key[i]) H.cnt[i] = __ldcg(S.ph_cnt + H.slot[i]);
            const uint32_t n_pending = use_cache ? min(__ldcg(&S.ctrl->n_hot), (unsigned int)S.ph_mask) : 0;  // pairs that rose past T in the heavy step
            __syncthreads();
            for (uint32_t i = tid; i < n_pending; i += NT) {
                const ulonglong2 rec = __ldcg(S.hot + i);
                if (hc_find(H, rec.x) < 0)
                    hc_append(H, rec.x, (uint32_t)rec.y, __ldcg(S.ph_cnt + (uint32_t)rec.y), __ldcg(S.tk_pre + (uint32_t)((rec.x - 1) >> 32)));
            }
            n_heavy++;
            if (tracer) {
                const long long h5 = clock64();
                S.ctrl->heavy_clk[0] += h1 - h0; S.ctrl->heavy_clk[1] += h2 - h1; S.ctrl->heavy_clk[2] += h3 - h2; S.ctrl->heavy_clk[3] += h4 - h3;
                S.ctrl->heavy_clk[4] += h5 - h4; S.ctrl->heavy_clk[5] += h0 - c0;
            }
        }
        if (tracer) c4 = clock64();
        // ---- close the step ---------------------------------------------------------------------------------------------
        if (beats && tid == 0) S.beat[1] = 11;
        if (leader) {
            if (use_cache) {
                H.cnt[win.slot] = RETIRED;                 // the winner never occurs again (bpe_trainer.py:243);
                S.ph_cnt[H.slot[win.slot]] = RETIRED;      // a pair without any occurrence is retired the same way
            } else {
                S.ph_cnt[win.slot] = RETIRED;
            }
            if (matched) S.wl_len[x] = matched;
            // no fence here: this CTA sees the stores after the barrier below, and every hand-over to other CTAs (heavy
            // step) or re-read of global counts (rebuild) starts with deferred_drain, whose fence covers them
        }
        if (matched) {  // bpe_trainer.py:309-316
            wl_cur = wl_start + matched;
            next_id++;
        }
        __syncthreads();  // cache and leader stores are ordered before the next argmax of this CTA
        if (sh_blob_full) break;  // written by the bookkeeping warp before the barrier: every thread sees the same value
        if (tracer) {
            long long c5 = clock64();
            if (list_tok != NONE && list_len <= LIGHT_MAX) {
                S.ctrl->phase_clk[0] += c1 - c0; S.ctrl->phase_clk[1] += c2 - c1; S.ctrl->phase_clk[2] += c3 - c2;
                S.ctrl->phase_clk[3] += c4 - c3; S.ctrl->phase_clk[4] += c5 - c4; S.ctrl->phase_clk[6]++;
            } else {
                S.ctrl->phase_clk[5] += c5 - c0; S.ctrl->phase_clk[7]++;
            }
        }
        if (S.trace && leader && iter < S.vocab_size) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
            S.trace[4 * iter + 0] = (uint32_t)(now - t_prev); S.trace[4 * iter + 1] = list_len;
            S.trace[4 * iter + 2] = list_tok == NONE; S.trace[4 * iter + 3] = matched;
            t_prev = now;
        }
        iter++;
    }
    if (tid == 0) {
        S.ctrl->n_iters = iter; S.ctrl->n_merges = next_id - S.first_new_id; S.ctrl->n_heavy = n_heavy;
        unsigned int smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        S.ctrl->dbg[7] = smid;
        S.ctrl->cmd.op = 0;
        unsigned long long t_now;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_now));
        S.ctrl->t_begin = t_start; S.ctrl->t_end = t_now;
        __threadfence();
        S.beat[1] = 12;
    }
    driver_wake();  // releases the helpers, which read op = 0 and leave
}

size_t next_pow2(size_t v) { size_t p = 1; while (p < v) p <<= 1; return p; }

}  // namespace

static int run_train_once(bpe_ctx* ctx, uint32_t vocab_size, uint32_t first_new_id, const uint8_t* h_init_blob, const uint32_t* h_init_off,
                          uint32_t* h_merge_a, uint32_t* h_merge_b, uint32_t* n_merges, uint64_t blob_per_token, bool* blob_full, cudaStream_t s);

// The pool of token byte strings starts at 256 bytes per vocabulary entry; a corpus whose tokens grow longer than that
// (long runs of one character double a token's length with every merge) fills it, the loop stops, and the training
// runs again — it is deterministic and rebuilds its words from the context's table — with a pool 16 times larger.
int run_train(bpe_ctx* ctx, uint32_t vocab_size, uint32_t first_new_id, const uint8_t* h_init_blob, const uint32_t* h_init_off,
              uint32_t* h_merge_a, uint32_t* h_merge_b, uint32_t* n_merges, cudaStream_t s) {
    uint64_t per_token = 256;
    if (const char* f = getenv("BPE_B200_TRAIN_BLOB_PER_TOKEN")) per_token = (uint64_t)std::max(1, atoi(f));  // tests: force the retry
    for (;;) {
        bool blob_full = false;
        const int rc = run_train_once(ctx, vocab_size, first_new_id, h_init_blob, h_init_off, h_merge_a, h_merge_b, n_merges, per_token, &blob_full, s);
        if (!blob_full) return rc;
        if ((uint64_t)h_init_off[first_new_id] + (uint64_t)vocab_size * per_token * 16 >= 0xFFFFFFF0ull) return ctx_status_to_code(ctx, ST_BLOB_FULL);
        per_token *= 16;
    }
}

static int run_train_once(bpe_ctx* ctx, uint32_t vocab_size, uint32_t first_new_id, const uint8_t* h_init_blob, const uint32_t* h_init_off,
                          uint32_t* h_merge_a, uint32_t* h_merge_b, uint32_t* n_merges, uint64_t blob_per_token, bool* blob_full, cudaStream_t s) {
    *n_merges = 0;
    if (!ctx->ut_alloc) return ctx_fail(ctx, BPE_ERR_NO_PAIRS, "no pretokens counted");
    uint64_t h_counters[4];
    CUDA_TRY(ctx, cudaMemcpyAsync(h_counters, ctx->ut.counters, sizeof(h_counters), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    const uint64_t n_unique = h_counters[1], n_sym = h_counters[0];
    if (n_unique == 0 || n_sym + 4 * n_unique >= 0xFFFFFFF0ull || n_unique > W_MASK) return ctx_fail(ctx, n_unique ? BPE_ERR_CAPACITY : BPE_ERR_NO_PAIRS, "word table size");
    if (vocab_size < first_new_id) vocab_size = first_new_id;  // nothing to learn, but "no pairs" is still an error

    // Progress words in pinned host memory (see below); allocated here, in front of everything timed: the first cudaHostAlloc of a
    // process can take milliseconds, and the GPU would idle between the pair histogram and the loop's launch.
    if (!ctx->h_beat) CUDA_TRY(ctx, cudaHostAlloc((void**)&ctx->h_beat, 2 * 4096 * sizeof(uint32_t), cudaHostAllocMapped));
    cudaEvent_t ev[3];
    for (auto& e : ev) cudaEventCreate(&e);
    cudaEventRecord(ev[0], s);

    TrainState S{};
    S.first_new_id = first_new_id; S.vocab_size = vocab_size; S.status = ctx->d_status;
    // heavy steps shared with the other GPUs of the box, if the caller has connected them (bpe_peer_connect) and asked for it
    const bool sharded = ctx->peer_world > 1 && ctx->train_sharded;
    S.rank = sharded ? (uint32_t)ctx->peer_rank : 0u;
    S.world = sharded ? (uint32_t)ctx->peer_world : 1u;
    if (sharded) {
        for (int p = 0; p < ctx->peer_world; p++) S.inbox[p] = (unsigned long long*)ctx->peer_ptr[p];
        S.inbox_cap = ctx->peer_cap;
        S.ut_key = (const unsigned long long*)ctx->ut.key;
        S.ut_mask = ctx->ut.cap_mask;
        if (n_unique / ctx->peer_world * 2 + 4096 > ctx->peer_cap) return ctx_fail(ctx, BPE_ERR_CAPACITY, "peer exchange buffers too small for this table (bpe_peer_alloc)");
    }
    const size_t ph_cap = next_pow2(std::max<size_t>(1u << 16, 2 * n_sym));
    S.ph_mask = ph_cap - 1;
    const uint32_t init_bytes = h_init_off[first_new_id];
    S.blob_cap = (uint32_t)std::min<uint64_t>(0xFFFFFFF0ull, init_bytes + (uint64_t)vocab_size * blob_per_token);
    S.wl_cap = (uint32_t)n_sym + 16;
    // one allocation, carved up (all sizes rounded to 256 B)
    struct Part { void** p; size_t bytes; };
    Part parts[] = {
        {(void**)&S.w_off, n_unique * 4}, {(void**)&S.w_len, n_unique * 4}, {(void**)&S.w_cnt, n_unique * 8},
        {(void**)&S.sym, (n_sym + 4 * n_unique + 8) * 4},
        {(void**)&S.ph_key, ph_cap * 8}, {(void**)&S.ph_cnt, ph_cap * 8}, {(void**)&S.live, ph_cap * 4}, {(void**)&S.hot, (ph_cap + 1) * 16},
        {(void**)&S.tk_off, (vocab_size + 2) * 4ull}, {(void**)&S.tk_blob, S.blob_cap}, {(void**)&S.tk_pre, (vocab_size + 1) * 8ull},
        {(void**)&S.wl_off, (vocab_size + 1) * 4ull}, {(void**)&S.wl_len, (vocab_size + 1) * 4ull}, {(void**)&S.wl_pool, (size_t)S.wl_cap * 8},
        {(void**)&S.m_a, (vocab_size + 1) * 4ull}, {(void**)&S.m_b, (vocab_size + 1) * 4ull},
        {(void**)&S.ctrl, sizeof(Ctrl)}, {(void**)&S.trace, (vocab_size + 1) * 16ull},
        {(void**)&S.w_owner, sharded ? n_unique : 0}, {(void**)&S.w_key, sharded ? n_unique * 8 : 0},
        {(void**)&S.slot_word, sharded ? (ctx->ut.cap_mask + 1) * 4 : 0},
    };
    size_t total = 0;
    for (auto& p : parts) total += (p.bytes + 255) / 256 * 256;
    uint8_t* base = (uint8_t*)ctx_scratch(ctx, 1, total);
    if (!base) return BPE_ERR_CUDA;
    size_t at = 0;
    for (auto& p : parts) { *p.p = base + at; at += (p.bytes + 255) / 256 * 256; }
    if (const char* f = getenv("BPE_B200_TRAIN_FLAGS")) S.flags = (uint32_t)atoi(f);
    S.watchdog_s = 120;
    S.cache_fill = 0;  // chosen below, once the number of distinct pairs is known
    S.driver_sm = -1;
    if (const char* f = getenv("BPE_B200_TRAIN_DRIVER_SM")) S.driver_sm = atoi(f);  // only meaningful with one CTA on every SM
    if (const char* f = getenv("BPE_B200_TRAIN_WATCHDOG")) S.watchdog_s = (uint32_t)std::max(1, atoi(f));
    const char* trace_path = getenv("BPE_B200_TRAIN_TRACE");
    uint32_t* d_trace = S.trace;
    if (!trace_path) S.trace = nullptr;
    // zero the pair table (keys, counts), the posting offsets / lengths and the control block
    CUDA_TRY(ctx, cudaMemsetAsync(S.ph_key, 0, (uint8_t*)S.live - (uint8_t*)S.ph_key, s));
    CUDA_TRY(ctx, cudaMemsetAsync(S.wl_off, 0, (uint8_t*)S.wl_pool - (uint8_t*)S.wl_off, s));
    CUDA_TRY(ctx, cudaMemsetAsync(S.ctrl, 0, sizeof(Ctrl), s));

    // initial tokens: bytes, offsets, prefix keys
    std::vector<unsigned long long> pre(first_new_id, 0);
    for (uint32_t t = 0; t < first_new_id; t++)
        for (uint32_t i = 0; i < 8; i++) pre[t] = (pre[t] << 8) | (i < h_init_off[t + 1] - h_init_off[t] ? h_init_blob[h_init_off[t] + i] : 0u);
    CUDA_TRY(ctx, cudaMemcpyAsync(S.tk_blob, h_init_blob, init_bytes, cudaMemcpyHostToDevice, s));
    CUDA_TRY(ctx, cudaMemcpyAsync(S.tk_off, h_init_off, (first_new_id + 1) * 4ull, cudaMemcpyHostToDevice, s));
    CUDA_TRY(ctx, cudaMemcpyAsync(S.tk_pre, pre.data(), first_new_id * 8ull, cudaMemcpyHostToDevice, s));

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
    // Size of the hot-pair cache.  Fewer cached pairs make every step cheaper (argmax scan, reload behind a heavy step) and
    // rebuilds more frequent, and a rebuild walks all live pairs.  Measured (tools/cache_fill_sweep.py), 512 / 1024 / 2048 / 3072
    // entries: 2.1 GB TinyStories-shaped, vocab 10k (2*10^5 words): 85.4 / 83.3 / 86.4 / 91.8 ms with 20 / 12 / 8 / 8 rebuilds;
    // 11 GB OpenWebText-shaped, vocab 32k (1.9*10^7 words): 4.11 / 3.14 / 2.63 / 2.55 s with 65 / 35 / 19 / 16 rebuilds.
    // (The table of pairs grows with the words, not with the byte pairs it starts from: the words decide.)
    S.cache_fill = n_words < (1u << 20) ? 1024u : n_words < (1u << 23) ? 2048u : 3072u;
    if (const char* f = getenv("BPE_B200_TRAIN_CACHE_FILL")) S.cache_fill = (uint32_t)std::min(3584, std::max(64, atoi(f)));
    if (getenv("BPE_B200_TRAIN_VERBOSE")) fprintf(stderr, "[bpe_train] %u words, %u distinct pairs at the start, hot-pair cache aims for %u entries\n", n_words, h_ctrl.n_live, S.cache_fill);
    if (h_ctrl.n_live == 0) {
        for (auto& e : ev) cudaEventDestroy(e);
        return ctx_fail(ctx, BPE_ERR_NO_PAIRS, "no byte pairs to merge");
    }
    if (vocab_size == first_new_id) {
        for (auto& e : ev) cudaEventDestroy(e);
        return BPE_OK;
    }

    // Progress words in pinned host memory: if the loop ever stalls the host says where instead of hanging silently.
    uint32_t* h_beat = ctx->h_beat;  // per context: two engines training from two threads must not share it
    memset(h_beat, 0, 2 * 4096 * sizeof(uint32_t));
    uint32_t* d_beat = nullptr;
    CUDA_TRY(ctx, cudaHostGetDevicePointer((void**)&d_beat, h_beat, 0));
    S.beat = d_beat;
    int launched_ctas = 0;
    // Default: a cooperative grid, one CTA per SM, helpers polling a flag (measured faster than the one-cluster variant:
    // heavy steps get all SMs).  BPE_B200_TRAIN_MODE=cluster selects the single-cluster variant (experiments, tests).
    bool launched = false;
    const char* mode = getenv("BPE_B200_TRAIN_MODE");
    if (mode && !strcmp(mode, "cluster") && !sharded) {
        auto kern = merge_loop_kernel<true, 1024>;
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(DriverShared));
        for (int csize = 16; csize >= 8 && !launched && e == cudaSuccess; csize -= 8) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(csize); cfg.blockDim = dim3(1024); cfg.dynamicSmemBytes = sizeof(DriverShared); cfg.stream = s;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = csize; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr; cfg.numAttrs = 1;
            int n_clusters = 0;
            if (cudaOccupancyMaxActiveClusters(&n_clusters, kern, &cfg) != cudaSuccess || n_clusters < 1) { cudaGetLastError(); continue; }
            KernelTimer t(ctx, "merge_loop_kernel", s);
            if (cudaLaunchKernelEx(&cfg, kern, S) == cudaSuccess) { launched = true; launched_ctas = csize; } else cudaGetLastError();
        }
    }
    if (!launched) {
        int nt = 512;
        if (const char* env = getenv("BPE_B200_TRAIN_NT")) nt = atoi(env);
        void (*kern)(TrainState) = nt == 1024 ? merge_loop_kernel<false, 1024> : nt == 256 ? merge_loop_kernel<false, 256> : merge_loop_kernel<false, 512>;
        if (nt != 1024 && nt != 256) nt = 512;
        if (sharded) { kern = merge_loop_kernel<false, 512, true>; nt = 512; }
        // Between the pair histogram and this launch the GPU is idle: nothing slow may sit here.  cudaGetDeviceProperties did
        // (2.5 ms on a good call, up to 100 ms on a bad one: the run-to-run outliers of round 1); the SM count is the
        // context's, the attribute and the occupancy answer are kept per kernel variant.
        if (ctx->train_kern != (void*)kern) {
            int per_sm = 0;
            CUDA_TRY(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(DriverShared)));
            CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, nt, sizeof(DriverShared)));
            ctx->train_kern = (void*)kern;
            ctx->train_per_sm = per_sm;
        }
        int grid = ctx->train_per_sm >= 1 ? ctx->n_sm : 1;  // helpers poll a flag: every CTA must be resident (cooperative launch)
        if (const char* env = getenv("BPE_B200_TRAIN_CTAS")) {
            int v = atoi(env);
            if (v >= 1 && v <= grid) grid = v;
        }
        KernelTimer t(ctx, "merge_loop_kernel", s);
        void* args[] = {&S};
        CUDA_TRY(ctx, cudaLaunchCooperativeKernel((void*)kern, dim3(grid), dim3(nt), args, sizeof(DriverShared), s));
        launched_ctas = grid;
    }
    double host_wait = 0, host_max_gap = 0;
    unsigned host_polls = 0;
    {   // wait with a deadline
        cudaEvent_t done;
        cudaEventCreateWithFlags(&done, cudaEventDisableTiming);
        cudaEventRecord(done, s);
        auto t0 = std::chrono::steady_clock::now();
        double last_poll = 0;
        while (cudaEventQuery(done) == cudaErrorNotReady) {
            double waited = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            host_max_gap = std::max(host_max_gap, waited - last_poll);
            last_poll = waited;
            host_polls++;
            if (waited > S.watchdog_s + 10.0) {
                fprintf(stderr, "[bpe_train rank %s] merge loop stalled: driver at iteration %u phase %u;", getenv("RANK") ? getenv("RANK") : "-", h_beat[0], h_beat[1]);
                for (int c = 1; c < launched_ctas && c < 16; c++) fprintf(stderr, " cta%d round %u phase %u;", c, h_beat[2 * c], h_beat[2 * c + 1]);
                fprintf(stderr, "\n");
                if (S.flags & 16u) {
                    fprintf(stderr, "[bpe_train] driver sees done %u of %u (hand-overs %u); helper CTAs not waiting for hand-over %u:", h_beat[8000], h_beat[8001], h_beat[8002], h_beat[8002]);
                    for (int c = 1; c < launched_ctas; c++)
                        if (h_beat[64 + 2 * c] != h_beat[8002] || h_beat[64 + 2 * c + 1] != 20) fprintf(stderr, " cta%d round %u phase %u;", c, h_beat[64 + 2 * c], h_beat[64 + 2 * c + 1]);
                    fprintf(stderr, "\n");
                }
                cudaEventDestroy(done);
                return ctx_fail(ctx, BPE_ERR_CUDA, "merge loop stalled (internal error); the process must exit");
            }
            if (waited > 0.002) std::this_thread::sleep_for(std::chrono::microseconds(200));
        }
        host_wait = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        cudaEventDestroy(done);
    }
    CUDA_TRY(ctx, cudaMemcpyAsync(&h_ctrl, S.ctrl, sizeof(Ctrl), cudaMemcpyDeviceToHost, s));
    cudaEventRecord(ev[2], s);
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    const uint32_t learned = h_ctrl.n_merges;
    CUDA_TRY(ctx, cudaMemcpyAsync(h_merge_a, S.m_a, learned * 4ull, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(ctx, cudaMemcpyAsync(h_merge_b, S.m_b, learned * 4ull, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    *n_merges = learned;
    if (h_ctrl.dbg[5])
        fprintf(stderr, "[train watchdog] stopped at iteration %u after %u loop passes; merges %u, rebuilds %u, heavy %u\n", h_ctrl.dbg[4], h_ctrl.dbg[5],
                h_ctrl.n_merges, h_ctrl.n_rebuilds, h_ctrl.n_heavy);
    if (S.flags & 4u) fprintf(stderr, "[train check] token comparisons %u over %u iterations\n", h_ctrl.dbg[6], h_ctrl.n_iters);
    if (S.flags & 2u)
        fprintf(stderr, "[train check] stale cache copies %u (first at iteration %u), hot pairs missing from the cache %u (first at iteration %u), rebuilds %u, heavy %u, iterations %u\n",
                h_ctrl.dbg[0], h_ctrl.dbg[1], h_ctrl.dbg[2], h_ctrl.dbg[3], h_ctrl.n_rebuilds, h_ctrl.n_heavy, h_ctrl.n_iters);
    if (trace_path) {  // developer aid: per-iteration timings of the loop
        const double nl = (double)std::max<unsigned long long>(1, h_ctrl.phase_clk[6]), nh = (double)std::max<unsigned long long>(1, h_ctrl.phase_clk[7]);
        fprintf(stderr, "[train trace] driver thread 0, SM cycles per LIGHT step (%llu steps): argmax %.0f, select+bookkeeping %.0f, filter %.0f, rewrite %.0f, "
                        "close %.0f; per HEAVY step (%llu steps): %.0f; rebuilds %u\n", h_ctrl.phase_clk[6], h_ctrl.phase_clk[0] / nl, h_ctrl.phase_clk[1] / nl,
                h_ctrl.phase_clk[2] / nl, h_ctrl.phase_clk[3] / nl, h_ctrl.phase_clk[4] / nl, h_ctrl.phase_clk[7], h_ctrl.phase_clk[5] / nh, h_ctrl.n_rebuilds);
        fprintf(stderr, "[train trace] HEAVY step, cycles: argmax+select %.0f, drain %.0f, wake %.0f, own share %.0f, join %.0f, reload+pending %.0f\n",
                h_ctrl.heavy_clk[5] / nh, h_ctrl.heavy_clk[0] / nh, h_ctrl.heavy_clk[1] / nh, h_ctrl.heavy_clk[2] / nh, h_ctrl.heavy_clk[3] / nh, h_ctrl.heavy_clk[4] / nh);
        const double ni = (double)std::max(1u, h_ctrl.n_iters);
        fprintf(stderr, "[train trace] argmax, cycles per step: scan %.0f, reduce+barrier %.0f, ties %.0f, barrier %.0f, order %.0f (candidates %.2f); entries scanned %.0f\n",
                h_ctrl.sub_clk[0] / ni, h_ctrl.sub_clk[1] / ni, h_ctrl.sub_clk[2] / ni, h_ctrl.sub_clk[3] / ni, h_ctrl.sub_clk[4] / ni,
                h_ctrl.sub_clk[6] / ni, h_ctrl.sub_clk[5] / ni);
        std::vector<uint32_t> tr(4ull * h_ctrl.n_iters);
        cudaMemcpy(tr.data(), d_trace, tr.size() * 4, cudaMemcpyDeviceToHost);
        if (FILE* f = fopen(trace_path, "w")) {
            for (uint32_t k = 0; k < h_ctrl.n_iters && k < vocab_size; k++) fprintf(f, "%u %u %u %u\n", tr[4 * k], tr[4 * k + 1], tr[4 * k + 2], tr[4 * k + 3]);
            fclose(f);
        }
    }
    float ms0 = 0, ms1 = 0;
    cudaEventElapsedTime(&ms0, ev[0], ev[1]);
    cudaEventElapsedTime(&ms1, ev[1], ev[2]);
    ctx->train_ms[0] = ms0; ctx->train_ms[1] = ms1; ctx->train_ms[2] = h_ctrl.n_iters;
    if (getenv("BPE_B200_TRAIN_VERBOSE"))
        fprintf(stderr, "[bpe_train] merge loop %.2f ms, %u iterations, %u heavy steps, %u rebuilds, driver CTA on SM %u; driver CTA inside the kernel %.2f ms; "
                        "host waited %.2f ms in %u polls, longest gap between polls %.2f ms\n", ms1, h_ctrl.n_iters,
                h_ctrl.n_heavy, h_ctrl.n_rebuilds, h_ctrl.dbg[7], (h_ctrl.t_end - h_ctrl.t_begin) / 1e6, host_wait * 1e3, host_polls, host_max_gap * 1e3);
    for (auto& e : ev) cudaEventDestroy(e);
    const uint32_t st = ctx_read_status(ctx, s);
    if (st == ST_BLOB_FULL) { *blob_full = true; return BPE_OK; }  // nothing else went wrong: the caller retries with a larger pool
    return ctx_status_to_code(ctx, st);
}
