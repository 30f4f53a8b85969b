// api.cu — the extern "C" surface declared in include/bpe_b200.h: context life cycle, argument checks, status
// translation.  The kernels live in scan.cu, count.cu, train.cu and encode.cu.
#include <algorithm>
#include <cstring>

#include "../../include/bpe_unicode_tables.h"
#include <mutex>

#include "common.cuh"

static std::string g_create_error;

// Closes the peers' exchange buffers (bpe_peer_connect); this GPU's own buffer stays.
static void peer_release(bpe_ctx* ctx) {
    for (int p = 0; p < ctx->peer_world; p++)
        if (ctx->peer_ptr[p] && ctx->peer_ptr[p] != ctx->peer_buf) cudaIpcCloseMemHandle(ctx->peer_ptr[p]);
    for (auto& p : ctx->peer_ptr) p = nullptr;
    ctx->peer_world = 1; ctx->peer_rank = 0; ctx->train_sharded = false;
}

int ctx_fail(bpe_ctx* ctx, int code, const char* what, cudaError_t e) {
    static std::mutex mu;  // bpe_load_file / bpe_store_to_host may run beside another call on the same context
    std::lock_guard<std::mutex> lock(mu);
    std::string msg = what ? what : "";
    if (e != cudaSuccess) { msg += ": "; msg += cudaGetErrorString(e); }
    if (ctx) ctx->err = msg; else g_create_error = msg;
    return code;
}

void* ctx_scratch(bpe_ctx* ctx, int which, size_t bytes) {
    if (ctx->scratch_bytes[which] >= bytes) return ctx->scratch[which];
    if (ctx->scratch[which]) cudaFree(ctx->scratch[which]);
    ctx->scratch[which] = nullptr; ctx->scratch_bytes[which] = 0;
    size_t want = bytes + bytes / 4 + 4096;
    cudaError_t e = cudaMalloc(&ctx->scratch[which], want);
    if (e != cudaSuccess) { ctx_fail(ctx, BPE_ERR_CUDA, "cudaMalloc(scratch)", e); return nullptr; }
    ctx->scratch_bytes[which] = want;
    return ctx->scratch[which];
}

// A few words from device memory to the host WITHOUT the copy engine: one tiny kernel stores them into page-locked,
// device-mapped host memory.  A 4-byte cudaMemcpyAsync would queue on the device-to-host copy engine behind whatever
// bulk copy is in flight there (bpe_store_to_host of the previous piece's ids: more than a millisecond), and every
// call of the encode pipeline reads a status word and a total.
__global__ void read_words_kernel(const uint32_t* __restrict__ src, int n_words, const uint32_t* __restrict__ status, volatile uint32_t* host) {
    const int t = threadIdx.x;
    if (t < n_words) host[1 + t] = src[t];
    if (t == 0) host[0] = status ? *status : 0;
    __threadfence_system();
}

// status word (0x80000000 on a CUDA error) + n_words 32-bit words of d_src (may be NULL / 0) into h_dst; synchronises s
uint32_t ctx_read_words(bpe_ctx* ctx, cudaStream_t s, const void* d_src, int n_words, void* h_dst) {
    if (n_words > 30) return 0x80000000u;
    if (!ctx->h_read) {
        if (cudaHostAlloc((void**)&ctx->h_read, 32 * sizeof(uint32_t), cudaHostAllocMapped) != cudaSuccess) return 0x80000000u;
        if (cudaHostGetDevicePointer((void**)&ctx->d_read, (void*)ctx->h_read, 0) != cudaSuccess) return 0x80000000u;
    }
    read_words_kernel<<<1, 32, 0, s>>>((const uint32_t*)d_src, n_words, ctx->d_status, ctx->d_read);
    if (cudaGetLastError() != cudaSuccess || cudaStreamSynchronize(s) != cudaSuccess) return 0x80000000u;
    if (n_words) memcpy(h_dst, (const void*)(ctx->h_read + 1), (size_t)n_words * 4);
    return ctx->h_read[0];
}

uint32_t ctx_read_status(bpe_ctx* ctx, cudaStream_t s, bool clear) {
    uint32_t st = ctx_read_words(ctx, s, nullptr, 0, nullptr);
    if (st && st != 0x80000000u && clear) cudaMemsetAsync(ctx->d_status, 0, 4, s);
    return st;
}

int ctx_status_to_code(bpe_ctx* ctx, uint32_t st) {
    if (!st) return BPE_OK;
    if (st & 0x80000000u) return ctx_fail(ctx, BPE_ERR_CUDA, "reading the status word", cudaGetLastError());
    if (st & ST_INVALID_UTF8) return ctx_fail(ctx, BPE_ERR_INVALID_UTF8, "text is not valid UTF-8");
    if (st & ST_COLLISION) return ctx_fail(ctx, BPE_ERR_HASH_COLLISION, "two pretokens share a 64-bit key");
    if (st & ST_MISSING_ID) return ctx_fail(ctx, BPE_ERR_MISSING_ID, "bytes without an id in the vocabulary");
    if (st & ST_TABLE_FULL) return ctx_fail(ctx, BPE_ERR_CAPACITY, "pretoken table full (bpe_ctx_configure: table_log2)");
    if (st & ST_POOL_FULL) return ctx_fail(ctx, BPE_ERR_CAPACITY, "pretoken byte / id pool full (bpe_ctx_configure: pool_bytes)");
    if (st & ST_PAIR_TABLE_FULL) return ctx_fail(ctx, BPE_ERR_CAPACITY, "pair table full");
    if (st & ST_WATCHDOG) return ctx_fail(ctx, BPE_ERR_CUDA, "merge loop watchdog: no progress for 60 s (internal error)");
    if (st & ST_LIST_FULL) return ctx_fail(ctx, BPE_ERR_CAPACITY, "posting pool full");
    if (st & ST_BLOB_FULL) return ctx_fail(ctx, BPE_ERR_CAPACITY, "token byte pool full");
    if (st & ST_HAS_CR) return ctx_fail(ctx, BPE_ERR_CARRIAGE_RETURN, "the text holds a carriage return (BPE_FAIL_ON_CR)");
    return ctx_fail(ctx, BPE_ERR_CUDA, "unknown status bits");
}

static void free_unique(bpe_ctx* ctx) {
    if (!ctx->ut_alloc) return;
    cudaFree(ctx->ut.key); cudaFree(ctx->ut.cnt); cudaFree(ctx->ut.pool_off); cudaFree(ctx->ut.len); cudaFree(ctx->ut.chk);
    cudaFree(ctx->ut.enc_off); cudaFree(ctx->ut.enc_len); cudaFree(ctx->ut.enc); cudaFree(ctx->ut.pool); cudaFree(ctx->ut.counters);
    ctx->ut = UniqueTable{};
    ctx->ut_alloc = false;
}

static int clear_unique(bpe_ctx* ctx) {
    size_t cap = ctx->ut.cap_mask + 1;
    CUDA_TRY(ctx, cudaMemset(ctx->ut.key, 0, cap * 8));
    CUDA_TRY(ctx, cudaMemset(ctx->ut.cnt, 0, cap * 8));
    CUDA_TRY(ctx, cudaMemset(ctx->ut.chk, 0, cap * 4));
    CUDA_TRY(ctx, cudaMemset(ctx->ut.enc_len, 0xFF, cap * 4));
    CUDA_TRY(ctx, cudaMemset(ctx->ut.enc, 0, cap * 16));
    CUDA_TRY(ctx, cudaMemset(ctx->ut.counters, 0, 4 * 8));
    ctx->enc_done = 0;
    ctx->enc_warm = false;
    return BPE_OK;
}

static int ensure_unique(bpe_ctx* ctx) {
    if (ctx->ut_alloc) return BPE_OK;
    size_t cap = (size_t)1 << ctx->table_log2;
    UniqueTable& t = ctx->ut;
    CUDA_TRY(ctx, cudaMalloc(&t.key, cap * 8));
    CUDA_TRY(ctx, cudaMalloc(&t.cnt, cap * 8));
    CUDA_TRY(ctx, cudaMalloc(&t.pool_off, cap * 8));
    CUDA_TRY(ctx, cudaMalloc(&t.len, cap * 4));
    CUDA_TRY(ctx, cudaMalloc(&t.chk, cap * 4));
    CUDA_TRY(ctx, cudaMalloc(&t.enc_off, cap * 4));
    CUDA_TRY(ctx, cudaMalloc(&t.enc_len, cap * 4));
    CUDA_TRY(ctx, cudaMalloc(&t.enc, cap * 16));
    CUDA_TRY(ctx, cudaMalloc(&t.pool, ctx->pool_bytes + BPE_TEXT_PAD));
    CUDA_TRY(ctx, cudaMalloc(&t.counters, 4 * 8));
    t.cap_mask = cap - 1;
    t.pool_cap = ctx->pool_bytes;
    ctx->ut_alloc = true;
    return clear_unique(ctx);
}

extern "C" {

int bpe_ctx_create(int device, bpe_ctx** out) {
    if (!out) return BPE_ERR_ARGUMENT;
    *out = nullptr;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return ctx_fail(nullptr, BPE_ERR_CUDA, "cudaSetDevice", e);
    bpe_ctx* ctx = new bpe_ctx();
    ctx->device = device;
    cudaDeviceGetAttribute(&ctx->n_sm, cudaDevAttrMultiProcessorCount, device);
    if (ctx->n_sm <= 0) ctx->n_sm = BPE_N_SM;
    auto fail = [&](const char* what, cudaError_t err) { ctx_fail(nullptr, BPE_ERR_CUDA, what, err); bpe_ctx_destroy(ctx); return BPE_ERR_CUDA; };
    if ((e = cudaMalloc(&ctx->d_status, 4)) != cudaSuccess) return fail("cudaMalloc(status)", e);
    if ((e = cudaMemset(ctx->d_status, 0, 4)) != cudaSuccess) return fail("cudaMemset(status)", e);
    if ((e = cudaMalloc(&ctx->d_stage1, sizeof(BPE_UC_STAGE1))) != cudaSuccess) return fail("cudaMalloc(stage1)", e);
    if ((e = cudaMalloc(&ctx->d_stage2, sizeof(BPE_UC_STAGE2))) != cudaSuccess) return fail("cudaMalloc(stage2)", e);
    if ((e = cudaMemcpy(ctx->d_stage1, BPE_UC_STAGE1, sizeof(BPE_UC_STAGE1), cudaMemcpyHostToDevice)) != cudaSuccess) return fail("upload stage1", e);
    if ((e = cudaMemcpy(ctx->d_stage2, BPE_UC_STAGE2, sizeof(BPE_UC_STAGE2), cudaMemcpyHostToDevice)) != cudaSuccess) return fail("upload stage2", e);
    ctx->cls.stage1 = ctx->d_stage1; ctx->cls.stage2 = ctx->d_stage2; ctx->cls.n_stage1 = BPE_UC_N_STAGE1;
    memset(&ctx->sp, 0, sizeof(ctx->sp));
    *out = ctx;
    return BPE_OK;
}

void bpe_ctx_destroy(bpe_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    free_unique(ctx);
    io_staging_free(ctx);
    cudaFree(ctx->d_status); cudaFree(ctx->d_stage1); cudaFree(ctx->d_stage2);
    cudaFree(ctx->d_mkey); cudaFree(ctx->d_mrank); cudaFree(ctx->d_mid); cudaFree(ctx->d_byte_to_id); cudaFree(ctx->d_special_id);
    cudaFree(ctx->d_idpool);
    for (auto& p : ctx->scratch) cudaFree(p);
    if (ctx->h_beat) cudaFreeHost(ctx->h_beat);
    if (ctx->h_read) cudaFreeHost((void*)ctx->h_read);
    peer_release(ctx);
    cudaFree(ctx->peer_buf);
    delete ctx;
}

const char* bpe_last_error(const bpe_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int bpe_ctx_configure(bpe_ctx* ctx, int table_log2, uint64_t pool_bytes) {
    if (!ctx || table_log2 < 10 || table_log2 > 30 || pool_bytes < 4096) return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_ctx_configure") : BPE_ERR_ARGUMENT;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    free_unique(ctx);
    ctx->table_log2 = table_log2;
    ctx->pool_bytes = pool_bytes;
    if (ctx->d_idpool) { cudaFree(ctx->d_idpool); ctx->d_idpool = nullptr; ctx->idpool_cap = 0; }
    return BPE_OK;
}

int bpe_set_special_tokens(bpe_ctx* ctx, const uint8_t* h_blob, const uint32_t* h_off, int n) {
    if (!ctx || n < 0 || n > BPE_MAX_SPECIALS) return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "at most 16 special tokens") : BPE_ERR_ARGUMENT;
    Specials sp;
    memset(&sp, 0, sizeof(sp));
    sp.n = n;
    for (int k = 0; k < n; k++) {
        uint32_t len = h_off[k + 1] - h_off[k];
        // scan.cu searches SCAN_HALO = 2 chunks (64 bytes) in front of a tile for tokens that reach into it
        if (len == 0 || len > 60 || h_off[k + 1] - h_off[0] > BPE_MAX_SPECIAL_BYTES)
            return ctx_fail(ctx, BPE_ERR_ARGUMENT, "special tokens must be 1..60 bytes, 256 bytes in total");
        sp.off[k] = h_off[k] - h_off[0];
        sp.off[k + 1] = h_off[k + 1] - h_off[0];
        memcpy(sp.blob + sp.off[k], h_blob + h_off[k], len);
        memcpy(sp.words[k], h_blob + h_off[k], len);  // little-endian words, zero padded (memset above)
        if (len > 1 && memchr(h_blob + h_off[k], '\n', len - 1)) sp.newline_inside |= 1u << k;
        uint8_t b0 = h_blob[h_off[k]];
        sp.first_byte_mask[b0 >> 5] |= 1u << (b0 & 31);
        sp.max_len = std::max<int>(sp.max_len, (int)len);
    }
    ctx->sp = sp;
    return BPE_OK;
}

int bpe_scan(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint64_t* d_cuts, uint64_t n_cuts, uint32_t flags,
             uint32_t* d_start, void* stream) {
    if (!ctx || !d_start || (n && !d_text) || ((uintptr_t)d_text & 15)) return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_scan arguments") : BPE_ERR_ARGUMENT;
    cudaStream_t s = (cudaStream_t)stream;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    uint32_t* cutmask = nullptr;
    if (d_cuts && n_cuts) {
        cutmask = (uint32_t*)ctx_scratch(ctx, 5, (n / 32 + 2) * 4);
        if (!cutmask) return BPE_ERR_CUDA;
        int rc = launch_cuts_to_mask(ctx, d_cuts, n_cuts, n, cutmask, s);
        if (rc) return rc;
    }
    int rc = launch_scan(ctx, d_text, n, cutmask, flags, d_start, s);
    if (rc) return rc;
    return ctx_status_to_code(ctx, ctx_read_status(ctx, s));
}

int bpe_count(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint32_t* d_start, void* stream) {
    if (!ctx || (n && (!d_text || !d_start)) || ((uintptr_t)d_text & 15)) return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_count arguments") : BPE_ERR_ARGUMENT;
    cudaStream_t s = (cudaStream_t)stream;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    int rc = ensure_unique(ctx);
    if (rc) return rc;
    rc = launch_count(ctx, d_text, n, d_start, s);
    if (rc) return rc;
    return ctx_status_to_code(ctx, ctx_read_status(ctx, s));
}

int bpe_pretokenize_count(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint64_t* d_cuts, uint64_t n_cuts, uint32_t flags,
                          uint32_t* d_start, void* stream) {
    if (!ctx || (n && !d_text) || ((uintptr_t)d_text & 15)) return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_pretokenize_count arguments") : BPE_ERR_ARGUMENT;
    cudaStream_t s = (cudaStream_t)stream;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    int rc = ensure_unique(ctx);
    if (rc) return rc;
    uint32_t* cutmask = nullptr;
    if (d_cuts && n_cuts) {
        cutmask = (uint32_t*)ctx_scratch(ctx, 5, (n / 32 + 2) * 4);
        if (!cutmask) return BPE_ERR_CUDA;
        rc = launch_cuts_to_mask(ctx, d_cuts, n_cuts, n, cutmask, s);
        if (rc) return rc;
    }
    rc = launch_fused_count(ctx, d_text, n, cutmask, flags, d_start, s);
    if (rc) return rc;
    return ctx_status_to_code(ctx, ctx_read_status(ctx, s));
}

int bpe_unique_stats(bpe_ctx* ctx, uint64_t* n_unique, uint64_t* n_bytes, uint64_t* n_tokens) {
    if (!ctx) return BPE_ERR_ARGUMENT;
    uint64_t c[4] = {0, 0, 0, 0};
    if (ctx->ut_alloc) {
        CUDA_TRY(ctx, cudaSetDevice(ctx->device));
        CUDA_TRY(ctx, cudaMemcpy(c, ctx->ut.counters, sizeof(c), cudaMemcpyDeviceToHost));
    }
    if (n_bytes) *n_bytes = c[0];
    if (n_unique) *n_unique = c[1];
    if (n_tokens) *n_tokens = c[2];
    return BPE_OK;
}

int bpe_unique_export(bpe_ctx* ctx, uint8_t* d_blob, uint64_t* d_off, uint32_t* d_len, uint64_t* d_cnt, void* stream) {
    if (!ctx) return BPE_ERR_ARGUMENT;
    if (!ctx->ut_alloc) return BPE_OK;
    cudaStream_t s = (cudaStream_t)stream;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    int rc = launch_export(ctx, d_blob, d_off, d_len, d_cnt, s);
    if (rc) return rc;
    CUDA_TRY(ctx, cudaStreamSynchronize(s));
    return BPE_OK;
}

int bpe_unique_import(bpe_ctx* ctx, const uint8_t* d_blob, const uint64_t* d_off, const uint32_t* d_len, const uint64_t* d_cnt,
                      uint64_t n_unique, void* stream) {
    if (!ctx || ((uintptr_t)d_blob & 15)) return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_unique_import arguments") : BPE_ERR_ARGUMENT;
    cudaStream_t s = (cudaStream_t)stream;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    int rc = ensure_unique(ctx);
    if (rc) return rc;
    rc = launch_import(ctx, d_blob, d_off, d_len, d_cnt, n_unique, s);
    if (rc) return rc;
    return ctx_status_to_code(ctx, ctx_read_status(ctx, s));
}

int bpe_unique_reset(bpe_ctx* ctx) {
    if (!ctx) return BPE_ERR_ARGUMENT;
    if (!ctx->ut_alloc) return BPE_OK;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaDeviceSynchronize());
    return clear_unique(ctx);
}

int bpe_train(bpe_ctx* ctx, uint32_t vocab_size, uint32_t first_new_id, const uint8_t* h_init_blob, const uint32_t* h_init_off,
              uint32_t* h_merge_a, uint32_t* h_merge_b, uint32_t* n_merges, void* stream) {
    if (!ctx || !h_init_blob || !h_init_off || !n_merges || first_new_id < 256) return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_train arguments") : BPE_ERR_ARGUMENT;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    return run_train(ctx, vocab_size, first_new_id, h_init_blob, h_init_off, h_merge_a, h_merge_b, n_merges, (cudaStream_t)stream);
}

int bpe_peer_alloc(bpe_ctx* ctx, uint32_t hits_per_sender, uint8_t h_handle[64]) {
    if (!ctx || !h_handle || hits_per_sender == 0) return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_peer_alloc arguments") : BPE_ERR_ARGUMENT;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    peer_release(ctx);
    if (ctx->peer_buf) { cudaFree(ctx->peer_buf); ctx->peer_buf = nullptr; }
    const size_t bytes = (2 * 8 + 2 * 8 * (size_t)hits_per_sender) * 8;  // flags + hit areas: 2 parities x 8 senders (train.cu: peer_flag / peer_entry)
    CUDA_TRY(ctx, cudaMalloc(&ctx->peer_buf, bytes));
    CUDA_TRY(ctx, cudaMemset(ctx->peer_buf, 0, bytes));
    ctx->peer_cap = hits_per_sender;
    cudaIpcMemHandle_t h;
    CUDA_TRY(ctx, cudaIpcGetMemHandle(&h, ctx->peer_buf));
    memcpy(h_handle, &h, 64);
    return BPE_OK;
}

int bpe_peer_connect(bpe_ctx* ctx, int rank, int world, const uint8_t* h_handles) {
    if (!ctx || !h_handles || world < 2 || world > 8 || rank < 0 || rank >= world || !ctx->peer_buf)
        return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_peer_connect arguments (2..8 GPUs, after bpe_peer_alloc)") : BPE_ERR_ARGUMENT;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    peer_release(ctx);
    for (int p = 0; p < world; p++) {
        if (p == rank) { ctx->peer_ptr[p] = ctx->peer_buf; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, h_handles + 64 * p, 64);
        cudaError_t e = cudaIpcOpenMemHandle(&ctx->peer_ptr[p], h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            ctx->peer_world = p;  // what has been opened so far
            peer_release(ctx);
            return ctx_fail(ctx, BPE_ERR_CUDA, "cudaIpcOpenMemHandle (peer exchange buffer)", e);
        }
    }
    ctx->peer_rank = rank; ctx->peer_world = world; ctx->train_sharded = true;
    return BPE_OK;
}

int bpe_peer_disconnect(bpe_ctx* ctx) {
    if (!ctx) return BPE_ERR_ARGUMENT;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaDeviceSynchronize());
    peer_release(ctx);
    if (ctx->peer_buf) { cudaFree(ctx->peer_buf); ctx->peer_buf = nullptr; }
    ctx->peer_cap = 0;
    return BPE_OK;
}

int bpe_train_timing(bpe_ctx* ctx, double out[3]) {
    if (!ctx || !out) return BPE_ERR_ARGUMENT;
    for (int k = 0; k < 3; k++) out[k] = ctx->train_ms[k];
    return BPE_OK;
}

int bpe_load_model(bpe_ctx* ctx, const uint32_t* h_a, const uint32_t* h_b, const uint32_t* h_merged, uint32_t n_merges,
                   const uint32_t* h_byte_to_id, const uint32_t* h_special_id) {
    if (!ctx || !h_byte_to_id || (n_merges && (!h_a || !h_b || !h_merged))) return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_load_model arguments") : BPE_ERR_ARGUMENT;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaDeviceSynchronize());
    size_t cap = 1024;
    while (cap < (size_t)n_merges * 2 + 2) cap <<= 1;
    std::vector<uint64_t> key(cap, 0);
    std::vector<uint32_t> rank(cap, 0), id(cap, 0);
    for (uint32_t r = 0; r < n_merges; r++) {  // later duplicate wins, like the reference's dict (bpe_tokenizer.py:55)
        uint64_t k = (((uint64_t)h_a[r] << 32) | h_b[r]) + 1ull;
        size_t slot = (size_t)((k * 0x9E3779B97F4A7C15ull >> 20) & (cap - 1));
        while (key[slot] && key[slot] != k) slot = (slot + 1) & (cap - 1);
        key[slot] = k; rank[slot] = r; id[slot] = h_merged[r];
    }
    cudaFree(ctx->d_mkey); cudaFree(ctx->d_mrank); cudaFree(ctx->d_mid);
    ctx->d_mkey = nullptr; ctx->d_mrank = nullptr; ctx->d_mid = nullptr;
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_mkey, cap * 8));
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_mrank, cap * 4));
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_mid, cap * 4));
    CUDA_TRY(ctx, cudaMemcpy(ctx->d_mkey, key.data(), cap * 8, cudaMemcpyHostToDevice));
    CUDA_TRY(ctx, cudaMemcpy(ctx->d_mrank, rank.data(), cap * 4, cudaMemcpyHostToDevice));
    CUDA_TRY(ctx, cudaMemcpy(ctx->d_mid, id.data(), cap * 4, cudaMemcpyHostToDevice));
    ctx->m_mask = cap - 1;
    if (!ctx->d_byte_to_id) CUDA_TRY(ctx, cudaMalloc(&ctx->d_byte_to_id, 256 * 4));
    if (!ctx->d_special_id) CUDA_TRY(ctx, cudaMalloc(&ctx->d_special_id, BPE_MAX_SPECIALS * 4));
    CUDA_TRY(ctx, cudaMemcpy(ctx->d_byte_to_id, h_byte_to_id, 256 * 4, cudaMemcpyHostToDevice));
    uint32_t sid[BPE_MAX_SPECIALS] = {0};
    for (int k = 0; k < ctx->sp.n && h_special_id; k++) sid[k] = h_special_id[k];
    CUDA_TRY(ctx, cudaMemcpy(ctx->d_special_id, sid, sizeof(sid), cudaMemcpyHostToDevice));
    int rc = ensure_unique(ctx);
    if (rc) return rc;
    if (!ctx->d_idpool) {
        ctx->idpool_cap = ctx->pool_bytes / 2 + 1024;
        CUDA_TRY(ctx, cudaMalloc(&ctx->d_idpool, ctx->idpool_cap * 4));
    }
    // cached encodings belong to the previous model
    size_t ucap = ctx->ut.cap_mask + 1;
    CUDA_TRY(ctx, cudaMemset(ctx->ut.enc_len, 0xFF, ucap * 4));
    CUDA_TRY(ctx, cudaMemset(ctx->ut.enc, 0, ucap * 16));
    CUDA_TRY(ctx, cudaMemset(ctx->ut.counters + 3, 0, 8));
    ctx->enc_done = 0;
    ctx->enc_warm = false;
    ctx->model_loaded = true;
    return BPE_OK;
}

int bpe_encode(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint64_t* d_cuts, uint64_t n_cuts, uint32_t flags,
               const uint32_t* d_start, void* d_out, int out_elem_bytes, uint64_t cap, uint64_t* n_out, void* stream) {
    if (!ctx || !n_out || (out_elem_bytes != 2 && out_elem_bytes != 4) || (n && !d_text) || ((uintptr_t)d_text & 15))
        return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_encode arguments") : BPE_ERR_ARGUMENT;
    if (!ctx->model_loaded) return ctx_fail(ctx, BPE_ERR_STATE, "bpe_encode before bpe_load_model");
    cudaStream_t s = (cudaStream_t)stream;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    *n_out = 0;
    if (n == 0) return BPE_OK;
    if (flags & BPE_FAIL_ON_CR) {
        int rc = launch_find_byte(ctx, d_text, n, '\r', ST_HAS_CR, s);
        if (rc) return rc;
        if (d_start) {  // no scan of ours follows whose status read would report it
            rc = ctx_status_to_code(ctx, ctx_read_status(ctx, s));
            if (rc) return rc;
        }
    }
    if (!d_start) {
        uint32_t* mask = (uint32_t*)ctx_scratch(ctx, 6, (n / 32 + 2) * 4);
        if (!mask) return BPE_ERR_CUDA;
        // cold table: split and count in one pass over the text; warm table: the split alone, the emit pass does the rest
        const bool cold = !ctx->enc_warm;
        int rc = cold ? bpe_pretokenize_count(ctx, d_text, n, d_cuts, n_cuts, flags, mask, stream)
                      : bpe_scan(ctx, d_text, n, d_cuts, n_cuts, flags, mask, stream);
        if (rc) return rc;
        return run_encode(ctx, d_text, n, mask, d_out, out_elem_bytes, cap, n_out, s, cold);
    }
    return run_encode(ctx, d_text, n, d_start, d_out, out_elem_bytes, cap, n_out, s, false);
}

int bpe_decode(bpe_ctx* ctx, const void* d_ids, int id_elem_bytes, uint64_t n_ids, const uint8_t* h_vocab_blob,
               const uint64_t* h_vocab_off, uint32_t n_vocab, uint8_t* d_bytes, uint64_t cap, uint64_t* n_bytes, void* stream) {
    if (!ctx || !n_bytes || !h_vocab_off || (id_elem_bytes != 2 && id_elem_bytes != 4)) return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_decode arguments") : BPE_ERR_ARGUMENT;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    return run_decode(ctx, d_ids, id_elem_bytes, n_ids, h_vocab_blob, h_vocab_off, n_vocab, d_bytes, cap, n_bytes, (cudaStream_t)stream);
}

int bpe_debug_key_mask(bpe_ctx* ctx, uint64_t mask) {
    if (!ctx) return BPE_ERR_ARGUMENT;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaDeviceSynchronize());
    CUDA_TRY(ctx, set_key_mask_count(mask));
    CUDA_TRY(ctx, set_key_mask_fused(mask));
    CUDA_TRY(ctx, set_key_mask_encode(mask));
    return BPE_OK;
}

int bpe_kernel_times(bpe_ctx* ctx, bpe_kernel_time* out, int max) {
    if (!ctx) return 0;
    int n = 0;
    for (auto& t : ctx->times) {
        if (n >= max) break;
        out[n++] = t;
    }
    return n;
}

int bpe_set_profiling(bpe_ctx* ctx, int on) {
    if (!ctx) return BPE_ERR_ARGUMENT;
    ctx->profiling = on != 0;  // on: CUDA-event timing per launch; either way the launch counters restart
    ctx->times.clear();
    return BPE_OK;
}

}  // extern "C"
