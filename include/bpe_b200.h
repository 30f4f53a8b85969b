/* bpe_b200.h — C-ABI of libbpe_b200.so: the BPE tokenizer hot path of milasd/BPE-Transformer on B200 (sm_100a).
 *
 * Plain C, `int` status (0 = ok), plain pointers and sizes; no torch types.  Pointers named d_* are DEVICE
 * pointers (e.g. torch.Tensor.data_ptr()), h_* are host pointers.  `stream` is a cudaStream_t passed as
 * void* (NULL = the legacy default stream).  One context per GPU; calls on one context must be serialised
 * by the caller.  The library never frees caller buffers; everything it allocates belongs to the context
 * and is released by bpe_ctx_destroy().  Every entry point returns only after its work on `stream` has
 * completed (it synchronises the stream to read its status word), unless stated otherwise.
 *
 * Citations give the reference code (paths relative to the reference repo) each entry point replaces.
 * The Python host side that mirrors the reference's own API on top of this is bpe_transformer_b200/.
 */
#ifndef BPE_B200_H
#define BPE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bpe_ctx bpe_ctx;

/* status codes */
#define BPE_OK 0
#define BPE_ERR_CUDA 1          /* a CUDA runtime call failed; see bpe_last_error() */
#define BPE_ERR_INVALID_UTF8 2  /* the text is not valid UTF-8 (the reference decodes with errors="ignore",
                                   pretokenization.py:194: the host side cleans the bytes and calls again) */
#define BPE_ERR_CAPACITY 3      /* a context-owned table or pool is full; enlarge with bpe_ctx_configure() */
#define BPE_ERR_ARGUMENT 4
#define BPE_ERR_HASH_COLLISION 5 /* two different pretokens share a 64-bit key and differ in the 32-bit check */
#define BPE_ERR_MISSING_ID 6    /* a merge whose joined bytes have no id was applied (reference: KeyError,
                                   bpe_tokenizer.py:290) or a byte / special token has no id (:175, :206) */
#define BPE_ERR_NO_PAIRS 7      /* nothing to merge (reference: AttributeError, bpe_trainer.py:237-240) */
#define BPE_ERR_STATE 8         /* call order: e.g. bpe_encode() before bpe_load_model() */
#define BPE_ERR_CARRIAGE_RETURN 9 /* bpe_encode() with BPE_FAIL_ON_CR met a '\r' in the text; nothing was encoded */

/* flags for bpe_scan() / bpe_encode() */
#define BPE_CUT_AFTER_NEWLINE 1u /* a segment boundary after every '\n': BPETokenizer.encode_iterable over
                                    an iterable of lines (bpe_tokenizer.py:368-379) */
#define BPE_FAIL_ON_CR 2u        /* bpe_encode() only: return BPE_ERR_CARRIAGE_RETURN if the text holds a '\r'.  For callers
                                    that stand in for a text-mode file handle, which would translate "\r\n" and "\r" to
                                    "\n" before the reference sees the lines: they translate on the host and call again */

/* Every text buffer handed to the library must be readable for BPE_TEXT_PAD bytes past its end
 * (contents irrelevant), and aligned to 16 bytes. */
#define BPE_TEXT_PAD 64

/* ---- context ---------------------------------------------------------------------------------------- */

/* Creates a context on CUDA device `device`.  Uploads the code-point class tables (include/
 * bpe_unicode_tables.h, generated from the `regex` module the reference scans with: pretokenization.py:2). */
int bpe_ctx_create(int device, bpe_ctx** out);
void bpe_ctx_destroy(bpe_ctx* ctx);
/* Text of the last error on this context (never NULL).  With ctx == NULL: the last bpe_ctx_create() error. */
const char* bpe_last_error(const bpe_ctx* ctx);

/* Sizes of the context-owned pretoken table: 2^table_log2 slots (32 B each) and a byte pool holding one
 * copy of every distinct pretoken.  Call before the first bpe_count(); drops any counted pretokens.
 * Defaults: table_log2 = 22, pool_bytes = 256 MiB. */
int bpe_ctx_configure(bpe_ctx* ctx, int table_log2, uint64_t pool_bytes);

/* Special tokens, concatenated in `h_blob`, token k = h_blob[h_off[k] .. h_off[k+1]).  The ORDER is the match
 * priority at one position; the host passes them longest first, which is what the reference's
 * split_on_special_tokens does (pretokenization.py:228-232).  n = 0 clears them. */
int bpe_set_special_tokens(bpe_ctx* ctx, const uint8_t* h_blob, const uint32_t* h_off, int n);

/* ---- host I/O staging (the step before and after the path) ----------------------------------- */

/* Reads bytes [offset, offset + n) of the file at `path` into the device buffer d_dst: up to 16 host threads, each
 * with its own copy stream and pair of persistent pinned buffers, read 2 MB pieces and copy them to the GPU.
 * Replaces the f.seek / f.read of the reference's workers (pretokenization.py:192-194).  The copy has completed on
 * return.  The two I/O calls are the exception to "one call at a time per context": ONE bpe_load_file and ONE
 * bpe_store_to_host may run (from other host threads) while another call works on the same context, which is how the
 * Python layer overlaps reading the next piece of a file with encoding this one and storing the previous one.
 * Thread count: BPE_B200_IO_THREADS, default min(16, cores / LOCAL_WORLD_SIZE). */
int bpe_load_file(bpe_ctx* ctx, const char* path, uint64_t offset, uint64_t n, uint8_t* d_dst, void* stream);

/* The same for n_ranges byte ranges of the file at once: range r = file bytes [file_off[r], file_off[r] + len[r]) goes to
 * d_base + dev_off[r].  The reference's workers each read their own byte range (pretokenization.py:96-109, :192-194);
 * here the ranges are the pieces of a large corpus, cut at special-token starts, that the GPU scans and counts while later
 * pieces are still on their way.  *h_ranges_done (host memory, may be NULL) is updated while the call runs: the number of
 * leading ranges that are completely in device memory.  The caller polls it from another thread and may use range r as
 * soon as the value exceeds r.  Same concurrency rule as bpe_load_file (and not beside a bpe_load_file). */
int bpe_load_file_ranges(bpe_ctx* ctx, const char* path, uint32_t n_ranges, const uint64_t* file_off, const uint64_t* len,
                         const uint64_t* dev_off, uint8_t* d_base, volatile uint64_t* h_ranges_done, void* stream);

/* Copies n bytes from device memory to an ordinary (pageable) host array through staging buffers of its own, e.g. the
 * ids of bpe_encode into a NumPy array — the 1-D token dataset format.  Waits for `stream` first.  A destination that
 * is page-locked already (cudaHostAlloc / cudaHostRegister) receives the DMA directly, without staging. */
int bpe_store_to_host(bpe_ctx* ctx, const void* d_src, uint64_t n, void* h_dst, void* stream);

/* ---- pretokenisation ---------------------------------------------------------------------------------- */

/* Token-start bitmask of d_text[0, n): bit (i & 31) of d_start[i >> 5] is set iff a token starts at byte i.
 * A token is a match of the GPT-2 split pattern (settings.py:8) as regex.finditer finds it inside one
 * segment (pretokenization.py:238-252), or one special token; segments are what is left after the text is
 * cut at `d_cuts` (sorted byte positions with a boundary in front of them; may be NULL) and/or after every
 * newline (flag), and then split on the special tokens (pretokenization.py:211-235).
 * d_start must hold n/32 + 2 words.  Returns BPE_ERR_INVALID_UTF8 if the text is not valid UTF-8. */
int bpe_scan(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint64_t* d_cuts, uint64_t n_cuts,
             uint32_t flags, uint32_t* d_start, void* stream);

/* Adds every pretoken of d_text (given its start mask) to the context's table of distinct pretokens with
 * occurrence counts — the Counter that pretokenize_chunk builds and parallel_pretokenization folds
 * (pretokenization.py:73-111, 171-208).  Special tokens are left out.  Callable once per shard / chunk. */
int bpe_count(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint32_t* d_start, void* stream);

/* bpe_scan + bpe_count in ONE pass over the text: pretokenize_chunk as a whole (pretokenization.py:171-208 — split on the
 * special tokens, the split pattern, the Counter).  The token starts stay in registers; every pretoken is hashed and
 * counted where it is found, so the text is read once and no start mask is written or read back.  Same segmentation
 * arguments as bpe_scan; same table, same results as bpe_scan followed by bpe_count.  d_start may be NULL; if it is not,
 * the start mask is written as well (n/32 + 2 words).  On BPE_ERR_INVALID_UTF8 the table may already hold counts of this
 * text: the caller resets it (bpe_unique_reset) before counting the repaired bytes. */
int bpe_pretokenize_count(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint64_t* d_cuts, uint64_t n_cuts,
                          uint32_t flags, uint32_t* d_start, void* stream);

/* Number of distinct pretokens counted so far, the sum of their lengths, and the total occurrences. */
int bpe_unique_stats(bpe_ctx* ctx, uint64_t* n_unique, uint64_t* n_bytes, uint64_t* n_tokens);

/* Copies the table out: pretoken k = d_blob[d_off[k] .. d_off[k] + d_len[k]) with count d_cnt[k] (order
 * unspecified).  d_blob holds n_bytes + BPE_TEXT_PAD bytes, the other arrays n_unique entries.  Used to
 * exchange per-GPU tables (all-gather) and to hand the Counter to Python. */
int bpe_unique_export(bpe_ctx* ctx, uint8_t* d_blob, uint64_t* d_off, uint32_t* d_len, uint64_t* d_cnt, void* stream);

/* Adds pretokens (same layout as bpe_unique_export; d_blob 16-byte aligned and padded) with their counts to
 * the table: the fold of the per-worker Counters (pretokenization.py:109), here across GPUs. */
int bpe_unique_import(bpe_ctx* ctx, const uint8_t* d_blob, const uint64_t* d_off, const uint32_t* d_len,
                      const uint64_t* d_cnt, uint64_t n_unique, void* stream);

/* Forgets all counted pretokens (and cached encodings). */
int bpe_unique_reset(bpe_ctx* ctx);

/* ---- training ------------------------------------------------------------------------------------------ */

/* Learns merges from the counted pretokens until the vocabulary holds `vocab_size` entries or no pair
 * qualifies: BPETrainer._initialize_pairs_cache + _merge_tokens + _merge_pair + _get_adjacency_pairs
 * (bpe_trainer.py:169-426), including its neighbour double counting and its tie-break "highest count, then
 * the lexicographically greater (bytes, bytes) pair" (bpe_trainer.py:36-67).
 * Initial vocabulary: ids 0..255 are the single bytes, ids 256..first_new_id-1 the special tokens, whose
 * bytes come in h_init_blob / h_init_off (first_new_id + 1 offsets, all first_new_id tokens).
 * Merge k creates id first_new_id + k = (h_merge_a[k], h_merge_b[k]); both arrays hold vocab_size -
 * first_new_id entries.  *n_merges receives the number learned. */
int bpe_train(bpe_ctx* ctx, uint32_t vocab_size, uint32_t first_new_id, const uint8_t* h_init_blob,
              const uint32_t* h_init_off, uint32_t* h_merge_a, uint32_t* h_merge_b, uint32_t* n_merges,
              void* stream);

/* ---- training on several GPUs of one box: heavy steps of the merge loop shared over NVLink --------------------------
 * (No counterpart in the reference, whose merge loop is one process: bpe_trainer.py:242-325.  BASELINE.json configs[2].)
 * Every GPU holds all pretokens (after the table exchange) and runs the same loop; in a step whose candidate list is long,
 * each GPU filters only the words it owns (a function of the pretoken's key), rewrites what it finds, and stores the keys of
 * its hits straight into the other GPUs' exchange buffers (peer memory); then all rewrite the others' hits too.  Results are
 * those of one GPU.  Set-up: every process calls bpe_peer_alloc (room for `hits_per_sender` hits per sender and step: at
 * least 2 * distinct pretokens / GPUs + 4096), the 64-byte handles are exchanged by the host (an all-gather), every process
 * calls bpe_peer_connect with all handles in rank order; bpe_train then shares its heavy steps.  All processes must call
 * bpe_train together.  bpe_peer_disconnect (after a host barrier) returns to the single-GPU loop and frees the buffer. */
int bpe_peer_alloc(bpe_ctx* ctx, uint32_t hits_per_sender, uint8_t h_handle[64]);
int bpe_peer_connect(bpe_ctx* ctx, int rank, int world, const uint8_t* h_handles /* world x 64 bytes */);
int bpe_peer_disconnect(bpe_ctx* ctx);

/* Timing of the last bpe_train() on this context, in milliseconds of device time:
 * out[0] word table + pair histogram, out[1] merge loop, out[2] number of loop iterations. */
int bpe_train_timing(bpe_ctx* ctx, double out[3]);

/* ---- encoding ------------------------------------------------------------------------------------------ */

/* Loads a tokenizer: BPETokenizer.__init__ (bpe_tokenizer.py:38-60).  Merge r joins ids (h_a[r], h_b[r])
 * into h_merged[r] (0xFFFFFFFF = the joined bytes have no id); rows are in priority order and are over
 * canonical ids, i.e. the host has already resolved bytes -> id the way the reference's dicts do
 * (:55 later duplicate wins, :85 later id wins).  h_byte_to_id[256] (0xFFFFFFFF = byte not in the vocab),
 * h_special_id[k] = id of special token k of bpe_set_special_tokens().  Clears cached encodings. */
int bpe_load_model(bpe_ctx* ctx, const uint32_t* h_a, const uint32_t* h_b, const uint32_t* h_merged,
                   uint32_t n_merges, const uint32_t* h_byte_to_id, const uint32_t* h_special_id);

/* BPETokenizer.encode (bpe_tokenizer.py:139-290) over d_text, segmented as bpe_scan() describes: every
 * pretoken becomes its byte ids, merged by repeatedly joining the leftmost pair of minimum merge rank;
 * special tokens become their id.  Ids are written in text order to d_out as uint16 (out_elem_bytes = 2)
 * or uint32 (4); `cap` = capacity in elements; *n_out = ids produced (if > cap nothing was written and
 * BPE_ERR_CAPACITY is returned).  d_start may be NULL (the library scans into its own scratch) or a mask
 * made by bpe_scan() with the same cuts and flags.  Encodings of distinct pretokens stay cached in the
 * context across calls (the reference's _encoding_cache, bpe_tokenizer.py:57-60, 232-236).  With BPE_FAIL_ON_CR in
 * `flags` a text that holds a '\r' is refused (BPE_ERR_CARRIAGE_RETURN, nothing written). */
int bpe_encode(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint64_t* d_cuts, uint64_t n_cuts,
               uint32_t flags, const uint32_t* d_start, void* d_out, int out_elem_bytes, uint64_t cap,
               uint64_t* n_out, void* stream);

/* BPETokenizer.decode's byte concatenation (bpe_tokenizer.py:119-137): writes vocab[id] for every id
 * (U+FFFD bytes EF BF BD for ids without an entry) to d_bytes; the vocabulary comes as h_vocab_blob /
 * h_vocab_off over ids 0..n_vocab-1 (empty entry = no such id).  *n_bytes = bytes produced; with d_bytes == NULL
 * only the size is computed. */
int bpe_decode(bpe_ctx* ctx, const void* d_ids, int id_elem_bytes, uint64_t n_ids, const uint8_t* h_vocab_blob,
               const uint64_t* h_vocab_off, uint32_t n_vocab, uint8_t* d_bytes, uint64_t cap, uint64_t* n_bytes,
               void* stream);

/* (No counterpart in the reference.)  bpe_decode for callers that cannot allocate device memory: a worker that
 * turns id batches back into text next to a training or serving process which has already handed all of the
 * GPU's memory to its own allocator (a raw cudaMalloc then fails or fragments that arena), or a latency-critical
 * loop that cannot afford the device-wide synchronisation a cudaMalloc / cudaFree implies.  It needs no context
 * and calls no cudaMalloc, cudaFree or any other allocating routine: inputs and outputs are the caller's
 * (the vocabulary too — d_vocab_blob / d_vocab_off are DEVICE pointers here, laid out as in bpe_decode), and the
 * per-tile sizes and offsets go into a fixed 768 KiB array inside the library image, which every GPU that loads
 * the library holds until the process exits.  Hence the limit: at most BPE_NOALLOC_MAX_IDS ids per call; more
 * returns BPE_ERR_CAPACITY with a message, and the caller splits the batch.
 *   - Works on the calling thread's current device.  The first call on a device looks the array's device address
 *     up (cudaGetSymbolAddress); make that first call before any stream capture begins on that device.
 *   - All calls on a GPU share the one array, which still holds whatever the previous call wrote.  Nothing in it has
 *     to start at zero (every word is written before it is read within a call), but two calls that overlap on the
 *     GPU — different streams or host threads — race on it: serialise them.
 *   - d_bytes == NULL only computes *n_bytes.  h_err (may be NULL) receives a message of at most h_err_cap bytes on
 *     failure.  Returns after the work on `stream` has completed. */
#define BPE_NOALLOC_MAX_IDS (1u << 24)
int bpe_decode_noalloc(const void* d_ids, int id_elem_bytes, uint64_t n_ids, const uint8_t* d_vocab_blob,
                       const uint64_t* d_vocab_off, uint32_t n_vocab, uint8_t* d_bytes, uint64_t cap,
                       uint64_t* n_bytes, void* stream, char* h_err, size_t h_err_cap);

/* Test hook (no counterpart in the reference): the 64-bit keys the library names pretokens by are ANDed with `mask` from now on
 * (all contexts of this process on this device; all ones = normal operation).  With a narrow mask distinct pretokens share keys;
 * the calls that add pretokens to the table (bpe_pretokenize_count, bpe_count, bpe_unique_import) and bpe_encode then return
 * BPE_ERR_HASH_COLLISION instead of merging them: the table's identity is the BYTES (the reference's Counter is keyed by the
 * bytes, pretokenization.py:201-207), the key only finds them.  Reset the table (bpe_unique_reset) after changing the mask. */
int bpe_debug_key_mask(bpe_ctx* ctx, uint64_t mask);

/* Kernels launched on this context since the last bpe_set_profiling(): fills up to `max` entries of
 * (name, milliseconds, launches); returns the number of entries.  Launches are always counted; milliseconds
 * (CUDA events on the launching stream) only while profiling is on. */
typedef struct { const char* name; double ms; uint64_t launches; } bpe_kernel_time;
int bpe_kernel_times(bpe_ctx* ctx, bpe_kernel_time* out, int max);
/* Switch per-kernel CUDA-event timing on (1) or off (0, default) and restart the counters. */
int bpe_set_profiling(bpe_ctx* ctx, int on);

#ifdef __cplusplus
}
#endif
#endif
