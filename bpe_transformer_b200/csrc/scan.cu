// scan.cu — token-start bitmask of a UTF-8 text (the GPT-2 split pattern as a local, per-byte rule).
//
// Replaces regex.split on the special tokens + regex.finditer(PAT) of the reference
// (bpe_transformer/tokenization/preprocessing/pretokenization.py:211-252, pattern bpe_transformer/settings.py:8).
// The pattern needs no sequential state: whether a token starts at byte i depends on the classes of the
// characters from 4 before to 2 after i.  So every thread owns 32 bytes, classifies
// them SIMD-within-a-register (4 bytes per 32-bit op), reduces the classes to 32-bit masks and evaluates the
// rule with ~30 bitwise ops for all 32 bytes at once.  Rare things leave the bit-parallel path:
//   * a byte that may begin a special token  -> exact match + leftmost-non-overlapping resolution (scalar)
//   * non-ASCII bytes                        -> decode, two-stage class table, strict UTF-8 validation
//   * apostrophes                            -> contraction rule 's 'd 'm 't 'll 've 're (scalar)
// HBM traffic: N bytes read, N/8 bytes written.  Grid = persistent, a multiple of 148 SMs.
#define PRETOK_SLOW __forceinline__
#include "pretok.cuh"

#define SCAN_THREADS 256

namespace {

struct ScanParams : PretokParams {
    uint32_t* start;
    uint64_t n_tiles;
};

#define SCAN_LG 2                                    // ghost chunks in front of a tile's core (special tokens up to 60 bytes reach in)
#define SCAN_RG 1                                    // ghost chunk behind it
#define SCAN_CORE (SCAN_THREADS - SCAN_LG - SCAN_RG)  // chunks of 32 bytes a tile produces

// A tile is SCAN_THREADS chunks of 32 bytes, one per thread; the first two and the last one are ghosts: they are
// classified like the others (so that their neighbours can read the classes next to their own bytes from shared
// memory instead of classifying extra words) but their result belongs to the neighbouring tile.
template <bool NLMODE>
__global__ void __launch_bounds__(SCAN_THREADS) scan_kernel(const ScanParams P, const __grid_constant__ Specials sp) {
    __shared__ uint32_t wall[SCAN_THREADS + 2];  // bytes covered by a special token, per chunk of the tile
    __shared__ uint32_t sps[SCAN_THREADS + 2];   // first byte of a special token
    __shared__ uint32_t e_hi[SCAN_THREADS];      // classes of a chunk's last four bytes: 4 bits each of L, N, S, SP, NL, AP
    __shared__ uint32_t e_lo[SCAN_THREADS];      // ... of its first four bytes
    __shared__ uint32_t fbm[SCAN_THREADS];       // bytes that can begin a special token
    const uint32_t tid = threadIdx.x;
    const SpecialFirst sf = special_first(sp);

    for (uint64_t tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
        const int64_t chunk = (int64_t)tile * SCAN_CORE - SCAN_LG + tid;  // may lie in front of the text (first tile)
        const bool exists = chunk >= 0 && (uint64_t)chunk * 32 < P.n;
        const uint64_t c0 = exists ? (uint64_t)chunk * 32 : 0;
        const uint32_t valid = !exists ? 0u : (P.n - c0 >= 32 ? 0xFFFFFFFFu : ((1u << (uint32_t)(P.n - c0)) - 1u));
        wall[tid] = 0; sps[tid] = 0;
        if (tid < 2) { wall[SCAN_THREADS + tid] = 0; sps[SCAN_THREADS + tid] = 0; }

        // ---- phase A: 32 bytes -> class masks -------------------------------------------------------------------
        uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (exists) {
            const uint4* q = reinterpret_cast<const uint4*>(P.text + c0);
            uint4 a = __ldg(q), b = make_uint4(0, 0, 0, 0);
            if (c0 + 16 < P.n) b = __ldg(q + 1);
            w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
        }
        const ChunkClasses c = classify_chunk<NLMODE, false>(P, sp, sf, w, c0, valid);
        e_hi[tid] = edge_hi(c);
        e_lo[tid] = edge_lo(c);
        fbm[tid] = c.FB;
        __syncthreads();

        // ---- phase B: special tokens that start in this chunk ---------------------------------------------------------
        if (sp.n) {
            const uint64_t before = tid >= 2 ? ((uint64_t)fbm[tid - 2]) | ((uint64_t)fbm[tid - 1] << 32) : 0;  // 64 bytes in front of the chunk
            for (uint32_t m = c.FB; m; m &= m - 1) {
                const uint32_t b = __ffs(m) - 1;
                const uint32_t len = special_match(P, sp, c0, b, c.FB, before, tid >= 2);
                if (!len) continue;
                atomicOr(&sps[tid], 1u << b);
                const uint64_t bits = ((1ull << len) - 1ull);  // len <= 60: spans at most three chunks
                const uint64_t lo64 = bits << b;
                atomicOr(&wall[tid], (uint32_t)lo64);
                if (b + len > 32) atomicOr(&wall[tid + 1], (uint32_t)(lo64 >> 32));
                if (b + len > 64) atomicOr(&wall[tid + 2], (uint32_t)(bits >> (64 - b)));
            }
        }
        __syncthreads();

        // ---- phase C: the split pattern, 32 bytes at a time -------------------------------------------------------------------
        if (exists && tid >= SCAN_LG && tid < SCAN_THREADS - SCAN_RG)
            P.start[c0 >> 5] = split_chunk<NLMODE>(P, c, (uint64_t)chunk, c0, valid, e_hi[tid - 1], e_lo[tid + 1], wall[tid], wall[tid - 1], wall[tid + 1], sps[tid]);
        __syncthreads();  // shared arrays are rewritten by the next tile
    }
}

__global__ void cuts_to_mask_kernel(const uint64_t* __restrict__ cuts, uint64_t n_cuts, uint64_t n, uint32_t* __restrict__ mask) {
    uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (i < n_cuts) {
        uint64_t p = cuts[i];
        if (p < n) atomicOr(&mask[p >> 5], 1u << (p & 31));
    }
}

}  // namespace

// Does the text hold `byte`?  (bpe_encode's BPE_FAIL_ON_CR.)  16 bytes per thread and step, grid-stride; the buffer's
// padding lets the last load run past n, the tail bytes are masked out.
__global__ void __launch_bounds__(256) find_byte_kernel(const uint8_t* __restrict__ text, uint64_t n, uint32_t pattern, uint32_t bit, uint32_t* status) {
    const uint64_t n_vec = (n + 15) / 16;
    uint32_t hit = 0;
    for (uint64_t v = blockIdx.x * 256ull + threadIdx.x; v < n_vec; v += (uint64_t)gridDim.x * 256) {
        const uint4 q = __ldg(reinterpret_cast<const uint4*>(text) + v);
        uint32_t w[4] = {q.x ^ pattern, q.y ^ pattern, q.z ^ pattern, q.w ^ pattern};
        #pragma unroll
        for (int k = 0; k < 4; k++) {
            uint32_t z = (w[k] - 0x01010101u) & ~w[k] & 0x80808080u;  // 0x80 in (at least) every byte that was zero
            const uint64_t at = v * 16 + k * 4;
            if (at + 4 > n) z &= at >= n ? 0u : (0xFFFFFFFFu >> (8 * (at + 4 - n)));
            hit |= z;
        }
    }
    if (__any_sync(0xFFFFFFFFu, hit != 0) && (threadIdx.x & 31) == 0) atomicOr(status, bit);
}

int launch_find_byte(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, uint8_t byte, uint32_t status_bit, cudaStream_t s) {
    if (!n) return BPE_OK;
    KernelTimer t(ctx, "find_byte_kernel", s);
    const uint64_t want = (n / 16 + 255) / 256;
    find_byte_kernel<<<(unsigned)std::max<uint64_t>(1, std::min<uint64_t>(want, (uint64_t)BPE_N_SM * 8)), 256, 0, s>>>(d_text, n, byte * 0x01010101u, status_bit, ctx->d_status);
    CUDA_TRY(ctx, cudaGetLastError());
    return BPE_OK;
}

int launch_scan(bpe_ctx* ctx, const uint8_t* d_text, uint64_t n, const uint32_t* d_cutmask, uint32_t flags, uint32_t* d_start, cudaStream_t s) {
    CUDA_TRY(ctx, cudaMemsetAsync(d_start + (n / 32), 0, 2 * sizeof(uint32_t), s));
    if (n == 0) return BPE_OK;
    ScanParams P{};
    P.text = d_text; P.n = n; P.cutmask = d_cutmask; P.flags = flags; P.start = d_start; P.status = ctx->d_status; P.cls = ctx->cls;
    P.n_tiles = (n + SCAN_CORE * 32 - 1) / (SCAN_CORE * 32);
    uint64_t grid = P.n_tiles < (uint64_t)BPE_N_SM * 8 ? P.n_tiles : (uint64_t)BPE_N_SM * 8;
    KernelTimer t(ctx, "scan_kernel", s);
    if (flags & BPE_CUT_AFTER_NEWLINE) scan_kernel<true><<<(unsigned)grid, SCAN_THREADS, 0, s>>>(P, ctx->sp);
    else scan_kernel<false><<<(unsigned)grid, SCAN_THREADS, 0, s>>>(P, ctx->sp);
    CUDA_TRY(ctx, cudaGetLastError());
    return BPE_OK;
}

int launch_cuts_to_mask(bpe_ctx* ctx, const uint64_t* d_cuts, uint64_t n_cuts, uint64_t n, uint32_t* d_cutmask, cudaStream_t s) {
    CUDA_TRY(ctx, cudaMemsetAsync(d_cutmask, 0, (n / 32 + 2) * sizeof(uint32_t), s));
    if (n_cuts) {
        cuts_to_mask_kernel<<<(unsigned)((n_cuts + 255) / 256), 256, 0, s>>>(d_cuts, n_cuts, n, d_cutmask);
        CUDA_TRY(ctx, cudaGetLastError());
    }
    return BPE_OK;
}
