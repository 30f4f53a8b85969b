// io.cu — host I/O staging for the tokenizer path: file -> pinned staging -> HBM and
// HBM -> pinned staging -> host array, double-buffered so that reading / copying on the host overlaps the DMA.
//
// The reference reads its corpus with f.read() per worker (pretokenization.py:192-194); here the whole byte range goes to
// one device buffer.  Reading 2 GB through a single thread (~5 GB/s from the page cache) and pinning 2 GB of fresh host
// memory per call cost more than every kernel together, hence: a few persistent pinned buffers per context, a handful of
// reader threads per chunk, cudaMemcpyAsync per chunk on a copy stream.
#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <cstring>
#include <thread>
#include <vector>

#include "common.cuh"

namespace {

constexpr int kStageBuffers = 3;
constexpr size_t kStageBytes = 32u << 20;

struct Staging {
    uint8_t* buf[kStageBuffers] = {nullptr, nullptr, nullptr};
    cudaEvent_t done[kStageBuffers] = {nullptr, nullptr, nullptr};
    cudaStream_t copy = nullptr;
};

Staging* staging_of(bpe_ctx* ctx) {
    if (ctx->io_staging) return static_cast<Staging*>(ctx->io_staging);
    Staging* st = new Staging();
    bool ok = cudaStreamCreateWithFlags(&st->copy, cudaStreamNonBlocking) == cudaSuccess;
    for (int k = 0; ok && k < kStageBuffers; k++)
        ok = cudaHostAlloc((void**)&st->buf[k], kStageBytes, cudaHostAllocDefault) == cudaSuccess &&
             cudaEventCreateWithFlags(&st->done[k], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) {
        ctx_fail(ctx, BPE_ERR_CUDA, "allocating pinned staging buffers", cudaGetLastError());
        delete st;
        return nullptr;
    }
    ctx->io_staging = st;
    return st;
}

int n_io_threads() {
    unsigned hw = std::thread::hardware_concurrency();
    return (int)std::max(1u, std::min(8u, hw ? hw : 4u));
}

// Fills dst[0, n) from the file at offset `off` with several threads; returns false on a short read.
bool parallel_pread(int fd, uint8_t* dst, size_t n, uint64_t off) {
    const int T = n < (4u << 20) ? 1 : n_io_threads();
    std::vector<std::thread> th;
    std::vector<int> bad(T, 0);
    const size_t per = (n + T - 1) / T;
    for (int t = 0; t < T; t++) {
        const size_t a = std::min(n, t * per), b = std::min(n, a + per);
        auto work = [fd, dst, off, a, b, &bad, t] {
            size_t got = a;
            while (got < b) {
                ssize_t k = pread(fd, dst + got, b - got, (off_t)(off + got));
                if (k <= 0) { bad[t] = 1; return; }
                got += (size_t)k;
            }
        };
        if (t + 1 == T) work(); else th.emplace_back(work);
    }
    for (auto& x : th) x.join();
    for (int v : bad) if (v) return false;
    return true;
}

void parallel_memcpy(uint8_t* dst, const uint8_t* src, size_t n) {
    const int T = n < (4u << 20) ? 1 : n_io_threads();
    std::vector<std::thread> th;
    const size_t per = (n + T - 1) / T;
    for (int t = 0; t < T; t++) {
        const size_t a = std::min(n, t * per), b = std::min(n, a + per);
        auto work = [dst, src, a, b] { memcpy(dst + a, src + a, b - a); };
        if (t + 1 == T) work(); else th.emplace_back(work);
    }
    for (auto& x : th) x.join();
}

}  // namespace

void io_staging_free(bpe_ctx* ctx) {
    if (!ctx->io_staging) return;
    Staging* st = static_cast<Staging*>(ctx->io_staging);
    for (int k = 0; k < kStageBuffers; k++) {
        if (st->buf[k]) cudaFreeHost(st->buf[k]);
        if (st->done[k]) cudaEventDestroy(st->done[k]);
    }
    if (st->copy) cudaStreamDestroy(st->copy);
    delete st;
    ctx->io_staging = nullptr;
}

extern "C" int bpe_load_file(bpe_ctx* ctx, const char* path, uint64_t offset, uint64_t n, uint8_t* d_dst, void* stream) {
    if (!ctx || !path || (n && !d_dst)) return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_load_file arguments") : BPE_ERR_ARGUMENT;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (n == 0) return BPE_OK;
    Staging* st = staging_of(ctx);
    if (!st) return BPE_ERR_CUDA;
    int fd = open(path, O_RDONLY);
    if (fd < 0) return ctx_fail(ctx, BPE_ERR_ARGUMENT, "cannot open file");
    int rc = BPE_OK;
    uint64_t done = 0;
    for (int i = 0; done < n; i++) {
        const int b = i % kStageBuffers;
        const size_t len = (size_t)std::min<uint64_t>(kStageBytes, n - done);
        if (i >= kStageBuffers && cudaEventSynchronize(st->done[b]) != cudaSuccess) { rc = ctx_fail(ctx, BPE_ERR_CUDA, "staging event", cudaGetLastError()); break; }
        if (!parallel_pread(fd, st->buf[b], len, offset + done)) { rc = ctx_fail(ctx, BPE_ERR_ARGUMENT, "short read"); break; }
        if (cudaMemcpyAsync(d_dst + done, st->buf[b], len, cudaMemcpyHostToDevice, st->copy) != cudaSuccess ||
            cudaEventRecord(st->done[b], st->copy) != cudaSuccess) { rc = ctx_fail(ctx, BPE_ERR_CUDA, "host-to-device copy", cudaGetLastError()); break; }
        done += len;
    }
    close(fd);
    cudaError_t e = cudaStreamSynchronize(st->copy);
    if (rc == BPE_OK && e != cudaSuccess) rc = ctx_fail(ctx, BPE_ERR_CUDA, "host-to-device copy", e);
    (void)stream;  // the copy has completed on return: any stream may read the buffer afterwards
    return rc;
}

extern "C" int bpe_store_to_host(bpe_ctx* ctx, const void* d_src, uint64_t n, void* h_dst, void* stream) {
    if (!ctx || (n && (!d_src || !h_dst))) return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_store_to_host arguments") : BPE_ERR_ARGUMENT;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (n == 0) return BPE_OK;
    Staging* st = staging_of(ctx);
    if (!st) return BPE_ERR_CUDA;
    CUDA_TRY(ctx, cudaStreamSynchronize((cudaStream_t)stream));  // the producer of d_src has finished
    const uint8_t* src = static_cast<const uint8_t*>(d_src);
    uint8_t* dst = static_cast<uint8_t*>(h_dst);
    const uint64_t n_chunks = (n + kStageBytes - 1) / kStageBytes;
    // chunk i is copied to staging buffer i % K; it is drained to the destination while chunk i + 1 .. i + K - 1 are in flight
    for (uint64_t i = 0; i < n_chunks + kStageBuffers - 1; i++) {
        if (i >= (uint64_t)(kStageBuffers - 1)) {
            const uint64_t j = i - (kStageBuffers - 1);  // drain chunk j
            if (j < n_chunks) {
                const int b = (int)(j % kStageBuffers);
                CUDA_TRY(ctx, cudaEventSynchronize(st->done[b]));
                parallel_memcpy(dst + j * kStageBytes, st->buf[b], (size_t)std::min<uint64_t>(kStageBytes, n - j * kStageBytes));
            }
        }
        if (i < n_chunks) {
            const int b = (int)(i % kStageBuffers);
            const size_t len = (size_t)std::min<uint64_t>(kStageBytes, n - i * kStageBytes);
            CUDA_TRY(ctx, cudaMemcpyAsync(st->buf[b], src + i * kStageBytes, len, cudaMemcpyDeviceToHost, st->copy));
            CUDA_TRY(ctx, cudaEventRecord(st->done[b], st->copy));
        }
    }
    return BPE_OK;
}
