// io.cu — host I/O staging for the tokenizer path: file -> pinned staging -> HBM and
// HBM -> pinned staging -> host array, double-buffered so that reading / copying on the host overlaps the DMA.
//
// The reference reads its corpus with f.read() per worker (pretokenization.py:192-194); here the whole byte range goes to
// one device buffer.  Reading 2 GB through a single thread (~5 GB/s from the page cache) and pinning 2 GB of fresh host
// memory per call cost more than every kernel together, hence: persistent pinned buffers per context and up to 16 host
// threads ("lanes"), each reading 2 MB chunks into its own pair of buffers and copying them on its own stream.
#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "common.cuh"

namespace {

// Every lane is one host thread with its own copy stream and two pinned buffers: while the DMA engine drains one buffer the
// thread fills (or empties) the other.  Lanes take chunks off a shared counter and never wait for one another.
constexpr int kMaxLanes = 16;
// small enough that the DMA engine starts early in every call, large enough for full PCIe rate (BPE_B200_IO_CHUNK_KB: developer override)
static size_t lane_bytes_from_env() {
    if (const char* e = getenv("BPE_B200_IO_CHUNK_KB")) { long v = atol(e); if (v >= 64 && v <= (64 << 10)) return (size_t)v << 10; }
    return 2u << 20;
}
static const size_t kLaneBytes = lane_bytes_from_env();

struct Staging {
    int lanes = 0;
    int device = 0;
    uint8_t* block = nullptr;  // lanes * 2 * kLaneBytes of pinned memory
    cudaStream_t copy[kMaxLanes] = {};
    cudaEvent_t done[kMaxLanes][2] = {};
    uint8_t* buf(int lane, int b) const { return block + ((size_t)lane * 2 + b) * kLaneBytes; }

    // Lane threads 1 .. lanes-1 live as long as the context (lane 0 is the calling thread): a call publishes a job, the
    // threads run job(lane) and report back.  Starting threads per call cost about 2 ms per 128 MB piece.
    std::vector<std::thread> pool;
    std::mutex mu;
    std::condition_variable wake, finished;
    std::function<int(int)> job;
    uint64_t generation = 0;
    int active = 0, pending = 0, first_error = 0;
    bool quit = false;

    void lane_main(int lane) {
        cudaSetDevice(device);
        uint64_t seen = 0;
        for (;;) {
            std::function<int(int)> mine;
            {
                std::unique_lock<std::mutex> lock(mu);
                wake.wait(lock, [&] { return quit || generation != seen; });
                if (quit) return;
                seen = generation;
                if (lane >= active) continue;
                mine = job;
            }
            const int rc = mine(lane);
            {
                std::lock_guard<std::mutex> lock(mu);
                if (rc && !first_error) first_error = rc;
                if (--pending == 0) finished.notify_all();
            }
        }
    }

    // Runs work(lane) on min(lanes, n_chunks) lanes; returns the first non-zero result.
    int run(uint64_t n_chunks, std::function<int(int)> work) {
        const int T = (int)std::min<uint64_t>((uint64_t)lanes, std::max<uint64_t>(1, n_chunks));
        if (T > 1) {
            std::lock_guard<std::mutex> lock(mu);
            job = work;
            active = T;
            pending = T - 1;
            first_error = 0;
            generation++;
        }
        if (T > 1) wake.notify_all();
        const int rc0 = work(0);
        if (T == 1) return rc0;
        std::unique_lock<std::mutex> lock(mu);
        finished.wait(lock, [&] { return pending == 0; });
        return rc0 ? rc0 : first_error;
    }
};

int n_io_lanes() {
    if (const char* e = getenv("BPE_B200_IO_THREADS")) {
        int v = atoi(e);
        if (v >= 1) return std::min(v, kMaxLanes);
    }
    unsigned hw = std::thread::hardware_concurrency();
    if (!hw) hw = 4;
    unsigned share = 1;  // processes of this node that load at the same time (one per GPU under torchrun)
    if (const char* e = getenv("LOCAL_WORLD_SIZE")) { int v = atoi(e); if (v > 1) share = (unsigned)v; }
    return (int)std::max(2u, std::min((unsigned)kMaxLanes, hw / share));
}

void staging_free(Staging* st) {
    {
        std::lock_guard<std::mutex> lock(st->mu);
        st->quit = true;
    }
    st->wake.notify_all();
    for (auto& t : st->pool) t.join();
    for (int l = 0; l < kMaxLanes; l++) {
        if (st->copy[l]) cudaStreamDestroy(st->copy[l]);
        for (int b = 0; b < 2; b++) if (st->done[l][b]) cudaEventDestroy(st->done[l][b]);
    }
    if (st->block) cudaFreeHost(st->block);
    delete st;
}

Staging* staging_of(bpe_ctx* ctx, int which) {
    if (ctx->io_staging[which]) return static_cast<Staging*>(ctx->io_staging[which]);
    Staging* st = new Staging();
    st->lanes = n_io_lanes();
    st->device = ctx->device;
    bool ok = cudaHostAlloc((void**)&st->block, (size_t)st->lanes * 2 * kLaneBytes, cudaHostAllocDefault) == cudaSuccess;
    for (int l = 0; ok && l < st->lanes; l++)
        ok = cudaStreamCreateWithFlags(&st->copy[l], cudaStreamNonBlocking) == cudaSuccess &&
             cudaEventCreateWithFlags(&st->done[l][0], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&st->done[l][1], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) {
        ctx_fail(ctx, BPE_ERR_CUDA, "allocating pinned staging buffers", cudaGetLastError());
        staging_free(st);
        return nullptr;
    }
    for (int l = 1; l < st->lanes; l++) st->pool.emplace_back([st, l] { st->lane_main(l); });
    ctx->io_staging[which] = st;
    return st;
}

bool pread_all(int fd, uint8_t* dst, size_t n, uint64_t off) {
    size_t got = 0;
    while (got < n) {
        ssize_t k = pread(fd, dst + got, n - got, (off_t)(off + got));
        if (k <= 0) return false;
        got += (size_t)k;
    }
    return true;
}

}  // namespace

void io_staging_free(bpe_ctx* ctx) {
    for (int which = 0; which < 2; which++) {
        if (ctx->io_staging[which]) staging_free(static_cast<Staging*>(ctx->io_staging[which]));
        ctx->io_staging[which] = nullptr;
    }
}

// Lane results: 1 = short read, 2 = CUDA error.
extern "C" int bpe_load_file(bpe_ctx* ctx, const char* path, uint64_t offset, uint64_t n, uint8_t* d_dst, void* stream) {
    if (!ctx || !path || (n && !d_dst)) return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_load_file arguments") : BPE_ERR_ARGUMENT;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (n == 0) return BPE_OK;
    Staging* st = staging_of(ctx, 0);
    if (!st) return BPE_ERR_CUDA;
    int fd = open(path, O_RDONLY);
    if (fd < 0) return ctx_fail(ctx, BPE_ERR_ARGUMENT, "cannot open file");
    const uint64_t n_chunks = (n + kLaneBytes - 1) / kLaneBytes;
    std::atomic<uint64_t> next{0};
    int rc = st->run(n_chunks, [&](int lane) -> int {
        int bad = 0;
        for (int it = 0;; it++) {
            const uint64_t c = next.fetch_add(1);
            if (c >= n_chunks) break;
            const int b = it & 1;
            const size_t len = (size_t)std::min<uint64_t>(kLaneBytes, n - c * kLaneBytes);
            if (it >= 2 && cudaEventSynchronize(st->done[lane][b]) != cudaSuccess) { bad = 2; break; }
            if (!pread_all(fd, st->buf(lane, b), len, offset + c * kLaneBytes)) { bad = 1; break; }
            if (cudaMemcpyAsync(d_dst + c * kLaneBytes, st->buf(lane, b), len, cudaMemcpyHostToDevice, st->copy[lane]) != cudaSuccess ||
                cudaEventRecord(st->done[lane][b], st->copy[lane]) != cudaSuccess) { bad = 2; break; }
        }
        if (cudaStreamSynchronize(st->copy[lane]) != cudaSuccess && !bad) bad = 2;
        return bad;
    });
    close(fd);
    (void)stream;  // the copy has completed on return: any stream may read the buffer afterwards
    if (rc == 1) return ctx_fail(ctx, BPE_ERR_ARGUMENT, "short read");
    if (rc) return ctx_fail(ctx, BPE_ERR_CUDA, "host-to-device copy", cudaGetLastError());
    return BPE_OK;
}

// Several byte ranges of one file, each to its own place in a device buffer, in one pass of the lanes; *h_ranges_done
// (ordinary host memory, polled by the caller from another thread) counts the leading ranges that have landed completely.
extern "C" int bpe_load_file_ranges(bpe_ctx* ctx, const char* path, uint32_t n_ranges, const uint64_t* file_off, const uint64_t* len,
                                    const uint64_t* dev_off, uint8_t* d_base, volatile uint64_t* h_ranges_done, void* stream) {
    if (!ctx || !path || (n_ranges && (!file_off || !len || !dev_off || !d_base)))
        return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_load_file_ranges arguments") : BPE_ERR_ARGUMENT;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (h_ranges_done) *h_ranges_done = 0;
    Staging* st = staging_of(ctx, 0);
    if (!st) return BPE_ERR_CUDA;
    struct Chunk { uint64_t file_off, dev_off; uint32_t len, range; };
    std::vector<Chunk> chunks;
    std::vector<uint64_t> range_end(n_ranges);  // chunks [0, range_end[r]) cover ranges 0..r
    for (uint32_t r = 0; r < n_ranges; r++) {
        for (uint64_t o = 0; o < len[r]; o += kLaneBytes)
            chunks.push_back(Chunk{file_off[r] + o, dev_off[r] + o, (uint32_t)std::min<uint64_t>(kLaneBytes, len[r] - o), r});
        range_end[r] = chunks.size();
    }
    const uint64_t n_chunks = chunks.size();
    auto publish = [&](uint64_t landed_chunks) {
        if (!h_ranges_done) return;
        uint64_t r = *h_ranges_done;
        while (r < n_ranges && range_end[r] <= landed_chunks) r++;
        __atomic_store_n(const_cast<uint64_t*>(h_ranges_done), r, __ATOMIC_RELEASE);
    };
    if (n_chunks == 0) { publish(0); return BPE_OK; }
    int fd = open(path, O_RDONLY);
    if (fd < 0) return ctx_fail(ctx, BPE_ERR_ARGUMENT, "cannot open file");
    std::unique_ptr<std::atomic<uint8_t>[]> landed(new std::atomic<uint8_t>[n_chunks]);
    for (uint64_t c = 0; c < n_chunks; c++) landed[c].store(0, std::memory_order_relaxed);
    std::mutex mark_mu;
    uint64_t watermark = 0;  // chunks [0, watermark) are in device memory
    auto mark = [&](uint64_t c) {
        landed[c].store(1, std::memory_order_release);
        std::lock_guard<std::mutex> lock(mark_mu);
        const uint64_t before = watermark;
        while (watermark < n_chunks && landed[watermark].load(std::memory_order_acquire)) watermark++;
        if (watermark != before) publish(watermark);
    };
    std::atomic<uint64_t> next{0};
    int rc = st->run(n_chunks, [&](int lane) -> int {
        int bad = 0;
        uint64_t in_flight[2] = {~0ull, ~0ull};  // chunk copied from buffer b and not yet known to have landed
        for (int it = 0;; it++) {
            const uint64_t c = next.fetch_add(1);
            if (c >= n_chunks) break;
            const int b = it & 1;
            const Chunk& ch = chunks[c];
            if (in_flight[b] != ~0ull) {
                if (cudaEventSynchronize(st->done[lane][b]) != cudaSuccess) { bad = 2; break; }
                mark(in_flight[b]);
                in_flight[b] = ~0ull;
            }
            if (!pread_all(fd, st->buf(lane, b), ch.len, ch.file_off)) { bad = 1; break; }
            if (cudaMemcpyAsync(d_base + ch.dev_off, st->buf(lane, b), ch.len, cudaMemcpyHostToDevice, st->copy[lane]) != cudaSuccess ||
                cudaEventRecord(st->done[lane][b], st->copy[lane]) != cudaSuccess) { bad = 2; break; }
            in_flight[b] = c;
            // the copy issued one round ago has had a whole pread to finish: report it now rather than a round later
            if (in_flight[b ^ 1] != ~0ull && cudaEventQuery(st->done[lane][b ^ 1]) == cudaSuccess) { mark(in_flight[b ^ 1]); in_flight[b ^ 1] = ~0ull; }
        }
        if (cudaStreamSynchronize(st->copy[lane]) != cudaSuccess && !bad) bad = 2;
        if (!bad)
            for (int b = 0; b < 2; b++)
                if (in_flight[b] != ~0ull) mark(in_flight[b]);
        return bad;
    });
    close(fd);
    (void)stream;
    if (rc == 1) return ctx_fail(ctx, BPE_ERR_ARGUMENT, "short read");
    if (rc) return ctx_fail(ctx, BPE_ERR_CUDA, "host-to-device copy", cudaGetLastError());
    return BPE_OK;
}

extern "C" int bpe_store_to_host(bpe_ctx* ctx, const void* d_src, uint64_t n, void* h_dst, void* stream) {
    if (!ctx || (n && (!d_src || !h_dst))) return ctx ? ctx_fail(ctx, BPE_ERR_ARGUMENT, "bpe_store_to_host arguments") : BPE_ERR_ARGUMENT;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (n == 0) return BPE_OK;
    Staging* st = staging_of(ctx, 1);
    if (!st) return BPE_ERR_CUDA;
    CUDA_TRY(ctx, cudaStreamSynchronize((cudaStream_t)stream));  // the producer of d_src has finished
    const uint8_t* src = static_cast<const uint8_t*>(d_src);
    uint8_t* dst = static_cast<uint8_t*>(h_dst);
    {   // a page-locked destination (cudaHostAlloc / cudaHostRegister, e.g. a pinned torch tensor) takes the DMA directly:
        // no staging buffer, no host memcpy, no first-touch page faults
        cudaPointerAttributes attr;
        if (cudaPointerGetAttributes(&attr, h_dst) == cudaSuccess && attr.type == cudaMemoryTypeHost) {
            CUDA_TRY(ctx, cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, st->copy[0]));
            CUDA_TRY(ctx, cudaStreamSynchronize(st->copy[0]));
            return BPE_OK;
        }
        cudaGetLastError();  // an ordinary pointer may be reported as an error by older drivers: not ours to keep
    }
    const uint64_t n_chunks = (n + kLaneBytes - 1) / kLaneBytes;
    std::atomic<uint64_t> next{0};
    int rc = st->run(n_chunks, [&](int lane) -> int {
        // the chunk copied in round it is moved to the destination in round it + 1, while the next chunk is in flight
        uint64_t prev_chunk = ~0ull;
        size_t prev_len = 0;
        for (int it = 0;; it++) {
            const uint64_t c = next.fetch_add(1);
            const int b = it & 1;
            if (c < n_chunks) {
                const size_t len = (size_t)std::min<uint64_t>(kLaneBytes, n - c * kLaneBytes);
                if (cudaMemcpyAsync(st->buf(lane, b), src + c * kLaneBytes, len, cudaMemcpyDeviceToHost, st->copy[lane]) != cudaSuccess ||
                    cudaEventRecord(st->done[lane][b], st->copy[lane]) != cudaSuccess) return 2;
            }
            if (prev_chunk != ~0ull) {
                if (cudaEventSynchronize(st->done[lane][b ^ 1]) != cudaSuccess) return 2;
                memcpy(dst + prev_chunk * kLaneBytes, st->buf(lane, b ^ 1), prev_len);
            }
            if (c >= n_chunks) break;
            prev_chunk = c;
            prev_len = (size_t)std::min<uint64_t>(kLaneBytes, n - c * kLaneBytes);
        }
        return 0;
    });
    if (rc) return ctx_fail(ctx, BPE_ERR_CUDA, "device-to-host copy", cudaGetLastError());
    return BPE_OK;
}
