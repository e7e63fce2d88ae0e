// cvx_stage.cpp -- pinned, chunked, multi-threaded staging between PAGEABLE host memory and the device, behind the
// host-pointer entry points (cvxgpu_compress / cvxgpu_decompress, hence the CvxCompress class and cvx_*).
//
// The reference's callers hand over posix_memalign / malloc memory (Test_With_Generated_Input.cpp:42-53,
// Compress_SEAM_Basin.cpp:52-71).  cudaMemcpy from such memory runs at the speed of ONE driver thread copying through a
// small internal bounce buffer (measured here: 14.8 GB/s each way, 700 ms per config-3 round trip against 190 ms from
// pinned memory).  Here T worker threads (default 12, at most the host's cores minus two; CVX_GPU_STAGE_THREADS) each own
// two pinned bounce buffers of 4 MiB (CVX_GPU_STAGE_CHUNK_MB), a private non-blocking stream and two events (sweep on the
// B200 box, 16 cores, config 3, ms per direction: 4 threads 214 / 222, 6: 118 / 138, 8: 106 / 102 with 4 MiB chunks,
// 12: 106 / 111; pinned memory: 95 / 95): thread t takes the chunks c = t, t + T, ...; while the DMA engine
// moves one of its buffers the thread memcpy()s the next chunk into (or the previous chunk out of) the other.  The host
// copies of the T threads run in parallel on the host cores, the DMA transfers of all threads queue on the copy engine
// back to back, so the PCIe link stays busy.  Pinned memory handed over by the caller skips all this (one cudaMemcpyAsync).
//
// Not done, on purpose: overlapping the global-RMS pass with the upload (0.8 ms of kernel against ~100 ms of copy) and
// downloading decoded slabs while later blocks decode (2.3 ms of kernel): the host-pointer path is bound by the PCIe
// transfer of the volume, 4*N bytes each way, which no reordering removes.
#include "../../include/cvx_gpu.h"
#include "cvx_internal.h"

#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

int cvx_internal_fail(int code, const char* fmt, ...);

namespace {
static size_t chunk_bytes()
{
    static size_t c = 0;
    if (!c) {
        long mb = 4;
        if (const char* e = getenv("CVX_GPU_STAGE_CHUNK_MB")) mb = atol(e);
        if (mb < 1) mb = 1;
        if (mb > 256) mb = 256;
        c = (size_t)mb << 20;
    }
    return c;
}
#define kChunk (chunk_bytes())
constexpr int kMaxThreads = 16;

int stage_threads()
{
    static int n = -1;
    if (n < 0) {
        const int hw = (int)std::thread::hardware_concurrency();
        int v = 12;
        if (hw > 0 && v > hw - 2) v = hw - 2;
        if (const char* e = getenv("CVX_GPU_STAGE_THREADS")) v = atoi(e);
        if (hw > 0 && v > hw) v = hw;
        n = std::max(1, std::min(v, kMaxThreads));
    }
    return n;
}

// 0 = pageable (or unknown) host memory, 1 = pinned / registered host memory
int host_is_pinned(const void* p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return 0; }
    return a.type == cudaMemoryTypeHost ? 1 : 0;
}
}  // namespace

static int stage_reserve(cvxgpu_workspace* ws, int T)
{
    const size_t want = (size_t)T * 2 * kChunk;
    if (ws->h_stage_bytes >= want) return CVXGPU_OK;
    if (ws->h_stage) cudaFreeHost(ws->h_stage);
    ws->h_stage = nullptr; ws->h_stage_bytes = 0;
    cudaError_t e = cudaHostAlloc(reinterpret_cast<void**>(&ws->h_stage), want, cudaHostAllocDefault);
    if (e != cudaSuccess) return cvx_internal_fail(CVXGPU_ERR_CUDA, "pinned staging buffers (%zu bytes): %s", want, cudaGetErrorString(e));
    ws->h_stage_bytes = want;
    return CVXGPU_OK;
}

// One direction of the staged transfer.  to_device: h -> d, else d -> h.  Blocks until every byte has arrived.
static int staged_copy(cvxgpu_workspace* ws, void* d_ptr, void* h_ptr, size_t bytes, bool to_device)
{
    const int T = (int)std::min<size_t>((size_t)stage_threads(), (bytes + kChunk - 1) / kChunk);
    int rc = stage_reserve(ws, T);
    if (rc) return rc;
    const size_t nchunks = (bytes + kChunk - 1) / kChunk;
    int dev = 0;
    cudaGetDevice(&dev);
    std::atomic<int> err{0};
    auto worker = [&](int t) {
        cudaSetDevice(dev);
        cudaStream_t st;
        cudaEvent_t ev[2];
        if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) { err.store((int)cudaGetLastError()); return; }
        cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming);
        unsigned char* buf[2] = {ws->h_stage + (size_t)(2 * t) * kChunk, ws->h_stage + (size_t)(2 * t + 1) * kChunk};
        unsigned char* d = static_cast<unsigned char*>(d_ptr);
        unsigned char* h = static_cast<unsigned char*>(h_ptr);
        cudaError_t e = cudaSuccess;
        if (to_device) {
            int k = 0;
            for (size_t c = (size_t)t; c < nchunks && e == cudaSuccess; c += (size_t)T, ++k) {
                const size_t off = c * kChunk, n = std::min(kChunk, bytes - off);
                const int b = k & 1;
                if (k >= 2) e = cudaEventSynchronize(ev[b]);              // the DMA that last read this buffer is done
                memcpy(buf[b], h + off, n);
                if (e == cudaSuccess) e = cudaMemcpyAsync(d + off, buf[b], n, cudaMemcpyHostToDevice, st);
                if (e == cudaSuccess) e = cudaEventRecord(ev[b], st);
            }
        } else {
            // DMA of chunk k+1 runs while chunk k is copied out of its bounce buffer
            size_t prev_off = 0, prev_n = 0;
            int k = 0, prev_b = -1;
            for (size_t c = (size_t)t; c < nchunks && e == cudaSuccess; c += (size_t)T, ++k) {
                const size_t off = c * kChunk, n = std::min(kChunk, bytes - off);
                const int b = k & 1;
                e = cudaMemcpyAsync(buf[b], d + off, n, cudaMemcpyDeviceToHost, st);
                if (e == cudaSuccess) e = cudaEventRecord(ev[b], st);
                if (prev_b >= 0 && e == cudaSuccess) {
                    e = cudaEventSynchronize(ev[prev_b]);
                    if (e == cudaSuccess) memcpy(h + prev_off, buf[prev_b], prev_n);
                }
                prev_off = off; prev_n = n; prev_b = b;
            }
            if (prev_b >= 0 && e == cudaSuccess) {
                e = cudaEventSynchronize(ev[prev_b]);
                if (e == cudaSuccess) memcpy(h + prev_off, buf[prev_b], prev_n);
            }
        }
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) err.store((int)e);
        cudaEventDestroy(ev[0]); cudaEventDestroy(ev[1]);
        cudaStreamDestroy(st);
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < T; ++t) pool.emplace_back(worker, t);
    worker(0);
    for (auto& th : pool) th.join();
    if (err.load()) return cvx_internal_fail(CVXGPU_ERR_CUDA, "staged %s copy: %s", to_device ? "host-to-device" : "device-to-host",
                                             cudaGetErrorString((cudaError_t)err.load()));
    return CVXGPU_OK;
}

// host -> device.  Work already enqueued on `st` that writes d_dst must be complete (callers upload first).
int cvx_stage_upload(cvxgpu_workspace* ws, void* d_dst, const void* h_src, size_t bytes, cudaStream_t st)
{
    if (bytes == 0) return CVXGPU_OK;
    if (host_is_pinned(h_src) || bytes < (4u << 20)) {
        cudaError_t e = cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) return cvx_internal_fail(CVXGPU_ERR_CUDA, "host-to-device copy: %s", cudaGetErrorString(e));
        return CVXGPU_OK;
    }
    return staged_copy(ws, d_dst, const_cast<void*>(h_src), bytes, true);
}

// device -> host, after everything enqueued on `st` so far has finished
int cvx_stage_download(cvxgpu_workspace* ws, void* h_dst, const void* d_src, size_t bytes, cudaStream_t st)
{
    if (bytes == 0) return CVXGPU_OK;
    if (host_is_pinned(h_dst) || bytes < (4u << 20)) {
        cudaError_t e = cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) return cvx_internal_fail(CVXGPU_ERR_CUDA, "device-to-host copy: %s", cudaGetErrorString(e));
        return CVXGPU_OK;
    }
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return cvx_internal_fail(CVXGPU_ERR_CUDA, "device-to-host copy: %s", cudaGetErrorString(e));
    return staged_copy(ws, const_cast<void*>(d_src), h_dst, bytes, false);
}
