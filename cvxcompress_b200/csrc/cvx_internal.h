// cvx_internal.h -- host-side interface between the C ABI (cvx_api.cpp) and the kernel
// launchers (cvx_kernels.cu).  Not installed; the public surface is include/cvx_gpu.h.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

struct cvxgpu_workspace {
    int device = -1;
    int sm_count = 0;
    // device scalars (one 256-byte slab): sumsq, mulfac, ticket, payload_bytes, length, err
    unsigned char* d_scalars = nullptr;
    double* d_sumsq = nullptr;
    float* d_mulfac = nullptr;
    unsigned int* d_ticket = nullptr;
    long* d_payload_bytes = nullptr;
    long* d_length = nullptr;
    int* d_err = nullptr;
    double* d_partial = nullptr;          // per-CTA partial sums of k_sumsq
    int n_partial = 0;
    double* d_rowpartial = nullptr;       // per-CTA partial sums of k_sumsq_rows (64 per block row)
    double* d_rowsum = nullptr;           // one sum of squares per block row
    int rows_cap = 0;
    unsigned long long* d_status = nullptr;   // look-back words
    long status_cap = 0;
    float* d_scratch = nullptr;           // staged path: dense coefficient blocks
    size_t scratch_bytes = 0;
    // pinned host mirror for the few bytes read back per call
    unsigned char* h_pinned = nullptr;
    // staging buffers of the host-pointer entry points
    unsigned char* d_stage_vol = nullptr;
    size_t stage_vol_bytes = 0;
    unsigned char* d_stage_stream = nullptr;
    size_t stage_stream_bytes = 0;
    unsigned char* h_stage = nullptr;     // pinned bounce buffers of the staged pageable-memory copies (cvx_stage.cpp)
    size_t h_stage_bytes = 0;
    cudaStream_t priv_stream = nullptr;   // non-blocking stream of the host-pointer entry points (never the legacy stream 0)
};

int cvx_internal_default_ws(cvxgpu_workspace** out, void* cuda_stream);
int cvx_stage_upload(cvxgpu_workspace* ws, void* d_dst, const void* h_src, size_t bytes, cudaStream_t st);
int cvx_stage_download(cvxgpu_workspace* ws, void* h_dst, const void* d_src, size_t bytes, cudaStream_t st);

// Device allocations made while some stream of the process is being captured into a CUDA graph: cudaMalloc is refused
// in the (default) global capture mode, so the calling thread switches to the relaxed mode for the allocation alone
// (cudaThreadExchangeStreamCaptureMode).  The memory is ordinary device memory that outlives the graph.
struct CvxRelaxCapture {
    cudaStreamCaptureMode mode = cudaStreamCaptureModeRelaxed;
    CvxRelaxCapture() { cudaThreadExchangeStreamCaptureMode(&mode); }
    ~CvxRelaxCapture() { cudaThreadExchangeStreamCaptureMode(&mode); }
};

namespace cvx {

struct Shape {
    int nx, ny, nz;       // volume (or slab)
    int bx, by, bz;
    int nbx, nby, nbz;
    long nblk;
};
Shape make_shape(int nx, int ny, int nz, int bx, int by, int bz);
bool valid_block_size(int bx, int by, int bz);
bool fused_ok(int bx, int by, int bz);    // block fits the shared-memory kernels

// all launchers return a cudaError_t (cudaSuccess == 0) and never synchronise
cudaError_t ws_reserve_status(cvxgpu_workspace* ws, long nblk);
cudaError_t ws_reserve_scratch(cvxgpu_workspace* ws, size_t bytes);

cudaError_t launch_sumsq(cvxgpu_workspace* ws, const float* d_vol, long ncells, double* d_sumsq, cudaStream_t st);
cudaError_t ws_reserve_rows(cvxgpu_workspace* ws, int rows);
cudaError_t launch_sumsq_rows(cvxgpu_workspace* ws, const float* d_vol, long plane_cells, int bz, int nz, double* d_rowsum, cudaStream_t st);
cudaError_t launch_sum_rows(const double* d_rows, int nseg, int seg, const int* d_cnt, int n_single, double* d_total, cudaStream_t st);
cudaError_t launch_rms_header(const double* d_rows, int nseg, int seg, const int* d_cnt, int n_single, double* d_sumsq,
                              void* d_stream, const Shape& s, float scale, int use_local, long ncells_total, float* d_mulfac_out,
                              cudaStream_t st);
cudaError_t launch_write_header(void* d_stream, const Shape& s, float scale, int use_local, const double* d_sumsq,
                                long ncells_total, float* d_mulfac_out, cudaStream_t st);
cudaError_t launch_encode(cvxgpu_workspace* ws, const float* d_vol, const Shape& s, float scale, int use_local,
                          const float* d_glob_mulfac, unsigned long long* d_offs, float* d_blkmulfac,
                          void* d_payload, long cap, long* d_payload_bytes, int* d_err, int exact, cudaStream_t st);
cudaError_t launch_finish_length(const long* d_payload_bytes, long table_bytes, long* d_length, cudaStream_t st);
cudaError_t launch_offsets_add(unsigned long long* d_offs, long n, long base, cudaStream_t st);
cudaError_t launch_decode(cvxgpu_workspace* ws, float* d_vol, const Shape& s, float glob_mulfac, int use_local,
                          const unsigned long long* d_offs, const float* d_blkmulfac, const void* d_payload,
                          long payload_bytes, int exact, cudaStream_t st, const float* d_glob_mulfac = nullptr,
                          const long* d_payload_bytes = nullptr, const long* d_offs_bias = nullptr);
cudaError_t launch_shard_fixup(const long* d_sizes, int world, int rank, long* d_bases, unsigned long long* d_offs, long n, cudaStream_t st);
cudaError_t launch_shard_push(const void* d_src, int n, int kind, unsigned long long* const* d_peers, int world, int cap, int rank,
                              unsigned long long* d_epoch, cudaStream_t st);
cudaError_t launch_rms_header_mailbox(unsigned long long* d_mailbox, int world, int cap, const unsigned long long* d_epoch, const int* d_cnt,
                                      double* d_sumsq, void* d_header, const Shape& s, float scale, int use_local, long ncells_total,
                                      float* d_mulfac_out, int* d_status, cudaStream_t st);
cudaError_t launch_shard_fixup_mailbox(unsigned long long* d_mailbox, int world, int cap, const unsigned long long* d_epoch, int rank,
                                       long* d_bases, unsigned long long* d_offs, long n, int* d_status, cudaStream_t st);
size_t shard_mailbox_bytes(int world, int cap);
cudaError_t launch_shard_extent(const long* d_bases, int rank, long* d_out2, cudaStream_t st);
cudaError_t launch_wavelet(cvxgpu_workspace* ws, float* d_blocks, long nblocks, int bx, int by, int bz, int inverse,
                           int exact, cudaStream_t st);

}  // namespace cvx
