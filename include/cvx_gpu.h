/*
 * cvx_gpu.h -- thin C ABI of the B200 (sm_100a) CvxCompress hot path.
 *
 * This is the layer the host C++ (`CvxCompress` class, include/CvxCompress.hxx) calls, and the
 * one an FFI user (ctypes / ccall / cgo) binds when the volume and the stream already live in
 * device memory.  Plain C types only; `cuda_stream` is a `cudaStream_t` passed as `void*`
 * (NULL = the legacy default stream).  All `d_*` pointers are DEVICE pointers unless noted.
 *
 * Reference interfaces replaced (file:line under ChevronETC/CvxCompress):
 *   cvxgpu_compress_device     <- CvxCompress::Compress      CvxCompress.cpp:231-427  (decl CvxCompress.hxx:32-45)
 *   cvxgpu_decompress_device   <- CvxCompress::Decompress    CvxCompress.cpp:463-571  (decl CvxCompress.hxx:105-113)
 *   cvxgpu_sumsq_device        <- Compute_Global_RMS         CvxCompress.cpp:73-117
 *   cvxgpu_encode_blocks_device<- the block loop             CvxCompress.cpp:318-420  (Copy_To_Block, Wavelet_Transform_Fast_Forward,
 *                                                            Compute_Local_RMS, Run_Length_Encode_Slow, raw fallback, offset table)
 *   cvxgpu_decode_blocks_device<- the block loop             CvxCompress.cpp:528-568  (Run_Length_Decode_Slow, Wavelet_Transform_Fast_Inverse,
 *                                                            Copy_From_Block)
 *   cvxgpu_write_header_device <- header words 0..7          CvxCompress.cpp:284-300
 *   cvxgpu_is_valid_block_size <- Is_Valid_Block_Size        CvxCompress.cpp:54-71
 *   cvxgpu_wavelet_blocks_device <- Wavelet_Transform_Fast_{Forward,Inverse} on an array of dense blocks
 *                                                            (the module-test sweep, CvxCompress.cpp:803-891)
 *
 * Stream format (identical to the reference): 32-byte header, int64 offset per
 * block (bit 63 = raw block), optional fp32 mulfac per block, payload; length = 32 + 8*nnn
 * [+ 4*nnn] + payload_bytes + 7.  Blocks are laid out in natural iBlk order (what the reference
 * produces with num_threads=1).
 *
 * Every function returns CVXGPU_OK (0) or a negative error code; cvxgpu_last_error() gives the
 * message of the calling thread's last failure.  There is NO CPU fallback: without a CUDA
 * device every compute entry point fails with CVXGPU_ERR_CUDA.
 */
#ifndef CVX_GPU_H
#define CVX_GPU_H

#ifdef __cplusplus
extern "C" {
#endif

#define CVXGPU_OK 0
#define CVXGPU_ERR_ARG (-1)       /* bad dimension / block size / null pointer          */
#define CVXGPU_ERR_CUDA (-2)      /* CUDA runtime failure (message has the CUDA error)  */
#define CVXGPU_ERR_CAPACITY (-3)  /* stream would not fit in `capacity` bytes           */
#define CVXGPU_ERR_STREAM (-4)    /* malformed / mismatching compressed stream          */

/* arithmetic of the 9/7 transform.  EXACT = direct-form filter bank in the reference's operation
 * order (bit-identical coefficients and streams); LIFTING = 4-step lifting with FMA (same
 * operator, ~4e-7 rms-relative coefficient difference, 2.3x fewer FP32 instructions). */
#define CVXGPU_TRANSFORM_EXACT 0
#define CVXGPU_TRANSFORM_LIFTING 1

typedef struct cvxgpu_workspace cvxgpu_workspace;

const char* cvxgpu_last_error(void);
int  cvxgpu_device_count(void);

/* process-wide default transform (initialised from env CVX_GPU_TRANSFORM=exact|lifting; default exact) */
int  cvxgpu_set_transform(int mode);
int  cvxgpu_get_transform(void);

/* scratch owned by the caller: look-back status words, ticket, partial sums, staging -- the state of
 * ONE call in flight.  A workspace must not be used by two host threads at once, nor by two calls
 * that can overlap on the GPU (calls enqueued on different CUDA streams).  NULL may be passed to
 * every call below: the library then uses a default workspace of the calling host thread, one per
 * (current device, cuda_stream) pair, so that calls a thread enqueues on different streams never
 * share state; calls on one stream are ordered by the stream.  A default workspace is created by
 * the first call that needs it (device allocations: make that call outside a CUDA-graph capture);
 * cvxgpu_workspace_release_thread_default() synchronises and frees the calling thread's. */
int  cvxgpu_workspace_create(cvxgpu_workspace** ws);
int  cvxgpu_workspace_destroy(cvxgpu_workspace* ws);
int  cvxgpu_workspace_release_thread_default(void);

int  cvxgpu_is_valid_block_size(int bx, int by, int bz);
long cvxgpu_num_blocks(int nx, int ny, int nz, int bx, int by, int bz);
/* worst case (every block raw): 32 + 8*nnn [+ 4*nnn] + 4*bx*by*bz*nnn + 7 */
long cvxgpu_max_compressed_bytes(int nx, int ny, int nz, int bx, int by, int bz, int use_local_rms);

/* ---- whole-volume calls (one GPU) ------------------------------------------------------- */

/* Compress d_vol (nx fastest, nz slowest) into d_stream.  *length receives compressed_length,
 * *ratio the reference's return value 4*N/length.  Synchronises `cuda_stream` before returning
 * (the length is an output).  d_stream must be 8-byte aligned. */
int cvxgpu_compress_device(cvxgpu_workspace* ws, const float* d_vol, int nx, int ny, int nz,
                           int bx, int by, int bz, float scale, int use_local_rms,
                           void* d_stream, long capacity, long* length, float* ratio, void* cuda_stream);

/* Same work, no host synchronisation: the length is left in *d_length (device int64) and an
 * overflow / error code in *d_status (device int32, 0 = ok).  Used by bench.py's device-resident
 * timing and by callers that keep everything on a stream. */
int cvxgpu_compress_device_async(cvxgpu_workspace* ws, const float* d_vol, int nx, int ny, int nz,
                                 int bx, int by, int bz, float scale, int use_local_rms,
                                 void* d_stream, long capacity, long* d_length, int* d_status, void* cuda_stream);

/* Decompress d_stream into d_vol.  nx,ny,nz must match the header (the reference prints and
 * asserts, CvxCompress.cpp:473-488; here CVXGPU_ERR_STREAM).  Asynchronous on `cuda_stream`
 * apart from a 32-byte header read. */
int cvxgpu_decompress_device(cvxgpu_workspace* ws, float* d_vol, int nx, int ny, int nz,
                             const void* d_stream, long length, void* cuda_stream);

/* header fields of a stream held in HOST memory (no GPU needed) */
int cvxgpu_parse_header(const void* h_stream, long length, int* nx, int* ny, int* nz,
                        int* bx, int* by, int* bz, float* glob_mulfac, int* flags);

/* Bounds-checked walk of a stream held in HOST memory (no GPU needed): SURVEY.md 8(f) row 3.
 * The reference decoder trusts the stream entirely (Run_Length_Encode_Slow.cpp:393-525 reads
 * tokens until the block is full, wherever that leads) and never looks at compressed_length
 * (CvxCompress.cpp:463-571); its authors left Compress_Safe/Decompress_Safe commented out
 * (Test_With_Generated_Input.cpp:69-71,89).  This is that check, done before any byte is decoded:
 *   level 0  header fields, length >= 32 + tables + 7, every table offset inside the payload,
 *            a raw block's 4*B bytes inside the payload, finite non-zero multipliers;
 *   level 1  additionally walks every coded block token by token: each token must lie inside the
 *            payload, the values must add up to exactly B (a run may not overshoot the block),
 *            and no block may run into the next block's first byte.
 * Returns CVXGPU_OK or CVXGPU_ERR_STREAM (cvxgpu_last_error() names the first offending block);
 * *info (may be NULL) is filled as far as the walk got. */
typedef struct cvxgpu_stream_info {
    int nx, ny, nz, bx, by, bz;
    int use_local_rms;
    float glob_mulfac;
    long nblocks;
    long raw_blocks;        /* table entries with bit 63 set                                  */
    long table_bytes;       /* 8*nblocks (+ 4*nblocks with local RMS)                         */
    long payload_bytes;     /* length - 32 - table_bytes - 7                                  */
    long used_payload_bytes;/* level 1: sum of the block payloads the walk measured, else -1 */
    long min_block_bytes, max_block_bytes; /* level 1, else -1                               */
    long tokens;            /* level 1: tokens walked, else -1                               */
    int natural_order;      /* 1 when table offsets increase with iBlk (num_threads=1 order) */
} cvxgpu_stream_info;
int cvxgpu_validate_stream(const void* h_stream, long length, int level, cvxgpu_stream_info* info);

/* Opt-in safe decode for the host-pointer path (cvxgpu_decompress, hence CvxCompress::Decompress and
 * cvx_decompress_*): level -1 = off, the reference's behaviour (default); 0 / 1 = run
 * cvxgpu_validate_stream at that level on a HOST stream before it is staged to the GPU and return
 * CVXGPU_ERR_STREAM instead of decoding (the class then reports and aborts like every other
 * failure).  The initial level comes from the environment variable CVX_GPU_SAFE_DECODE ("0" or
 * "1").  Device-resident streams are not walked: copy them to the host and call
 * cvxgpu_validate_stream when they come from an untrusted source. */
int cvxgpu_set_safe_decode(int level);
int cvxgpu_get_safe_decode(void);

/* ---- host-pointer convenience (what the CvxCompress class calls) ---------------------------
 * `vol` / `stream` may be pageable host, pinned host, managed or device pointers
 * (cudaPointerGetAttributes); host memory is staged through pinned chunks. */
int cvxgpu_compress(const float* vol, int nx, int ny, int nz, int bx, int by, int bz, float scale,
                    int use_local_rms, void* stream_out, long* length, float* ratio);
int cvxgpu_decompress(float* vol, int nx, int ny, int nz, const void* stream_in, long length);

/* ---- pieces, for z-slab sharding over several GPUs --------------------------
 * A rank owns the block rows [blkz0, blkz1) of the volume: a contiguous iBlk range and a
 * contiguous z-slab of memory.  Sequence on every rank:
 *   cvxgpu_sumsq_device(slab)                     -> d_sumsq (device double)
 *   all-gather the doubles, add them in rank order -> d_sumsq_total       (NCCL, 8 bytes/rank)
 *   cvxgpu_encode_blocks_device(slab, ...)        -> slab table slice + slab payload, local offsets
 *   all-gather payload byte counts                -> this rank's base     (NCCL, 8 bytes/rank)
 *   cvxgpu_offsets_add_device(slice, base)
 * Decompress needs no collective: every rank decodes its own iBlk range from the full stream. */
int cvxgpu_sumsq_device(cvxgpu_workspace* ws, const float* d_vol, long ncells, double* d_sumsq, void* cuda_stream);

/* The same reduction per BLOCK ROW (the planes [i*bz, (i+1)*bz) of a volume or of a z-slab that starts
 * on a block-row boundary): d_rowsums[(nz + bz - 1) / bz].  Every row's sum is formed in an order that
 * depends on the row alone, and cvxgpu_sum_rows_device adds rows one after the other, so the global RMS
 * -- hence glob_mulfac and every byte of the stream -- is the same for any number of slabs
 * (Compute_Global_RMS, CvxCompress.cpp:73-117).  cvxgpu_compress_device uses these two. */
int cvxgpu_sumsq_rows_device(cvxgpu_workspace* ws, const float* d_vol, long plane_cells, int bz, int nz,
                             double* d_rowsums, void* cuda_stream);
int cvxgpu_sum_rows_device(const double* d_rowsums, int nrows, double* d_total, void* cuda_stream);

/* cvxgpu_sumsq_rows_device without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same kernels, same summation order, bit-identical d_rowsums; the
 * per-CTA partial sums live in a file-scope device array of the library instead of a workspace:
 * CVXGPU_STATIC_MAX_ROWS block rows x 128 partials x 8 bytes = 4 MiB, held by every GPU that loads the
 * library until the process exits.  The call takes no workspace and calls no cudaMalloc / cudaFree or any
 * other allocating routine.  More than CVXGPU_STATIC_MAX_ROWS block rows ((nz + bz - 1) / bz) are refused
 * with CVXGPU_ERR_CAPACITY (cvxgpu_last_error() says so); use cvxgpu_sumsq_rows_device then.
 * All calls on one device share the array: calls that OVERLAP on the GPU (two streams without an ordering
 * between them, two host threads) race on it -- order them on one stream or with events.  The array keeps
 * the previous call's partial sums; none of them is read by the next call (every slot read was written by
 * the same call first), so nothing needs clearing.  Asynchronous on `cuda_stream`. */
#define CVXGPU_STATIC_MAX_ROWS 4096
int cvxgpu_sumsq_rows_static_device(const float* d_vol, long plane_cells, int bz, int nz,
                                    double* d_rowsums, void* cuda_stream);

/* Encode the blocks of a volume (or slab) of nx*ny*nz_slab cells.  d_sumsq_total / ncells_total
 * define the global RMS (ignored when use_local_rms).  d_offs receives nblk entries (payload
 * offsets relative to d_payload, bit 63 = raw), d_blkmulfac nblk floats when use_local_rms (else
 * may be NULL), *d_payload_bytes the payload size, *d_status 0 or CVXGPU_ERR_CAPACITY. */
int cvxgpu_encode_blocks_device(cvxgpu_workspace* ws, const float* d_vol, int nx, int ny, int nz_slab,
                                int bx, int by, int bz, float scale, int use_local_rms,
                                const double* d_sumsq_total, long ncells_total,
                                unsigned long long* d_offs, float* d_blkmulfac,
                                void* d_payload, long payload_capacity,
                                long* d_payload_bytes, int* d_status, void* cuda_stream);

int cvxgpu_write_header_device(void* d_stream, int nx, int ny, int nz, int bx, int by, int bz, float scale,
                               int use_local_rms, const double* d_sumsq_total, long ncells_total, void* cuda_stream);

int cvxgpu_offsets_add_device(unsigned long long* d_offs, long n, long base, void* cuda_stream);

/* Decode blocks [blk_first, blk_first + nblk_slab) of the full stream d_stream (nx,ny,nz and
 * block sizes come from its header, which is passed in by value to avoid a device read) into
 * the slab d_vol_slab whose first plane is z = (blk_first / (nbx*nby)) * bz. */
int cvxgpu_decode_blocks_device(cvxgpu_workspace* ws, float* d_vol_slab, int nx, int ny, int nz_slab,
                                int bx, int by, int bz, float glob_mulfac, int use_local_rms,
                                const unsigned long long* d_offs, const float* d_blkmulfac,
                                const void* d_payload, long payload_bytes, void* cuda_stream);

/* ---- z-slab sharding behind one call per direction (one process or host thread per GPU) -----------
 * The reference parallelises ONE volume over its workers (CvxCompress.cpp:318 and :528, OpenMP over the
 * blocks); here the workers are GPUs.  Rank r of `world` owns the block rows [r*nbz/world, (r+1)*nbz/world)
 * (cvxgpu_shard_range): a contiguous z-slab of the volume and a contiguous iBlk range of the tables.
 * A cvxgpu_shard holds the rank's NCCL communicator, workspace and protocol scratch.  NCCL is bound at run
 * time (symbols already in the process, else libnccl.so.2 / $CVX_GPU_NCCL_LIB): pass either the 128-byte
 * ncclUniqueId all ranks share (cvxgpu_shard_unique_id on rank 0, shipped by any means; the shard then
 * calls ncclCommInitRank on the current device and owns the communicator) or an existing `ncclComm_t`
 * as `nccl_comm`.  world == 1 needs neither.
 *
 * cvxgpu_compress_sharded enqueues on cuda_stream, without host synchronisation (CUDA-graph capturable
 * after one eager call):  block-row sums of squares of the slab -> exchange -> fixed-order total ->
 * header + glob_mulfac -> encode the slab's blocks -> exchange of the payload sizes -> exclusive
 * prefix, offsets made global.  "exchange" = a store into every rank's peer-memory mailbox, or ncclAllGather
 * (cvxgpu_shard_uses_peer_memory).  Outputs, all DEVICE memory of this rank:
 *   d_header   32 bytes, identical on every rank (CvxCompress.cpp:284-300)
 *   d_offs     this rank's slice of the offset table, nblk_slab entries, offsets relative to the GLOBAL
 *              payload start (bit 63 = raw);  d_blkmulfac  its slice of the multiplier table (local RMS)
 *   d_payload  this rank's payload bytes; they start at byte d_bases[rank] of the global payload
 *   d_bases    world + 1 int64: exclusive prefix of the ranks' payload sizes, d_bases[world] = total, so
 *              compressed_length = 32 + 8*nnn [+ 4*nnn] + d_bases[world] + 7
 *   d_status   0 or CVXGPU_ERR_CAPACITY (payload_capacity too small for this slab)
 * The concatenation header | tables in rank order | payloads in rank order is byte for byte the stream
 * cvxgpu_compress_device writes for the whole volume, for every `world`.
 * cvxgpu_decompress_sharded decodes the rank's slice back into its slab (no collective; d_payload must be
 * 16-byte aligned).  cvxgpu_decompress_slab_device decodes the rank's iBlk range out of a FULL stream every
 * rank holds (any block order, e.g. one written by the reference with several threads). */
typedef struct cvxgpu_shard cvxgpu_shard;
int cvxgpu_shard_range(int nz, int bz, int rank, int world, int* z0, int* nz_slab, int* block_row0, int* n_block_rows);
int cvxgpu_shard_unique_id(void* id128);
int cvxgpu_shard_create(cvxgpu_shard** sh, const void* id128, void* nccl_comm, int rank, int world);
int cvxgpu_shard_destroy(cvxgpu_shard* sh);
/* How the two per-rank scalars of cvxgpu_compress_sharded travel: 1 = through peer memory (every rank's mailbox is mapped
 * into every process with CUDA IPC at cvxgpu_shard_create; the producing kernel stores into all mailboxes over NVLink, the
 * consuming kernel spins on its own -- no collective launch on the path), 0 = two ncclAllGather (ranks that cannot map
 * each other, e.g. several ranks in one process, or CVX_GPU_SHARD_EXCHANGE=nccl).  All ranks of a shard agree. */
int cvxgpu_shard_uses_peer_memory(const cvxgpu_shard* sh);
int cvxgpu_compress_sharded(cvxgpu_shard* sh, const float* d_slab, int nx, int ny, int nz, int bx, int by, int bz,
                            float scale, int use_local_rms, void* d_header, unsigned long long* d_offs,
                            float* d_blkmulfac, void* d_payload, long payload_capacity, long* d_bases,
                            int* d_status, void* cuda_stream);
int cvxgpu_decompress_sharded(cvxgpu_shard* sh, float* d_slab, int nx, int ny, int nz, int bx, int by, int bz,
                              int use_local_rms, const void* d_header, const unsigned long long* d_offs,
                              const float* d_blkmulfac, const void* d_payload, const long* d_bases, void* cuda_stream);
int cvxgpu_decompress_slab_device(cvxgpu_workspace* ws, float* d_slab, int nx, int ny, int nz, int bx, int by, int bz,
                                  int use_local_rms, int rank, int world, const void* d_stream, long length,
                                  void* cuda_stream);

/* ---- wavelet only: nblocks dense blocks (x fastest), in place (config 2 sweep) ------------- */
int cvxgpu_wavelet_blocks_device(cvxgpu_workspace* ws, float* d_blocks, long nblocks, int bx, int by, int bz,
                                 int inverse, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif
