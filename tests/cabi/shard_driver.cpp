// shard_driver.cpp -- C++ test driver of the sharded C ABI (include/cvx_gpu.h: cvxgpu_compress_sharded /
// cvxgpu_decompress_sharded), no MPI and no torch: the parent starts one process per rank, rank 0 draws the NCCL
// unique id and hands it over through a file, every rank takes GPU `rank`, compresses and decompresses its z-slab of one
// synthetic volume, and compares against the single-GPU calls on the whole volume (cvxgpu_compress_device /
// cvxgpu_decompress_device on the same GPU):
//   * the header, this rank's slice of the offset table (and multiplier table) and its payload bytes must equal the
//     corresponding bytes of the single-GPU stream -- for every world size (G-independent global RMS);
//   * the decoded slab must equal the single-GPU reconstruction bit for bit.
// Test infrastructure: links the product library only.
//
//   shard_driver <world> [nx ny nz bx by bz scale use_local_rms [lifting]]
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <string>
#include <vector>
#include <sys/wait.h>
#include <unistd.h>
#include "cvx_gpu.h"

#define CHECK(call)                                                                    \
    do {                                                                               \
        int rc__ = (call);                                                             \
        if (rc__ != 0) { fprintf(stderr, "rank %d: %s failed (%d): %s\n", g_rank, #call, rc__, cvxgpu_last_error()); exit(2); } \
    } while (0)
#define CUCHECK(call)                                                                  \
    do {                                                                               \
        cudaError_t e__ = (call);                                                      \
        if (e__ != cudaSuccess) { fprintf(stderr, "rank %d: %s: %s\n", g_rank, #call, cudaGetErrorString(e__)); exit(2); } \
    } while (0)
static int g_rank = -1;

static float cell(long i, int nx, int ny, int nz)       // smooth field + counter-hash noise (same on every rank)
{
    const long x = i % nx, y = (i / nx) % ny, z = i / ((long)nx * ny);
    const double r = sqrt((x - 0.5 * (nx - 1)) * (x - 0.5 * (nx - 1)) + (y - 0.5 * (ny - 1)) * (y - 0.5 * (ny - 1)) +
                          (z - 0.5 * (nz - 1)) * (z - 0.5 * (nz - 1)));
    uint64_t h = (uint64_t)i + 0x9E3779B97F4A7C15ull;
    h = (h ^ (h >> 30)) * 0xBF58476D1CE4E5B9ull; h = (h ^ (h >> 27)) * 0x94D049BB133111EBull; h ^= h >> 31;
    return (float)(sin(r / 10.0) + (double)(h >> 40) / 16777216.0 / 100.0);
}

static int run_rank(int rank, int world, const char* idfile, int nx, int ny, int nz, int bx, int by, int bz, float scale, int local)
{
    g_rank = rank;
    int ndev = 0;
    CUCHECK(cudaGetDeviceCount(&ndev));
    if (ndev < world) { fprintf(stderr, "rank %d: %d GPUs visible, %d needed\n", rank, ndev, world); return 3; }
    CUCHECK(cudaSetDevice(rank));
    unsigned char id[128] = {0};
    if (world > 1) {
        if (rank == 0) {
            CHECK(cvxgpu_shard_unique_id(id));
            std::string tmp = std::string(idfile) + ".tmp";
            FILE* f = fopen(tmp.c_str(), "wb");
            fwrite(id, 1, 128, f);
            fclose(f);
            rename(tmp.c_str(), idfile);
        } else {
            FILE* f = nullptr;
            for (int t = 0; t < 600 && !(f = fopen(idfile, "rb")); ++t) usleep(100000);
            if (!f || fread(id, 1, 128, f) != 128) { fprintf(stderr, "rank %d: no NCCL id\n", rank); return 3; }
            fclose(f);
        }
    }
    cvxgpu_shard* sh = nullptr;
    CHECK(cvxgpu_shard_create(&sh, world > 1 ? id : nullptr, nullptr, rank, world));
    int z0, nzs, b0, nb;
    CHECK(cvxgpu_shard_range(nz, bz, rank, world, &z0, &nzs, &b0, &nb));
    const long plane = (long)nx * ny, ncells = plane * nz;
    const long nbx = (nx + bx - 1) / bx, nby = (ny + by - 1) / by, nbz = (nz + bz - 1) / bz, nnn = nbx * nby * nbz;
    const long nblk_slab = (long)nb * nbx * nby, blk_first = (long)b0 * nbx * nby, B = (long)bx * by * bz;
    std::vector<float> vol((size_t)ncells);
    for (long i = 0; i < ncells; ++i) vol[(size_t)i] = cell(i, nx, ny, nz);
    float *d_vol, *d_out, *d_slab_out;
    CUCHECK(cudaMalloc(&d_vol, ncells * 4)); CUCHECK(cudaMalloc(&d_out, ncells * 4)); CUCHECK(cudaMalloc(&d_slab_out, (nzs > 0 ? plane * nzs : 1) * 4));
    CUCHECK(cudaMemcpy(d_vol, vol.data(), ncells * 4, cudaMemcpyHostToDevice));
    cudaStream_t st;
    CUCHECK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    // ---- single-GPU reference of this test: the whole volume through the whole-volume calls ----
    const long cap = cvxgpu_max_compressed_bytes(nx, ny, nz, bx, by, bz, local);
    unsigned char* d_full;
    CUCHECK(cudaMalloc(&d_full, cap + 64));
    long len = 0; float ratio = 0;
    CHECK(cvxgpu_compress_device(nullptr, d_vol, nx, ny, nz, bx, by, bz, scale, local, d_full, cap, &len, &ratio, st));
    CHECK(cvxgpu_decompress_device(nullptr, d_out, nx, ny, nz, d_full, len, st));
    std::vector<unsigned char> full((size_t)len);
    std::vector<float> rec((size_t)ncells);
    CUCHECK(cudaMemcpyAsync(full.data(), d_full, len - 7, cudaMemcpyDeviceToHost, st));
    CUCHECK(cudaMemcpyAsync(rec.data(), d_out, ncells * 4, cudaMemcpyDeviceToHost, st));
    CUCHECK(cudaStreamSynchronize(st));
    // ---- sharded ----
    unsigned char *d_hdr, *d_pay; unsigned long long* d_offs; float* d_bmf; long* d_bases; int* d_status;
    const long pcap = 4 * B * (nblk_slab > 0 ? nblk_slab : 1) + 64;
    CUCHECK(cudaMalloc(&d_hdr, 32)); CUCHECK(cudaMalloc(&d_pay, pcap)); CUCHECK(cudaMalloc(&d_offs, 8 * (nblk_slab + 1)));
    CUCHECK(cudaMalloc(&d_bmf, 4 * (nblk_slab + 1))); CUCHECK(cudaMalloc(&d_bases, 8 * (world + 1))); CUCHECK(cudaMalloc(&d_status, 4));
    const float* d_slab = d_vol + plane * z0;
    int bad = 0;
    for (int rep = 0; rep < 2; ++rep) {     // the second round re-uses every buffer (steady state)
        CHECK(cvxgpu_compress_sharded(sh, d_slab, nx, ny, nz, bx, by, bz, scale, local, d_hdr, d_offs, local ? d_bmf : nullptr, d_pay, pcap,
                                      d_bases, d_status, st));
        CHECK(cvxgpu_decompress_sharded(sh, d_slab_out, nx, ny, nz, bx, by, bz, local, d_hdr, d_offs, local ? d_bmf : nullptr, d_pay, d_bases, st));
        unsigned char hdr[32]; std::vector<long> bases((size_t)world + 1); int status = -99;
        CUCHECK(cudaMemcpyAsync(hdr, d_hdr, 32, cudaMemcpyDeviceToHost, st));
        CUCHECK(cudaMemcpyAsync(bases.data(), d_bases, 8 * (world + 1), cudaMemcpyDeviceToHost, st));
        CUCHECK(cudaMemcpyAsync(&status, d_status, 4, cudaMemcpyDeviceToHost, st));
        CUCHECK(cudaStreamSynchronize(st));
        const long tables = 8 * nnn + (local ? 4 * nnn : 0);
        const long total = 32 + tables + bases[(size_t)world] + 7;
        if (status != 0) { fprintf(stderr, "rank %d: status %d\n", rank, status); ++bad; }
        if (total != len) { fprintf(stderr, "rank %d: sharded length %ld, single-GPU length %ld\n", rank, total, len); ++bad; }
        if (memcmp(hdr, full.data(), 32)) { fprintf(stderr, "rank %d: header differs (glob_mulfac depends on the sharding?)\n", rank); ++bad; }
        if (nblk_slab > 0 && !bad) {
            std::vector<unsigned long long> offs((size_t)nblk_slab);
            CUCHECK(cudaMemcpy(offs.data(), d_offs, 8 * nblk_slab, cudaMemcpyDeviceToHost));
            if (memcmp(offs.data(), full.data() + 32 + 8 * blk_first, 8 * nblk_slab)) { fprintf(stderr, "rank %d: offset table slice differs\n", rank); ++bad; }
            if (local) {
                std::vector<float> bmf((size_t)nblk_slab);
                CUCHECK(cudaMemcpy(bmf.data(), d_bmf, 4 * nblk_slab, cudaMemcpyDeviceToHost));
                if (memcmp(bmf.data(), full.data() + 32 + 8 * nnn + 4 * blk_first, 4 * nblk_slab)) { fprintf(stderr, "rank %d: multiplier table slice differs\n", rank); ++bad; }
            }
            const long mine = bases[(size_t)rank + 1] - bases[(size_t)rank];
            std::vector<unsigned char> pay((size_t)(mine > 0 ? mine : 1));
            CUCHECK(cudaMemcpy(pay.data(), d_pay, mine, cudaMemcpyDeviceToHost));
            const long lim = (32 + tables + bases[(size_t)rank] + mine <= len - 7) ? mine : 0;
            if (lim != mine || memcmp(pay.data(), full.data() + 32 + tables + bases[(size_t)rank], mine)) { fprintf(stderr, "rank %d: payload shard differs\n", rank); ++bad; }
            std::vector<float> slab((size_t)(plane * nzs));
            CUCHECK(cudaMemcpy(slab.data(), d_slab_out, plane * nzs * 4, cudaMemcpyDeviceToHost));
            if (memcmp(slab.data(), rec.data() + plane * z0, plane * nzs * 4)) { fprintf(stderr, "rank %d: decoded slab differs from the single-GPU reconstruction\n", rank); ++bad; }
        }
    }
    // every rank decodes its slab out of the FULL stream too (what config 5 does with a reference-written stream)
    if (nzs > 0) {
        CUCHECK(cudaMemsetAsync(d_slab_out, 0xFF, plane * nzs * 4, st));
        CHECK(cvxgpu_decompress_slab_device(nullptr, d_slab_out, nx, ny, nz, bx, by, bz, local, rank, world, d_full, len, st));
        std::vector<float> slab((size_t)(plane * nzs));
        CUCHECK(cudaMemcpyAsync(slab.data(), d_slab_out, plane * nzs * 4, cudaMemcpyDeviceToHost, st));
        CUCHECK(cudaStreamSynchronize(st));
        if (memcmp(slab.data(), rec.data() + plane * z0, plane * nzs * 4)) { fprintf(stderr, "rank %d: slab decoded from the full stream differs\n", rank); ++bad; }
    }
    double num = 0, den = 0;
    for (long i = 0; i < ncells; ++i) { const double d = (double)vol[(size_t)i] - rec[(size_t)i]; num += d * d; den += (double)vol[(size_t)i] * vol[(size_t)i]; }
    printf("rank %d/%d: slab z=[%d,%d) %ld blocks, stream %ld bytes (ratio %.2f), round-trip error %.3e: %s\n", rank, world, z0, z0 + nzs, nblk_slab,
           len, ratio, sqrt(num / den), bad ? "MISMATCH" : "OK");
    CHECK(cvxgpu_shard_destroy(sh));
    return bad ? 1 : 0;
}

int main(int argc, char** argv)
{
    const int world = argc > 1 ? atoi(argv[1]) : 1;
    int nx = 96, ny = 80, nz = 200, bx = 32, by = 32, bz = 32, local = 0;
    float scale = 1e-2f;
    if (argc > 8) { nx = atoi(argv[2]); ny = atoi(argv[3]); nz = atoi(argv[4]); bx = atoi(argv[5]); by = atoi(argv[6]); bz = atoi(argv[7]); scale = (float)atof(argv[8]); }
    if (argc > 9) local = atoi(argv[9]);
    if (argc > 10 && !strcmp(argv[10], "lifting")) cvxgpu_set_transform(CVXGPU_TRANSFORM_LIFTING);
    if (world < 1 || world > 64) { fprintf(stderr, "usage: shard_driver <world> [nx ny nz bx by bz scale use_local_rms [lifting]]\n"); return 2; }
    char idfile[256];
    snprintf(idfile, sizeof(idfile), "/tmp/cvx_shard_id_%d", (int)getpid());
    unlink(idfile);
    if (world == 1) return run_rank(0, 1, idfile, nx, ny, nz, bx, by, bz, scale, local);
    std::vector<pid_t> kids;
    for (int r = 0; r < world; ++r) {          // fork BEFORE any CUDA call: every rank initialises its own context
        pid_t p = fork();
        if (p == 0) _exit(run_rank(r, world, idfile, nx, ny, nz, bx, by, bz, scale, local));
        kids.push_back(p);
    }
    int worst = 0;
    for (pid_t p : kids) {
        int stt = 0;
        waitpid(p, &stt, 0);
        const int rc = WIFEXITED(stt) ? WEXITSTATUS(stt) : 99;
        if (rc > worst) worst = rc;
    }
    unlink(idfile);
    printf("shard_driver world=%d: %s\n", world, worst ? "FAILED" : "ALL OK");
    return worst;
}
