// cvxfile: raw fp32 volume <-> CvxCompress stream on disk, with the error report of the reference's file driver
// (Test_Compression.cpp:53-230: read volume, Compress, Decompress, "Error is rms(diff)/rms(orig)", write the stream).
// SURVEY.md 8(f) row 4.  Differences from the reference driver, on purpose:
//   * the volume really is read from the file (the reference's Read_Raw_Volume.cpp:12-43 ignores the file name and
//     synthesises a 151x101x51 volume), so nx ny nz are arguments; cells are headerless little-endian fp32, x fastest;
//   * a stream read back from disk is walked by cvxgpu_validate_stream before it is decoded;
//   * `info` needs no GPU.
// Links libcvxcompress.so only (the CvxCompress class and the C-ABI); no oracle, no CPU codec.
#include "../../include/CvxCompress.hxx"
#include "../../include/cvx_gpu.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

static int usage(const char* argv0)
{
    fprintf(stderr,
            "Usage : %s compress   <raw-volume-file> <nx> <ny> <nz> <scale> <bx> <by> <bz> <compressed-volume-file> [num-loops] [--local]\n"
            "        %s decompress <compressed-volume-file> <raw-volume-file>\n"
            "        %s info       <compressed-volume-file> [--shallow]\n", argv0, argv0, argv0);
    return 2;
}

static bool read_file(const char* name, std::vector<unsigned char>& buf, size_t pad)
{
    FILE* fp = fopen(name, "rb");
    if (!fp) { fprintf(stderr, "cannot open %s\n", name); return false; }
    fseek(fp, 0, SEEK_END);
    const long n = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    buf.assign((size_t)n + pad, 0);
    const bool ok = n >= 0 && fread(buf.data(), 1, (size_t)n, fp) == (size_t)n;
    fclose(fp);
    buf.resize((size_t)n + pad);
    if (!ok) fprintf(stderr, "short read from %s\n", name);
    return ok;
}

static bool write_file(const char* name, const void* p, size_t n)
{
    FILE* fp = fopen(name, "wb");
    if (!fp) { fprintf(stderr, "cannot create %s\n", name); return false; }
    const bool ok = fwrite(p, 1, n, fp) == n;
    fclose(fp);
    if (!ok) fprintf(stderr, "short write to %s\n", name);
    return ok;
}

static double rms(const float* a, const float* b, size_t n)   // rms(a) or rms(a - b); Test_Compression.cpp:41-52,121-140
{
    double acc = 0.0;
    for (size_t i = 0; i < n; ++i) {
        const double d = b ? (double)a[i] - (double)b[i] : (double)a[i];
        acc += d * d;
    }
    return std::sqrt(acc / (double)n);
}

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

static int print_info(const char* name, const std::vector<unsigned char>& s, long length, int level)
{
    cvxgpu_stream_info I;
    const int rc = cvxgpu_validate_stream(s.data(), length, level, &I);
    if (rc == CVXGPU_OK || I.nblocks > 0) {
        printf("%s: %d x %d x %d cells, %d x %d x %d blocks, %s RMS, global multiplier %.6e\n", name, I.nx, I.ny, I.nz,
               I.bx, I.by, I.bz, I.use_local_rms ? "local" : "global", (double)I.glob_mulfac);
        printf("  %ld blocks (%ld raw), tables %ld bytes, payload %ld bytes, %s block order\n", I.nblocks, I.raw_blocks,
               I.table_bytes, I.payload_bytes, I.natural_order ? "natural" : "permuted");
        printf("  compression ratio %.2f:1\n", 4.0 * I.nx * I.ny * I.nz / (double)length);
    }
    if (rc != CVXGPU_OK) {
        printf("  INVALID: %s\n", cvxgpu_last_error());
        return 1;
    }
    if (level > 0)
        printf("  walked %ld tokens; block payloads %ld .. %ld bytes, %ld of %ld payload bytes used\n", I.tokens,
               I.min_block_bytes, I.max_block_bytes, I.used_payload_bytes, I.payload_bytes);
    printf("  stream is well formed (level %d)\n", level);
    return 0;
}

int main(int argc, char* argv[])
{
    if (argc < 3) return usage(argv[0]);
    const std::string cmd = argv[1];

    if (cmd == "info") {
        std::vector<unsigned char> s;
        if (!read_file(argv[2], s, 0)) return 1;
        const int level = (argc > 3 && !strcmp(argv[3], "--shallow")) ? 0 : 1;
        return print_info(argv[2], s, (long)s.size(), level);
    }

    if (cmd == "decompress") {
        if (argc < 4) return usage(argv[0]);
        std::vector<unsigned char> s;
        if (!read_file(argv[2], s, 8)) return 1;
        const long length = (long)s.size() - 8;
        cvxgpu_stream_info I;
        if (cvxgpu_validate_stream(s.data(), length, 1, &I) != CVXGPU_OK) {
            fprintf(stderr, "%s is not a well-formed stream: %s\n", argv[2], cvxgpu_last_error());
            return 1;
        }
        if (cvxgpu_device_count() <= 0) { fprintf(stderr, "no CUDA device: %s\n", cvxgpu_last_error()); return 1; }
        CvxCompress c;
        int nx, ny, nz;
        const double t0 = now();
        float* vol = c.Decompress(nx, ny, nz, reinterpret_cast<unsigned int*>(s.data()), length);
        const double t1 = now();
        if (!vol) { fprintf(stderr, "Decompress failed: %s\n", cvxgpu_last_error()); return 1; }
        const size_t n = (size_t)nx * ny * nz;
        printf("Decompression throughput was %.0f MC/s (%d x %d x %d)\n", n / (t1 - t0) * 1e-6, nx, ny, nz);
        printf("Writing raw volume to %s...\n", argv[3]);
        const bool ok = write_file(argv[3], vol, n * 4);
        free(vol);
        return ok ? 0 : 1;
    }

    if (cmd == "compress") {
        if (argc < 11) return usage(argv[0]);
        const int nx = atoi(argv[3]), ny = atoi(argv[4]), nz = atoi(argv[5]);
        const float scale = (float)atof(argv[6]);
        const int bx = atoi(argv[7]), by = atoi(argv[8]), bz = atoi(argv[9]);
        int loops = 1;
        bool local = false;
        for (int i = 11; i < argc; ++i) {
            if (!strcmp(argv[i], "--local")) local = true;
            else loops = atoi(argv[i]) > 0 ? atoi(argv[i]) : 1;
        }
        printf("Using %s RMS.\n", local ? "local" : "global");     // Test_Compression.cpp:62
        printf("Scale is %.3e\n", (double)scale);
        CvxCompress c;
        if (nx <= 0 || ny <= 0 || nz <= 0 || !c.Is_Valid_Block_Size(bx, by, bz)) {
            fprintf(stderr, "bad volume %d x %d x %d or block size %d x %d x %d\n", nx, ny, nz, bx, by, bz);
            return 2;
        }
        if (cvxgpu_device_count() <= 0) { fprintf(stderr, "no CUDA device: %s\n", cvxgpu_last_error()); return 1; }
        const size_t n = (size_t)nx * ny * nz;
        printf("Reading raw volume from file %s...\n", argv[2]);
        std::vector<unsigned char> raw;
        if (!read_file(argv[2], raw, 0)) return 1;
        if (raw.size() != n * 4) {
            fprintf(stderr, "%s holds %zu bytes, %d x %d x %d fp32 cells need %zu\n", argv[2], raw.size(), nx, ny, nz, n * 4);
            return 1;
        }
        float* vol = reinterpret_cast<float*>(raw.data());
        const long cap = cvxgpu_max_compressed_bytes(nx, ny, nz, bx, by, bz, local ? 1 : 0);
        std::vector<unsigned long long> comp((size_t)cap / 8 + 2);
        std::vector<float> rec(n);
        long length = 0;
        float ratio = 0.f;
        for (int l = 0; l < loops; ++l) {
            printf("Compressing.\n");
            double t0 = now();
            ratio = c.Compress(scale, vol, nx, ny, nz, bx, by, bz, local, reinterpret_cast<unsigned int*>(comp.data()), length);
            double t1 = now();
            if (!(ratio > 0.f) || length <= 0) { fprintf(stderr, "Compress failed: %s\n", cvxgpu_last_error()); return 1; }
            printf("Compression throughput was %.0f MC/s - Compression ratio is %.2f:1\n", n / (t1 - t0) * 1e-6, (double)ratio);
            printf("Decompressing.\n");
            t0 = now();
            c.Decompress(rec.data(), nx, ny, nz, reinterpret_cast<unsigned int*>(comp.data()), length);
            t1 = now();
            printf("Decompression throughput was %.0f MC/s\n", n / (t1 - t0) * 1e-6);
        }
        printf("Generating error volume.\n");
        const double r0 = rms(vol, nullptr, n), rd = rms(vol, rec.data(), n);
        printf("Error is %.3e\n", r0 > 0.0 ? rd / r0 : rd);
        // the 7 slack bytes are never written by the encoder (CvxCompress.cpp:421-422); put zeros on disk
        memset(reinterpret_cast<unsigned char*>(comp.data()) + length - 7, 0, 7);
        printf("Writing compressed volume to %s...\n", argv[10]);
        return write_file(argv[10], comp.data(), (size_t)length) ? 0 : 1;
    }
    return usage(argv[0]);
}
