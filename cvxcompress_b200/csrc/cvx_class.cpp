// cvx_class.cpp -- the CvxCompress class and the cvx_* C functions (include/CvxCompress.hxx) on
// top of the C ABI of include/cvx_gpu.h.  Mirrors the overload funnel of the reference:
// CvxCompress.cpp:173-228 (Compress), :429-461 (Decompress), :1249-1322 (C wrappers).
#include "../../include/CvxCompress.hxx"
#include "../../include/cvx_gpu.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

[[noreturn]] static void die(const char* what)
{
    // the reference asserts / exits on these paths (CvxCompress.cpp:246-248, :265, :478-488)
    fprintf(stderr, "CvxCompress (B200): %s failed: %s\n", what, cvxgpu_last_error());
    abort();
}

CvxCompress::CvxCompress() {}
CvxCompress::~CvxCompress() {}

float CvxCompress::Compress(float scale, float* vol, int nx, int ny, int nz, int bx, int by, int bz,
                            bool use_local_RMS, unsigned int* compressed, int /*num_threads*/, long& compressed_length)
{
    float ratio = 0.0f;
    long len = 0;
    if (cvxgpu_compress(vol, nx, ny, nz, bx, by, bz, scale, use_local_RMS ? 1 : 0, compressed, &len, &ratio) != CVXGPU_OK) die("Compress");
    compressed_length = len;
    return ratio;
}

float CvxCompress::Compress(float scale, float* vol, int nx, int ny, int nz, int bx, int by, int bz,
                            bool use_local_RMS, unsigned int* compressed, long& compressed_length)
{
    return Compress(scale, vol, nx, ny, nz, bx, by, bz, use_local_RMS, compressed, 0, compressed_length);
}

float CvxCompress::Compress(float scale, float* vol, int nx, int ny, int nz, int bx, int by, int bz,
                            unsigned int* compressed, long& compressed_length)
{
    return Compress(scale, vol, nx, ny, nz, bx, by, bz, false, compressed, 0, compressed_length);
}

float CvxCompress::Compress(float scale, float* vol, int nx, int ny, int nz, int bx, int by, int bz,
                            unsigned int* compressed, int num_threads, long& compressed_length)
{
    return Compress(scale, vol, nx, ny, nz, bx, by, bz, false, compressed, num_threads, compressed_length);
}

float* CvxCompress::Decompress(int& nx, int& ny, int& nz, unsigned int* compressed, long compressed_length)
{
    // header words 0..2 (CvxCompress.cpp:436-438); the stream may be a device pointer
    int h[3];
    {
        int hx, hy, hz;
        unsigned char hdr[32];
        extern int cvx_internal_fetch_header(const void*, unsigned char*);
        if (cvx_internal_fetch_header(compressed, hdr) != 0) die("Decompress (header)");
        if (cvxgpu_parse_header(hdr, 32, &hx, &hy, &hz, nullptr, nullptr, nullptr, nullptr, nullptr) != CVXGPU_OK) die("Decompress (header)");
        h[0] = hx; h[1] = hy; h[2] = hz;
    }
    nx = h[0]; ny = h[1]; nz = h[2];
    float* vol = nullptr;
    if (posix_memalign(reinterpret_cast<void**>(&vol), 64, sizeof(float) * (size_t)nx * (size_t)ny * (size_t)nz) != 0 || !vol) {
        fprintf(stderr, "CvxCompress (B200): out of host memory for a %d x %d x %d volume\n", nx, ny, nz);
        abort();
    }
    Decompress(vol, nx, ny, nz, compressed, 0, compressed_length);
    return vol;
}

void CvxCompress::Decompress(float* vol, int nx, int ny, int nz, unsigned int* compressed, int /*num_threads*/, long compressed_length)
{
    if (cvxgpu_decompress(vol, nx, ny, nz, compressed, compressed_length) != CVXGPU_OK) die("Decompress");
}

void CvxCompress::Decompress(float* vol, int nx, int ny, int nz, unsigned int* compressed, long compressed_length)
{
    Decompress(vol, nx, ny, nz, compressed, 0, compressed_length);
}

bool CvxCompress::Is_Valid_Block_Size(int bx, int by, int bz) { return cvxgpu_is_valid_block_size(bx, by, bz) != 0; }

// ---- C wrappers (CvxCompress.cpp:1249-1322): a temporary object per call ----------------------
extern "C" float cvx_compress(float scale, float* vol, int nx, int ny, int nz, int bx, int by, int bz,
                              unsigned int* compressed, long* compressed_length)
{
    CvxCompress c;
    long len = 0;
    float r = c.Compress(scale, vol, nx, ny, nz, bx, by, bz, false, compressed, len);
    *compressed_length = len;
    return r;
}

extern "C" float* cvx_decompress_outofplace(int* nx, int* ny, int* nz, unsigned int* compressed, long compressed_length)
{
    CvxCompress c;
    int x = 0, y = 0, z = 0;
    float* v = c.Decompress(x, y, z, compressed, compressed_length);
    *nx = x; *ny = y; *nz = z;
    return v;
}

extern "C" void cvx_decompress_inplace(float* vol, int nx, int ny, int nz, unsigned int* compressed, long compressed_length)
{
    CvxCompress c;
    c.Decompress(vol, nx, ny, nz, compressed, compressed_length);
}

extern "C" float cvx_compress_th(float scale, float* vol, int nx, int ny, int nz, int bx, int by, int bz,
                                 bool use_local_RMS, unsigned int* compressed, int num_threads, long* compressed_length)
{
    CvxCompress c;
    long len = 0;
    float r = c.Compress(scale, vol, nx, ny, nz, bx, by, bz, use_local_RMS, compressed, num_threads, len);
    *compressed_length = len;
    return r;
}

extern "C" void cvx_decompress_inplace_th(float* vol, int nx, int ny, int nz, unsigned int* compressed,
                                          int num_threads, long compressed_length)
{
    CvxCompress c;
    c.Decompress(vol, nx, ny, nz, compressed, num_threads, compressed_length);
}
