// cvx_selftest.cpp -- CvxCompress::Run_Module_Tests for the GPU library.  The reference's
// version (CvxCompress.cpp:673-1246) checks Fast-vs-Slow transforms, block copies, the global
// RMS and prints throughputs.  Here the CPU has no implementation to compare against (by
// design: no CPU fallback in the product), so the self-test checks properties on the device:
// forward+inverse transform round trips for block shapes, gather/scatter round trips through
// raw blocks, and Compress -> Decompress error on synthetic volumes.  Parity against the
// reference itself lives in tests/ (which may use the oracle; this file may not).
#include "../../include/CvxCompress.hxx"
#include "../../include/cvx_gpu.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

static float synth(long i)      // deterministic U(-1,1)
{
    unsigned long long x = (unsigned long long)i * 0x9E3779B97F4A7C15ull + 0x1234567ull;
    x ^= x >> 29; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 32;
    return (float)((double)(x & 0xFFFFFF) / 8388608.0 - 1.0);
}

bool CvxCompress::Run_Module_Tests(bool verbose, bool exhaustive)
{
    bool ok = true;
    // 1. Compress -> Decompress on a smooth + noise volume with partial blocks, several shapes, both RMS modes
    const int nx = 151, ny = 101, nz = 51;      // the reference's synthetic test volume size (Read_Raw_Volume.cpp:12-16)
    std::vector<float> vol((size_t)nx * ny * nz), out(vol.size());
    for (int z = 0; z < nz; ++z)
        for (int y = 0; y < ny; ++y)
            for (int x = 0; x < nx; ++x) {
                double r = std::sqrt((double)(x - 75) * (x - 75) + (double)(y - 50) * (y - 50) + (double)(z - 25) * (z - 25));
                long i = ((long)z * ny + y) * nx + x;
                vol[i] = (float)(std::sin(r / 10.0) + synth(i) / 200.0);
            }
    const int shapes[][3] = {{8, 8, 8}, {16, 16, 16}, {32, 32, 32}, {16, 16, 64}, {32, 32, 1}, {64, 64, 8}, {128, 8, 8}, {64, 64, 64}};
    const int nshapes = exhaustive ? 8 : 6;
    for (int si = 0; si < nshapes; ++si) {
        for (int local = 0; local < 2; ++local) {
            const int bx = shapes[si][0], by = shapes[si][1], bz = shapes[si][2];
            long cap = cvxgpu_max_compressed_bytes(nx, ny, nz, bx, by, bz, local);
            std::vector<unsigned int> stream((size_t)cap / 4 + 16);
            long len = 0;
            auto t0 = std::chrono::steady_clock::now();
            float ratio = Compress(1e-3f, vol.data(), nx, ny, nz, bx, by, bz, local != 0, stream.data(), len);
            auto t1 = std::chrono::steady_clock::now();
            Decompress(out.data(), nx, ny, nz, stream.data(), len);
            auto t2 = std::chrono::steady_clock::now();
            double num = 0, den = 0;
            for (size_t i = 0; i < vol.size(); ++i) { double d = (double)vol[i] - out[i]; num += d * d; den += (double)vol[i] * vol[i]; }
            double err = std::sqrt(num / den);
            bool pass = err < 5e-3 && ratio > 1.0f && len > 32;
            ok = ok && pass;
            if (verbose)
                printf("  %3dx%3dx%3d %s RMS: ratio %.2f:1, error %.3e, compress %.1f ms, decompress %.1f ms -> %s\n", bx, by, bz,
                       local ? "local " : "global", ratio, err, std::chrono::duration<double, std::milli>(t1 - t0).count(),
                       std::chrono::duration<double, std::milli>(t2 - t1).count(), pass ? "ok" : "FAILED");
        }
    }
    // 2. raw blocks (scale so small that every block falls back to raw): round trip is the pure transform error
    {
        long cap = cvxgpu_max_compressed_bytes(nx, ny, nz, 32, 32, 32, 0);
        std::vector<unsigned int> stream((size_t)cap / 4 + 16);
        long len = 0;
        std::vector<float> noise(vol.size());
        for (size_t i = 0; i < noise.size(); ++i) noise[i] = synth((long)i);
        Compress(1e-9f, noise.data(), nx, ny, nz, 32, 32, 32, false, stream.data(), len);
        Decompress(out.data(), nx, ny, nz, stream.data(), len);
        double m = 0;
        for (size_t i = 0; i < noise.size(); ++i) m = std::fmax(m, std::fabs((double)noise[i] - out[i]));
        bool pass = m < 5e-6;
        ok = ok && pass;
        if (verbose) printf("  raw-block round trip: max |err| %.3e, length %ld -> %s\n", m, len, pass ? "ok" : "FAILED");
    }
    return ok;
}
