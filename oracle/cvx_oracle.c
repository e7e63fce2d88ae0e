/*
 * cvx_oracle.c -- plain-C restatement of the CvxCompress hot path (see cvx_oracle.h).
 * TEST INFRASTRUCTURE ONLY; parity PINNED against oracle/_ref and the README goldens.
 * Build: oracle/Makefile (gcc -O2 -ffp-contract=off: the shipped reference has no FMA).
 */
#include "cvx_oracle.h"
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <limits.h>

/* ---- filter taps: Wavelet_Transform_Slow.cpp:21-30 (analysis), :136-145 (synthesis) ---- */
static const float AL[5] = { 8.526986790094000e-001f, 3.774028556126500e-001f, -1.106244044184200e-001f,
                             -2.384946501938001e-002f, 3.782845550699501e-002f };
static const float AH[4] = { 7.884856164056601e-001f, -4.180922732222101e-001f, -4.068941760955800e-002f,
                             6.453888262893799e-002f };
static const float SL[4] = { 7.884856164056601e-001f, 4.180922732222101e-001f, -4.068941760955800e-002f,
                             -6.453888262893799e-002f };
static const float SH[5] = { 8.526986790094000e-001f, -3.774028556126500e-001f, -1.106244044184200e-001f,
                             2.384946501938001e-002f, 3.782845550699501e-002f };

/* Wavelet_Transform_Slow.cpp:59-67 -- whole-sample mirror about 0 and n-1, applied twice */
int cvxo_mirr(int v, int n)
{
    v = v < 0 ? -v : v;
    v = v >= n ? 2 * n - 2 - v : v;
    v = v < 0 ? -v : v;
    v = v >= n ? 2 * n - 2 - v : v;
    return v;
}
/* Wavelet_Transform_Slow.cpp:178-188 */
int cvxo_mirr_sl(int v, int nl)
{
    for (int r = 0; r < 3; ++r) {
        v = v < 0 ? -v : v;
        v = v >= nl ? 2 * nl - 1 - v : v;
    }
    return v;
}
/* Wavelet_Transform_Slow.cpp:189-199 */
int cvxo_mirr_sh(int v, int nl, int nh)
{
    v -= nl;
    for (int r = 0; r < 3; ++r) {
        v = v < 0 ? -v - 1 : v;
        v = v >= nh ? 2 * nh - 2 - v : v;
    }
    return nl + v;
}

/* One analysis level on t[0..n) -> out[0..n).  FAST association: generated code,
 * Wavelet_Transform_Slow.cpp:416-448 (low) and :465-490 (high): acc = a4*p4; acc += a3*p3; ... acc += a0*t0.
 * SLOW association: Wavelet_Transform_Slow.cpp:104-109, :119-123. */
static void ds79_level(const float* t, float* out, int ostride, int n, int assoc)
{
    int nh = n / 2, nl = n - nh;
#define T(i) t[cvxo_mirr((i), n)]
    for (int ix = 0; ix < nl; ++ix) {
        int i0 = 2 * ix;
        float p1 = T(i0 - 1) + T(i0 + 1), p2 = T(i0 - 2) + T(i0 + 2);
        float p3 = T(i0 - 3) + T(i0 + 3), p4 = T(i0 - 4) + T(i0 + 4);
        float r;
        if (assoc == CVXO_ASSOC_FAST) {
            float acc = AL[4] * p4;
            acc = acc + AL[3] * p3;
            acc = acc + AL[2] * p2;
            acc = acc + AL[1] * p1;
            acc = acc + AL[0] * t[i0];
            r = acc;
        } else {
            float acc1 = AL[4] * p4; acc1 += AL[1] * p1; acc1 += AL[0] * t[i0];
            float acc2 = AL[3] * p3; acc2 += AL[2] * p2;
            r = acc1 + acc2;
        }
        out[ix * ostride] = r;
    }
    for (int ix = 0; ix < nh; ++ix) {
        int i0 = 2 * ix + 1;
        float p1 = T(i0 - 1) + T(i0 + 1), p2 = T(i0 - 2) + T(i0 + 2), p3 = T(i0 - 3) + T(i0 + 3);
        float r;
        if (assoc == CVXO_ASSOC_FAST) {
            float acc = AH[3] * p3;
            acc = acc + AH[2] * p2;
            acc = acc + AH[1] * p1;
            acc = acc + AH[0] * t[i0];
            r = acc;
        } else {
            float acc1 = AH[3] * p3; acc1 += AH[0] * t[i0];
            float acc2 = AH[2] * p2; acc2 += AH[1] * p1;
            r = acc1 + acc2;
        }
        out[(nl + ix) * ostride] = r;
    }
#undef T
}

/* One synthesis level.  FAST: Wavelet_Transform_Slow.cpp:613-638 (even), :664-696 (odd).
 * SLOW: Wavelet_Transform_Slow.cpp:233-247 (left-to-right sum). */
static void us79_level(const float* t, float* out, int ostride, int n, int assoc)
{
    int nh = n / 2, nl = n - nh;
#define L(i) t[cvxo_mirr_sl((i), nl)]
#define H(i) t[cvxo_mirr_sh((i), nl, nh)]
    for (int k = 0; k < nl; ++k) {
        float pl2 = L(k - 1) + L(k + 1);
        float ph1 = H(nl + k - 1) + H(nl + k);
        float ph3 = H(nl + k - 2) + H(nl + k + 1);
        float r;
        if (assoc == CVXO_ASSOC_FAST) {
            float acc = SH[3] * ph3;
            acc = acc + SL[2] * pl2;
            acc = acc + SH[1] * ph1;
            acc = acc + SL[0] * t[k];
            r = acc;
        } else {
            r = SL[0] * t[k] + SL[2] * pl2 + SH[1] * ph1 + SH[3] * ph3;
        }
        out[(2 * k) * ostride] = r;
    }
    for (int k = 0; k < nh; ++k) {
        float pl1 = L(k) + L(k + 1);
        float pl3 = L(k - 1) + L(k + 2);
        float ph2 = H(nl + k - 1) + H(nl + k + 1);
        float ph4 = H(nl + k - 2) + H(nl + k + 2);
        float r;
        if (assoc == CVXO_ASSOC_FAST) {
            float acc = SH[4] * ph4;
            acc = acc + SL[3] * pl3;
            acc = acc + SH[2] * ph2;
            acc = acc + SL[1] * pl1;
            acc = acc + SH[0] * t[nl + k];
            r = acc;
        } else {
            r = SL[1] * pl1 + SL[3] * pl3 + SH[0] * t[nl + k] + SH[2] * ph2 + SH[4] * ph4;
        }
        out[(2 * k + 1) * ostride] = r;
    }
#undef L
#undef H
}

/* Wavelet_Transform_Slow.cpp:71-134: levels dim, dim/2, ..., 2 on the low half, in place */
void cvxo_ds79_line(float* p, int stride, int dim, int assoc)
{
    float t[256];
    for (int n = dim; n >= 2; n = n - n / 2) {
        for (int i = 0; i < n; ++i) t[i] = p[i * stride];
        ds79_level(t, p, stride, n, assoc);
    }
}
/* Wavelet_Transform_Slow.cpp:201-259: levels 2, ..., dim */
void cvxo_us79_line(float* p, int stride, int dim, int assoc)
{
    float t[256];
    int lv[16], nlv = 0;
    for (int n = dim; n >= 2; n = n - n / 2) lv[nlv++] = n;
    for (int li = nlv - 1; li >= 0; --li) {
        int n = lv[li];
        for (int i = 0; i < n; ++i) t[i] = p[i * stride];
        us79_level(t, p, stride, n, assoc);
    }
}

/* Wavelet_Transform_Fast.cpp:13-146 (== Wavelet_Transform_Slow.cpp:261-280 on a dense block) */
void cvxo_wavelet_forward(float* b, int bx, int by, int bz, int assoc)
{
    for (int iz = 0; iz < bz; ++iz) {
        if (bx > 1) for (int iy = 0; iy < by; ++iy) cvxo_ds79_line(b + ((long)iz * by + iy) * bx, 1, bx, assoc);
        if (by > 1) for (int ix = 0; ix < bx; ++ix) cvxo_ds79_line(b + (long)iz * by * bx + ix, bx, by, assoc);
    }
    if (bz > 1)
        for (int iy = 0; iy < by; ++iy)
            for (int ix = 0; ix < bx; ++ix) cvxo_ds79_line(b + (long)iy * bx + ix, bx * by, bz, assoc);
}
/* Wavelet_Transform_Fast.cpp:148-279 */
void cvxo_wavelet_inverse(float* b, int bx, int by, int bz, int assoc)
{
    for (int iz = 0; iz < bz; ++iz) {
        if (bx > 1) for (int iy = 0; iy < by; ++iy) cvxo_us79_line(b + ((long)iz * by + iy) * bx, 1, bx, assoc);
        if (by > 1) for (int ix = 0; ix < bx; ++ix) cvxo_us79_line(b + (long)iz * by * bx + ix, bx, by, assoc);
    }
    if (bz > 1)
        for (int iy = 0; iy < by; ++iy)
            for (int ix = 0; ix < bx; ++ix) cvxo_us79_line(b + (long)iy * bx + ix, bx * by, bz, assoc);
}

/* Block_Copy.cpp:21-116: gather with zero padding beyond the volume in x, y, z */
void cvxo_copy_to_block(const float* vol, int x0, int y0, int z0, int nx, int ny, int nz,
                        float* blk, int bx, int by, int bz)
{
    for (int z = 0; z < bz; ++z)
        for (int y = 0; y < by; ++y)
            for (int x = 0; x < bx; ++x) {
                int gx = x0 + x, gy = y0 + y, gz = z0 + z;
                float v = 0.0f;
                if (gx < nx && gy < ny && gz < nz) v = vol[((long)gz * ny + gy) * nx + gx];
                blk[((long)z * by + y) * bx + x] = v;
            }
}
/* Block_Copy.cpp:136-212: clipped scatter */
void cvxo_copy_from_block(const float* blk, int bx, int by, int bz,
                          float* vol, int x0, int y0, int z0, int nx, int ny, int nz)
{
    for (int z = 0; z < bz && z0 + z < nz; ++z)
        for (int y = 0; y < by && y0 + y < ny; ++y)
            for (int x = 0; x < bx && x0 + x < nx; ++x)
                vol[((long)(z0 + z) * ny + (y0 + y)) * nx + (x0 + x)] = blk[((long)z * by + y) * bx + x];
}

/* CvxCompress.cpp:73-117 with one thread: 4 double lanes over the first nn/4*4 cells,
 * hadd + cross-half add (:101-107), scalar tail (:109-113), sqrt of the mean, cast to float. */
float cvxo_global_rms(const float* vol, int nx, int ny, int nz)
{
    long nn = (long)nx * ny * nz, n4 = nn >> 2;
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    for (long i = 0; i < n4; ++i) {
        double v0 = vol[4 * i], v1 = vol[4 * i + 1], v2 = vol[4 * i + 2], v3 = vol[4 * i + 3];
        a0 = a0 + v0 * v0; a1 = a1 + v1 * v1; a2 = a2 + v2 * v2; a3 = a3 + v3 * v3;
    }
    double rms = 0.0;
    rms += (a0 + a1) + (a2 + a3);
    for (long i = n4 * 4; i < nn; ++i) { double d = vol[i]; rms += d * d; }
    rms = sqrt(rms / ((double)nx * (double)ny * (double)nz));
    return (float)rms;
}

/* CvxCompress.cpp:119-142: 8 fp32 lanes, acc = acc + v*v (no FMA), two hadds, halves added */
float cvxo_local_rms(const float* blk, int bx, int by, int bz)
{
    int nn = bz * by * (bx >> 3);
    float a[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    for (int i = 0; i < nn; ++i)
        for (int j = 0; j < 8; ++j) { float v = blk[8 * i + j]; a[j] = a[j] + v * v; }
    float lo = (a[0] + a[1]) + (a[2] + a[3]);
    float hi = (a[4] + a[5]) + (a[6] + a[7]);
    float s = lo + hi;
    return sqrtf(s / (float)(bx * by * bz));
}

/* ------------------------------------------------------------------------------------------
 * Quantise + run-length encode.  Run_Length_Encode_Slow.cpp:189-294 (TMJ_AVX_RLE path).
 * Escape codes: Run_Length_Escape_Codes.hxx:8-14.
 * ------------------------------------------------------------------------------------------ */
#define RLESC1 127
#define VLESC3_8x 126
#define RLESC3 125
#define VLESC2 (-125)
#define VLESC2_8x (-126)
#define VLESC3 (-127)
#define VLESC4 (-128)

/* _mm256_cvttps_epi32 (Run_Length_Encode_Slow.cpp:204): truncation; NaN / out of range -> 0x80000000 */
int cvxo_quantise(float f)
{
    if (!(f > -2147483904.0f && f < 2147483648.0f)) return INT_MIN;   /* also catches NaN */
    return (int)f;
}

/* EncodeRLE_Slow, Run_Length_Encode_Slow.cpp:21-71 */
static void flush_run(int* rle, unsigned char* dst, int* pos)
{
    int r = *rle;
    if (r <= 0) return;
    if (r == 1) {
        dst[(*pos)++] = 0;
    } else if (r < 256) {
        dst[(*pos)++] = (unsigned char)RLESC1;
        dst[(*pos)++] = (unsigned char)r;
    } else {
        dst[(*pos)++] = (unsigned char)RLESC3;
        dst[(*pos)++] = (unsigned char)(r & 0xFF);
        dst[(*pos)++] = (unsigned char)((r >> 8) & 0xFF);
        dst[(*pos)++] = (unsigned char)((r >> 16) & 0xFF);
    }
    *rle = 0;
}

int cvxo_rle_encode(float mulfac, const float* vals, int num, unsigned char* dst)
{
    int rle = 0, pos = 0;
    for (int i = 0; i < num; i += 8) {
        float f[8]; int q[8]; float fq[8];
        int Z = 0, nb = 0, ns = 0, n3 = 0;
        for (int j = 0; j < 8; ++j) {
            f[j] = mulfac * vals[i + j];                       /* :203 */
            q[j] = cvxo_quantise(f[j]);                        /* :204 */
            fq[j] = (float)q[j];                               /* :205, classes are tested on the float */
            if (fq[j] == 0.0f) ++Z;                            /* :206 */
            if (fq[j] > (float)VLESC2 && fq[j] < (float)RLESC3) ++nb;        /* :215 (zeros count as bytes) */
            if (fq[j] >= -32768.0f && fq[j] <= 32767.0f) ++ns;               /* :230 */
            if (fq[j] >= -8388608.0f && fq[j] <= 8388607.0f) ++n3;           /* :245 */
        }
        if (Z == 8) { rle += 8; continue; }                    /* :208-212 */
        if (Z == 0 && nb == 8) {                               /* :216-226 */
            flush_run(&rle, dst, &pos);
            for (int j = 0; j < 8; ++j) dst[pos++] = (unsigned char)(q[j] & 0xFF);
            continue;
        }
        if (Z == 0 && ns == 8 && (nb + (8 - nb) * 3) > 17) {   /* :231-241 */
            flush_run(&rle, dst, &pos);
            dst[pos++] = (unsigned char)VLESC2_8x;
            for (int j = 0; j < 8; ++j) { dst[pos++] = (unsigned char)(q[j] & 0xFF); dst[pos++] = (unsigned char)((q[j] >> 8) & 0xFF); }
            continue;
        }
        if (Z == 0 && n3 == 8 && (nb + (ns - nb) * 3 + (8 - ns) * 4) > 25) {   /* :246-256 */
            flush_run(&rle, dst, &pos);
            dst[pos++] = (unsigned char)VLESC3_8x;
            for (int j = 0; j < 8; ++j) {
                dst[pos++] = (unsigned char)(q[j] & 0xFF);
                dst[pos++] = (unsigned char)((q[j] >> 8) & 0xFF);
                dst[pos++] = (unsigned char)((q[j] >> 16) & 0xFF);
            }
            continue;
        }
        for (int j = 0; j < 8; ++j) {                          /* :259-288, Encode_One_Word :161-174 */
            if (fq[j] == 0.0f) { ++rle; continue; }
            flush_run(&rle, dst, &pos);
            if (fq[j] > (float)VLESC2 && fq[j] < (float)RLESC3) {
                dst[pos++] = (unsigned char)(q[j] & 0xFF);
            } else if (fq[j] >= -32768.0f && fq[j] <= 32767.0f) {
                dst[pos++] = (unsigned char)VLESC2;
                dst[pos++] = (unsigned char)(q[j] & 0xFF); dst[pos++] = (unsigned char)((q[j] >> 8) & 0xFF);
            } else if (fq[j] >= -8388608.0f && fq[j] <= 8388607.0f) {
                dst[pos++] = (unsigned char)VLESC3;
                dst[pos++] = (unsigned char)(q[j] & 0xFF); dst[pos++] = (unsigned char)((q[j] >> 8) & 0xFF);
                dst[pos++] = (unsigned char)((q[j] >> 16) & 0xFF);
            } else {
                uint32_t u; memcpy(&u, &f[j], 4);              /* payload is the SCALED float, :270 */
                dst[pos++] = (unsigned char)VLESC4;
                dst[pos++] = (unsigned char)(u & 0xFF); dst[pos++] = (unsigned char)((u >> 8) & 0xFF);
                dst[pos++] = (unsigned char)((u >> 16) & 0xFF); dst[pos++] = (unsigned char)((u >> 24) & 0xFF);
            }
        }
    }
    flush_run(&rle, dst, &pos);                                /* :294 */
    return pos;
}

/* Run_Length_Decode_Slow, Run_Length_Encode_Slow.cpp:388-527 (its 8-literal fast path :402-412 can
 * never fire -- the byte compares are built with set1_epi32 -- so it is not restated). */
int cvxo_rle_decode(float mulfac, float* vals, int num_expected, const unsigned char* src)
{
    const unsigned char* p = src;
    float scalefac = 1.0f / mulfac;                            /* :392 */
    int num = 0;
    for (; num < num_expected; ++p) {
        int c = (signed char)*p;
        if (c > VLESC2 && c < RLESC3) {                        /* :416-423 */
            vals[num++] = (float)c * scalefac;
        } else if (c == RLESC1) {                              /* :424-434 */
            int r = p[1];
            for (int j = 0; j < r && num + j < num_expected; ++j) vals[num + j] = 0.0f;
            num += r; p += 1;
        } else if (c == RLESC3) {                              /* :435-446 */
            int r = p[1] | (p[2] << 8) | (p[3] << 16);
            for (int j = 0; j < r && num + j < num_expected; ++j) vals[num + j] = 0.0f;
            num += r; p += 3;
        } else if (c == VLESC2) {                              /* :447-457 */
            int q = (int16_t)(p[1] | (p[2] << 8));
            vals[num++] = (float)q * scalefac; p += 2;
        } else if (c == VLESC3) {                              /* :458-468 */
            int q = (int32_t)((uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24) >> 8;
            vals[num++] = (float)q * scalefac; p += 3;
        } else if (c == VLESC2_8x) {                           /* :469-485 */
            for (int j = 0; j < 8; ++j) {
                int q = (int16_t)(p[1 + 2 * j] | (p[2 + 2 * j] << 8));
                if (num + j < num_expected) vals[num + j] = scalefac * (float)q;
            }
            num += 8; p += 16;
        } else if (c == VLESC3_8x) {                           /* :486-504 */
            for (int j = 0; j < 8; ++j) {
                const unsigned char* b = p + 1 + 3 * j;
                int q = (int32_t)((uint32_t)b[0] << 8 | (uint32_t)b[1] << 16 | (uint32_t)b[2] << 24) >> 8;
                if (num + j < num_expected) vals[num + j] = scalefac * (float)q;
            }
            num += 8; p += 24;
        } else {                                               /* VLESC4 :505-513 */
            uint32_t u = (uint32_t)p[1] | (uint32_t)p[2] << 8 | (uint32_t)p[3] << 16 | (uint32_t)p[4] << 24;
            float f; memcpy(&f, &u, 4);
            vals[num++] = f * scalefac; p += 4;
        }
    }
    return (int)(p - src);
}

/* CvxCompress.cpp:54-71 */
int cvxo_is_valid_block_size(int bx, int by, int bz)
{
    int ok = 1;
    int d[3] = { bx, by, bz };
    for (int i = 0; i < 3; ++i) {
        int v = d[i];
        if (v <= 0 || (v & (v - 1)) != 0) ok = 0;
        if (i < 2 ? (v < 8 || v > 256) : !(v == 1 || (v >= 8 && v <= 256))) ok = 0;
    }
    return ok;
}

long cvxo_max_compressed_bytes(int nx, int ny, int nz, int bx, int by, int bz, int local)
{
    long nnn = (long)((nx + bx - 1) / bx) * ((ny + by - 1) / by) * ((nz + bz - 1) / bz);
    return 32 + 8 * nnn + (local ? 4 * nnn : 0) + 4L * bx * by * bz * nnn + 7;
}

/* CvxCompress::Compress, CvxCompress.cpp:231-427, block payloads in natural iBlk order
 * (what the reference produces with num_threads=1; with more threads only the ORDER differs). */
float cvxo_compress(float scale, const float* vol, int nx, int ny, int nz,
                    int bx, int by, int bz, int use_local_rms,
                    unsigned char* compressed, long* compressed_length)
{
    if (!cvxo_is_valid_block_size(bx, by, bz)) { *compressed_length = -1; return 0.0f; }   /* asserts :246-248 */
    float global_rms = use_local_rms ? 1.0f : cvxo_global_rms(vol, nx, ny, nz);           /* :249 */
    int nbx = (nx + bx - 1) / bx, nby = (ny + by - 1) / by, nbz = (nz + bz - 1) / bz;      /* :279-281 */
    long nnn = (long)nbx * nby * nbz;
    int B = bx * by * bz;
    float glob_mulfac = global_rms != 0.0f ? 1.0f / (global_rms * scale) : 1.0f;           /* :291 */
    if (!isfinite(glob_mulfac)) glob_mulfac = 1.0f;                                        /* :294 */
    int32_t hdr[8] = { nx, ny, nz, bx, by, bz, 0, use_local_rms ? 1 : 0 };                 /* :284-300 */
    memcpy(&hdr[6], &glob_mulfac, 4);
    memcpy(compressed, hdr, 32);
    unsigned char* offs = compressed + 32;                                                 /* :302 */
    unsigned char* mulf = offs + 8 * nnn;                                                  /* :308 */
    unsigned char* bytes = use_local_rms ? mulf + 4 * nnn : mulf;                          /* :309,:314 */

    /* encode every block into a private buffer (parallel), then lay them out in order (serial) */
    unsigned char** bufs = (unsigned char**)calloc((size_t)nnn, sizeof(*bufs));
    int* lens = (int*)calloc((size_t)nnn, sizeof(int));
    char* raw = (char*)calloc((size_t)nnn, 1);
    float* bmf = (float*)calloc((size_t)nnn, sizeof(float));
#pragma omp parallel
    {
        float* work = (float*)malloc(sizeof(float) * (size_t)B);
        unsigned char* enc = (unsigned char*)malloc(5 * (size_t)B + 16);
#pragma omp for schedule(dynamic)
        for (long iBlk = 0; iBlk < nnn; ++iBlk) {
            long iiz = iBlk / ((long)nbx * nby), rem = iBlk - iiz * nbx * nby;             /* :321-324 */
            long iiy = rem / nbx, iix = rem - iiy * nbx;
            cvxo_copy_to_block(vol, (int)iix * bx, (int)iiy * by, (int)iiz * bz, nx, ny, nz, work, bx, by, bz);  /* :339 */
            cvxo_wavelet_forward(work, bx, by, bz, CVXO_ASSOC_FAST);                       /* :340 */
            float mulfac = glob_mulfac;
            if (use_local_rms) {                                                           /* :343-348 */
                float lr = cvxo_local_rms(work, bx, by, bz);
                mulfac = lr != 0.0f ? 1.0f / (lr * scale) : 1.0f;
                bmf[iBlk] = mulfac;
            }
            int n = cvxo_rle_encode(mulfac, work, B, enc);                                 /* :349 */
            if (n > 4 * B) {                                                               /* :350,:355-360 raw fallback */
                raw[iBlk] = 1; n = 4 * B;
                bufs[iBlk] = (unsigned char*)malloc((size_t)n); memcpy(bufs[iBlk], work, (size_t)n);
            } else {
                bufs[iBlk] = (unsigned char*)malloc((size_t)n + 1); memcpy(bufs[iBlk], enc, (size_t)n);
            }
            lens[iBlk] = n;
        }
        free(work); free(enc);
    }
    long byte_offset = 0;
    for (long i = 0; i < nnn; ++i) {
        uint64_t o = (uint64_t)byte_offset | (raw[i] ? 0x8000000000000000ull : 0ull);      /* :382-384 */
        memcpy(offs + 8 * i, &o, 8);
        if (use_local_rms) memcpy(mulf + 4 * i, &bmf[i], 4);
        memcpy(bytes + byte_offset, bufs[i], (size_t)lens[i]);
        byte_offset += lens[i];
        free(bufs[i]);
    }
    free(bufs); free(lens); free(raw); free(bmf);
    long len = 32 + 8 * nnn + byte_offset + 7;                                             /* :421 */
    if (use_local_rms) len += 4 * nnn;                                                     /* :422 */
    *compressed_length = len;
    double ratio = ((double)nx * (double)ny * (double)nz * 4.0) / (double)len;             /* :425 */
    return (float)ratio;
}

/* CvxCompress::Decompress, CvxCompress.cpp:463-571 */
int cvxo_decompress(float* vol, int nx, int ny, int nz, const unsigned char* compressed)
{
    int32_t hdr[8]; memcpy(hdr, compressed, 32);
    if (hdr[0] != nx || hdr[1] != ny || hdr[2] != nz) return -1;                           /* :473-488 */
    int bx = hdr[3], by = hdr[4], bz = hdr[5];
    float glob_mulfac; memcpy(&glob_mulfac, &hdr[6], 4);
    int local = hdr[7] & 1;
    int nbx = (nx + bx - 1) / bx, nby = (ny + by - 1) / by, nbz = (nz + bz - 1) / bz;
    long nnn = (long)nbx * nby * nbz;
    int B = bx * by * bz;
    const unsigned char* offs = compressed + 32;
    const unsigned char* mulf = offs + 8 * nnn;
    const unsigned char* bytes = local ? mulf + 4 * nnn : mulf;
#pragma omp parallel
    {
        float* work = (float*)malloc(sizeof(float) * (size_t)B);
#pragma omp for schedule(dynamic)
        for (long iBlk = 0; iBlk < nnn; ++iBlk) {
            long iiz = iBlk / ((long)nbx * nby), rem = iBlk - iiz * nbx * nby;
            long iiy = rem / nbx, iix = rem - iiy * nbx;
            uint64_t o; memcpy(&o, offs + 8 * iBlk, 8);
            int is_raw = (o >> 63) & 1; o &= 0x7FFFFFFFFFFFFFFFull;                         /* :545-547 */
            float mulfac = glob_mulfac;
            if (local) memcpy(&mulfac, mulf + 4 * iBlk, 4);
            if (is_raw) memcpy(work, bytes + o, sizeof(float) * (size_t)B);                /* :555 */
            else cvxo_rle_decode(mulfac, work, B, bytes + o);                              /* :561 */
            cvxo_wavelet_inverse(work, bx, by, bz, CVXO_ASSOC_FAST);                       /* :556/:563 */
            cvxo_copy_from_block(work, bx, by, bz, vol, (int)iix * bx, (int)iiy * by, (int)iiz * bz, nx, ny, nz);
        }
        free(work);
    }
    return 0;
}
