// cvx_bigblock.inl -- line passes for block shapes that do not fit shared memory (any dimension > 64, or more than
// 32768 cells: 64^3 = 1 MiB ... 256^3 = 64 MiB per block), included by cvx_kernels.cu inside namespace cvx.
//
// The reference transforms such blocks with a dedicated strided z pass and a 4 KiB-stride workaround
// (Wavelet_Transform_Fast.cpp:100-145).  Here a batch of blocks lives densely (x fastest) in global scratch and each
// axis is one kernel over all lines of the batch:
//   x lines (contiguous):   a CTA stages a tile of rows in shared memory with coalesced 128-byte accesses (pitch N + 1:
//                           a thread walking its own row hits every bank once per 32 rows), one thread owns one row
//   y / z lines (strided):  consecutive threads own consecutive x, so every element index j is one coalesced access
//                           straight between global memory and the thread's registers (N <= 64), or a shared-memory
//                           column tile (pitch = tile width) for the 128- and 256-long lines
// Lines of <= 64 samples run all levels in registers (the same fwd_line / inv_line the shared-memory kernels use, EXACT
// or LIFTING); longer lines run in place in shared memory (generic-length lifting, or the EXACT filter bank with a level
// copy).  Copy_To_Block is fused into the forward x pass (rows come from the volume, zero padded: Block_Copy.cpp:92-115)
// and Copy_From_Block into the inverse z pass (clipped row stores, :136-212), so a batch makes three scratch round trips
// per direction instead of five.
// Order of the passes: x, y, z in both directions, as the reference (Wavelet_Transform_Fast.cpp:24-145, :148-279).

// ---- generic-length lifting in place on a strided line (shared memory) ------------------------------------------
__device__ __forceinline__ float lift_scale_rt(int p, int logn, bool fwd)
{
    // forward: position 0 carries ZL^levels, position p = 2^k * odd carries ZL^k * ZH (inverse: the reciprocals)
    const float zl = fwd ? CVX_ZL : CVX_IZL, zh = fwd ? CVX_ZH : CVX_IZH;
    const int k = p ? __ffs(p) - 1 : logn;
    float s = p ? zh : 1.0f;
    for (int i = 0; i < k; ++i) s *= zl;
    return s;
}
__device__ __forceinline__ void fwd_lift_mem(float* v, int st, int n, int logn)
{
    for (int S = 1; S < n; S <<= 1) {
        const int m = n / S / 2;
#define CVX_MS(j) v[(S * 2 * (j)) * st]
#define CVX_MD(j) v[(S * (2 * (j) + 1)) * st]
        for (int j = 0; j < m; ++j) CVX_MD(j) = CVX_FMA(CVX_LA, CVX_ADD(CVX_MS(j), CVX_MS(j + 1 < m ? j + 1 : m - 1)), CVX_MD(j));
        for (int j = 0; j < m; ++j) CVX_MS(j) = CVX_FMA(CVX_LB, CVX_ADD(CVX_MD(j > 0 ? j - 1 : 0), CVX_MD(j)), CVX_MS(j));
        for (int j = 0; j < m; ++j) CVX_MD(j) = CVX_FMA(CVX_LG, CVX_ADD(CVX_MS(j), CVX_MS(j + 1 < m ? j + 1 : m - 1)), CVX_MD(j));
        for (int j = 0; j < m; ++j) CVX_MS(j) = CVX_FMA(CVX_LD, CVX_ADD(CVX_MD(j > 0 ? j - 1 : 0), CVX_MD(j)), CVX_MS(j));
    }
    for (int p = 0; p < n; ++p) v[p * st] = CVX_MUL(lift_scale_rt(p, logn, true), v[p * st]);
}
__device__ __forceinline__ void inv_lift_mem(float* v, int st, int n, int logn)
{
    for (int p = 0; p < n; ++p) v[p * st] = CVX_MUL(lift_scale_rt(p, logn, false), v[p * st]);
    for (int S = n / 2; S >= 1; S >>= 1) {
        const int m = n / S / 2;
        for (int j = 0; j < m; ++j) CVX_MS(j) = CVX_FMA(-CVX_LD, CVX_ADD(CVX_MD(j > 0 ? j - 1 : 0), CVX_MD(j)), CVX_MS(j));
        for (int j = 0; j < m; ++j) CVX_MD(j) = CVX_FMA(-CVX_LG, CVX_ADD(CVX_MS(j), CVX_MS(j + 1 < m ? j + 1 : m - 1)), CVX_MD(j));
        for (int j = 0; j < m; ++j) CVX_MS(j) = CVX_FMA(-CVX_LB, CVX_ADD(CVX_MD(j > 0 ? j - 1 : 0), CVX_MD(j)), CVX_MS(j));
        for (int j = 0; j < m; ++j) CVX_MD(j) = CVX_FMA(-CVX_LA, CVX_ADD(CVX_MS(j), CVX_MS(j + 1 < m ? j + 1 : m - 1)), CVX_MD(j));
#undef CVX_MS
#undef CVX_MD
    }
}
// sub-band index i -> interleaved position of an n-long lifted line (runtime n; lift_idx_to_pos is its constexpr twin)
__device__ __forceinline__ int lift_pos_rt(int i, int n)
{
    if (i == 0) return 0;
    const int h = 1 << (31 - __clz(i));
    return (n / (2 * h)) * (2 * (i - h) + 1);
}

struct BigParams {
    float* scratch;            // nb dense blocks (x fastest)
    long blk_first;            // first block (iBlk) of the batch
    int nb;
    BlkShape s;
    int nbx, nby;
    VolView g;
    const float* vol_in;       // forward x pass: rows are gathered from the volume (null: in place on scratch)
    float* vol_out;            // inverse z pass: results are scattered to the volume (null: in place on scratch)
};

__device__ __forceinline__ void big_block_origin(const BigParams& p, long k, int& x0, int& y0, int& z0)
{
    const long t = p.blk_first + k;
    const int iix = (int)(t % p.nbx);
    const long q = t / p.nbx;
    x0 = iix * p.s.bx; y0 = (int)(q % p.nby) * p.s.by; z0 = (int)(q / p.nby) * p.s.bz;
}

// ---- x pass: RT rows of N floats per CTA iteration, one thread per row -----------------------------------------------
template <int N, bool EXACT, bool FWD>
__global__ void __launch_bounds__(128) k_big_x(const BigParams p)
{
    constexpr int RT = N >= 256 ? 64 : 128;            // rows per tile
    constexpr int PITCH = N + 1;
    constexpr bool INREG = N <= 64;
    extern __shared__ __align__(16) float s_tile[];   // [RT][PITCH] (+ [RT][PITCH] level copies for EXACT long lines)
    const long rows_per_blk = p.s.B / N;
    const long ntiles = ((long)p.nb * rows_per_blk + RT - 1) / RT;
    const int tid = threadIdx.x;
    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long r0 = tile * RT;
        const int nrow = (int)min((long)RT, (long)p.nb * rows_per_blk - r0);
        // load (coalesced along x); gather variant reads the volume with zero padding
        for (int i = tid; i < nrow * N; i += 128) {
            const int r = i / N, c = i - r * N;
            const long row = r0 + r;
            float v;
            if (p.vol_in) {
                const long k = row / rows_per_blk, rr = row - k * rows_per_blk;
                int x0, y0, z0;
                big_block_origin(p, k, x0, y0, z0);
                const int gy = y0 + (int)(rr & (p.s.by - 1)), gz = z0 + (int)(rr >> p.s.lby), gx = x0 + c;
                v = (gx < p.g.nx && gy < p.g.ny && gz < p.g.nz) ? __ldg(p.vol_in + ((long)gz * p.g.ny + gy) * p.g.nx + gx) : 0.f;
            } else {
                v = p.scratch[row * N + c];
            }
            // LIFTING inverse wants the line at its interleaved positions
            const int at = (!EXACT && !FWD) ? lift_pos_rt(c, N) : c;
            s_tile[r * PITCH + at] = v;
        }
        __syncthreads();
        if (tid < nrow) {
            float* line = s_tile + tid * PITCH;
            if constexpr (INREG) {
                float v[N];
                if (EXACT && !FWD) {
#pragma unroll
                    for (int i = 0; i < N; ++i) v[i] = line[i];
                    inv_line<N, true>(v);
                } else {
#pragma unroll
                    for (int i = 0; i < N; ++i) v[i] = line[i];
                    if (FWD) fwd_line<N, EXACT>(v); else inv_line<N, EXACT>(v);
                }
#pragma unroll
                for (int i = 0; i < N; ++i) line[i] = v[i];
            } else {
                constexpr int LOGN = N == 128 ? 7 : 8;
                if (EXACT) {
                    float* tmp = s_tile + RT * PITCH + tid * PITCH;
                    if (FWD) fwd_exact_mem(line, 1, tmp, N); else inv_exact_mem(line, 1, tmp, N);
                } else {
                    if (FWD) fwd_lift_mem(line, 1, N, LOGN); else inv_lift_mem(line, 1, N, LOGN);
                }
            }
        }
        __syncthreads();
        for (int i = tid; i < nrow * N; i += 128) {
            const int r = i / N, c = i - r * N;
            const int at = (!EXACT && FWD) ? lift_pos_rt(c, N) : c;       // LIFTING forward leaves sub-band c at an interleaved position
            p.scratch[(r0 + r) * N + c] = s_tile[r * PITCH + at];
        }
        __syncthreads();
    }
}

// ---- y / z pass: lines of N elements `stride` apart; consecutive line ids are consecutive addresses inside a group --------
// axis 1 (y): line l = (x, z): first element at z * bx*by + x, stride bx.   axis 2 (z): line l = (x, y): first element at l.
template <int N, bool EXACT, bool FWD>
__global__ void __launch_bounds__(N <= 64 ? 128 : 64) k_big_strided(const BigParams p, const int axis)
{
    constexpr bool INREG = N <= 64;
    constexpr int T = INREG ? 128 : 64;
    const long lines_per_blk = p.s.B / N;
    const long total = (long)p.nb * lines_per_blk;
    const long stride = axis == 1 ? p.s.bx : (long)p.s.bx * p.s.by;
    const bool scatter = p.vol_out != nullptr;           // set by the host on the last axis of an inverse transform only
    extern __shared__ __align__(16) float s_col[];       // long lines: [N][T] (+ [N][T] level copy for EXACT)
    for (long l0 = (long)blockIdx.x * T; l0 < total; l0 += (long)gridDim.x * T) {
        const long l = l0 + threadIdx.x;
        const bool act = l < total;
        const long k = act ? l / lines_per_blk : 0, ll = act ? l - k * lines_per_blk : 0;
        float* blk = p.scratch + k * p.s.B;
        const long base = axis == 1 ? (ll >> p.s.lbx) * ((long)p.s.bx * p.s.by) + (ll & (p.s.bx - 1)) : ll;
        // destination of sample j of this line when scattering to the volume (inverse z pass)
        float* vdst = nullptr;
        int nclip = 0;
        long vstride = 0;
        if (scatter && act) {
            int x0, y0, z0;
            big_block_origin(p, k, x0, y0, z0);
            const int gx = x0 + (int)(ll & (p.s.bx - 1)), go = (int)(ll >> p.s.lbx);      // go: y (z lines) or z (y lines, bz == 1 blocks)
            if (axis == 2) {
                if (gx < p.g.nx && y0 + go < p.g.ny) {
                    vdst = p.vol_out + ((long)z0 * p.g.ny + y0 + go) * p.g.nx + gx;
                    nclip = min(N, p.g.nz - z0);
                    vstride = (long)p.g.nx * p.g.ny;
                }
            } else if (gx < p.g.nx && z0 + go < p.g.nz) {
                vdst = p.vol_out + ((long)(z0 + go) * p.g.ny + y0) * p.g.nx + gx;
                nclip = min(N, p.g.ny - y0);
                vstride = p.g.nx;
            }
        }
        if constexpr (INREG) {
            if (act) {
                float v[N];
                if (!EXACT && !FWD) {
#pragma unroll
                    for (int i = 0; i < N; ++i) v[sub_to_reg<N, false>(i)] = blk[base + i * stride];
                } else {
#pragma unroll
                    for (int i = 0; i < N; ++i) v[i] = blk[base + i * stride];
                }
                if (FWD) fwd_line<N, EXACT>(v); else inv_line<N, EXACT>(v);
                if (scatter) {
                    if (vdst) {
#pragma unroll
                        for (int i = 0; i < N; ++i) if (i < nclip) vdst[i * vstride] = v[i];
                    }
                } else if (!EXACT && FWD) {
#pragma unroll
                    for (int i = 0; i < N; ++i) blk[base + i * stride] = v[sub_to_reg<N, false>(i)];
                } else {
#pragma unroll
                    for (int i = 0; i < N; ++i) blk[base + i * stride] = v[i];
                }
            }
        } else {
            constexpr int LOGN = N == 128 ? 7 : 8;
            float* col = s_col + threadIdx.x;
            if (act) {
                for (int i = 0; i < N; ++i) col[((!EXACT && !FWD) ? lift_pos_rt(i, N) : i) * T] = blk[base + i * stride];
                if (EXACT) {
                    float* tmp = s_col + N * T + threadIdx.x * N;          // private, contiguous: only this thread touches it
                    if (FWD) fwd_exact_mem(col, T, tmp, N); else inv_exact_mem(col, T, tmp, N);
                } else {
                    if (FWD) fwd_lift_mem(col, T, N, LOGN); else inv_lift_mem(col, T, N, LOGN);
                }
                if (scatter) {
                    if (vdst) for (int i = 0; i < nclip; ++i) vdst[i * vstride] = col[i * T];
                } else {
                    for (int i = 0; i < N; ++i) blk[base + i * stride] = col[((!EXACT && FWD) ? lift_pos_rt(i, N) : i) * T];
                }
            }
        }
    }
}

// ---- long lines (128, 256 samples), LIFTING: four lanes of a warp share a line ----------------------------------------
// One thread per 128- or 256-long line means a serial walk through shared memory, ~32 shared-memory accesses per sample and
// 64 threads per CTA: 0.05 / 0.035 of the 16 N roofline at 128^3 / 256^3.  Here each of four lanes holds a contiguous
// quarter of the line in registers (C = N / 4 samples at positions g * C + i).  A level with stride S < C works inside
// the quarters; the ONE neighbour value a lifting step needs across a quarter boundary comes by shuffle (the lanes of a
// line are D apart).  The last two levels (one sample per quarter) are gathered by shuffles and computed by every lane.
// Same operations in the same order as fwd_lift_mem / inv_lift_mem: bit-identical results.
#define LCS(j) v[S * 2 * (j)]
#define LCD(j) v[S * (2 * (j) + 1)]
template <int C, int S>
__device__ __forceinline__ void lc_fwd_levels(float* v, const bool first, const bool last, const int D)
{
    constexpr int m = C / S / 2;
    float nb = __shfl_down_sync(0xffffffffu, LCS(0), D);           // s_0 of the next quarter (line end: mirror, s_m -> s_(m-1))
    nb = last ? LCS(m - 1) : nb;
#pragma unroll
    for (int j = 0; j < m; ++j) LCD(j) = CVX_FMA(CVX_LA, CVX_ADD(LCS(j), j + 1 < m ? LCS(j + 1) : nb), LCD(j));
    nb = __shfl_up_sync(0xffffffffu, LCD(m - 1), D);               // last d of the previous quarter (line start: d_(-1) -> d_0)
    nb = first ? LCD(0) : nb;
#pragma unroll
    for (int j = 0; j < m; ++j) LCS(j) = CVX_FMA(CVX_LB, CVX_ADD(j > 0 ? LCD(j - 1) : nb, LCD(j)), LCS(j));
    nb = __shfl_down_sync(0xffffffffu, LCS(0), D);
    nb = last ? LCS(m - 1) : nb;
#pragma unroll
    for (int j = 0; j < m; ++j) LCD(j) = CVX_FMA(CVX_LG, CVX_ADD(LCS(j), j + 1 < m ? LCS(j + 1) : nb), LCD(j));
    nb = __shfl_up_sync(0xffffffffu, LCD(m - 1), D);
    nb = first ? LCD(0) : nb;
#pragma unroll
    for (int j = 0; j < m; ++j) LCS(j) = CVX_FMA(CVX_LD, CVX_ADD(j > 0 ? LCD(j - 1) : nb, LCD(j)), LCS(j));
    if constexpr (2 * S < C) lc_fwd_levels<C, 2 * S>(v, first, last, D);
}
template <int C, int S>
__device__ __forceinline__ void lc_inv_levels(float* v, const bool first, const bool last, const int D)
{
    if constexpr (2 * S < C) lc_inv_levels<C, 2 * S>(v, first, last, D);       // coarse levels first
    constexpr int m = C / S / 2;
    float nb = __shfl_up_sync(0xffffffffu, LCD(m - 1), D);
    nb = first ? LCD(0) : nb;
#pragma unroll
    for (int j = 0; j < m; ++j) LCS(j) = CVX_FMA(-CVX_LD, CVX_ADD(j > 0 ? LCD(j - 1) : nb, LCD(j)), LCS(j));
    nb = __shfl_down_sync(0xffffffffu, LCS(0), D);
    nb = last ? LCS(m - 1) : nb;
#pragma unroll
    for (int j = 0; j < m; ++j) LCD(j) = CVX_FMA(-CVX_LG, CVX_ADD(LCS(j), j + 1 < m ? LCS(j + 1) : nb), LCD(j));
    nb = __shfl_up_sync(0xffffffffu, LCD(m - 1), D);
    nb = first ? LCD(0) : nb;
#pragma unroll
    for (int j = 0; j < m; ++j) LCS(j) = CVX_FMA(-CVX_LB, CVX_ADD(j > 0 ? LCD(j - 1) : nb, LCD(j)), LCS(j));
    nb = __shfl_down_sync(0xffffffffu, LCS(0), D);
    nb = last ? LCS(m - 1) : nb;
#pragma unroll
    for (int j = 0; j < m; ++j) LCD(j) = CVX_FMA(-CVX_LA, CVX_ADD(LCS(j), j + 1 < m ? LCS(j + 1) : nb), LCD(j));
}
#undef LCS
#undef LCD
// the whole line transform on the quarter v[0..C) of lane-quarter g (lane0 = lane of quarter 0 of this line)
template <int N, bool FWD>
__device__ __forceinline__ void lc_line(float* v, const int g, const int lane0, const int D)
{
    constexpr int C = N / 4, LOGN = N == 128 ? 7 : 8;
    const bool first = g == 0, last = g == 3;
    if constexpr (FWD) {
        lc_fwd_levels<C, 1>(v, first, last, D);
        float q[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) q[k] = __shfl_sync(0xffffffffu, v[0], lane0 + k * D);
        fwd_lift_levels<4, 1>(q);
        v[0] = g == 0 ? q[0] : g == 1 ? q[1] : g == 2 ? q[2] : q[3];
        v[0] = CVX_MUL(lift_scale_rt(g * C, LOGN, true), v[0]);
#pragma unroll
        for (int i = 1; i < C; ++i) v[i] = CVX_MUL(lift_scale_rt(i, LOGN, true), v[i]);     // ctz(g * C + i) == ctz(i)
    } else {
        v[0] = CVX_MUL(lift_scale_rt(g * C, LOGN, false), v[0]);
#pragma unroll
        for (int i = 1; i < C; ++i) v[i] = CVX_MUL(lift_scale_rt(i, LOGN, false), v[i]);
        float q[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) q[k] = __shfl_sync(0xffffffffu, v[0], lane0 + k * D);
        inv_lift_levels<4, 1>(q);
        v[0] = g == 0 ? q[0] : g == 1 ? q[1] : g == 2 ? q[2] : q[3];
        lc_inv_levels<C, 1>(v, first, last, D);
    }
}

// x pass, long lines, LIFTING: 32 rows per tile, four adjacent lanes per row.  Shared-memory row = four quarters of C + 1
// floats (pitch 4 C + 4 = 4 mod 32): the 32 lanes of a warp reading element i of their quarters hit 32 different banks.
template <int N, bool FWD>
__global__ void __launch_bounds__(128) k_big_x_coop(const BigParams p)
{
    constexpr int C = N / 4, RT = 32, PITCH = 4 * (C + 1);
    extern __shared__ __align__(16) float s_tile[];   // [RT][PITCH]
    const long rows_per_blk = p.s.B / N;
    const long nrows = (long)p.nb * rows_per_blk;
    const long ntiles = (nrows + RT - 1) / RT;
    const int tid = threadIdx.x;
    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long r0 = tile * RT;
        const int nrow = (int)min((long)RT, nrows - r0);
        for (int i = tid; i < nrow * N; i += 128) {
            const int r = i / N, c = i - r * N;
            const long row = r0 + r;
            float x;
            if (p.vol_in) {
                const long k = row / rows_per_blk, rr = row - k * rows_per_blk;
                int x0, y0, z0;
                big_block_origin(p, k, x0, y0, z0);
                const int gy = y0 + (int)(rr & (p.s.by - 1)), gz = z0 + (int)(rr >> p.s.lby), gx = x0 + c;
                x = (gx < p.g.nx && gy < p.g.ny && gz < p.g.nz) ? __ldg(p.vol_in + ((long)gz * p.g.ny + gy) * p.g.nx + gx) : 0.f;
            } else {
                x = p.scratch[row * N + c];
            }
            const int at = FWD ? c : lift_pos_rt(c, N);           // the inverse wants sub-band c at its interleaved position
            s_tile[r * PITCH + (at / C) * (C + 1) + (at % C)] = x;
        }
        __syncthreads();
        {
            const int g = tid & 3;
            float* q = s_tile + (tid >> 2) * PITCH + g * (C + 1);
            float v[C];
#pragma unroll
            for (int i = 0; i < C; ++i) v[i] = q[i];
            lc_line<N, FWD>(v, g, (tid & 31) & ~3, 1);
#pragma unroll
            for (int i = 0; i < C; ++i) q[i] = v[i];
        }
        __syncthreads();
        for (int i = tid; i < nrow * N; i += 128) {
            const int r = i / N, c = i - r * N;
            const int at = FWD ? lift_pos_rt(c, N) : c;           // the forward leaves sub-band c at an interleaved position
            p.scratch[(r0 + r) * N + c] = s_tile[r * PITCH + (at / C) * (C + 1) + (at % C)];
        }
        __syncthreads();
    }
}

// y / z pass, long lines, LIFTING: a warp owns 8 consecutive lines (consecutive x: 32-byte segments per row), lane = g * 8 + line
template <int N, bool FWD>
__global__ void __launch_bounds__(128) k_big_strided_coop(const BigParams p, const int axis)
{
    constexpr int C = N / 4;
    const long lines_per_blk = p.s.B / N;
    const long total = (long)p.nb * lines_per_blk;
    const long stride = axis == 1 ? p.s.bx : (long)p.s.bx * p.s.by;
    const bool scatter = p.vol_out != nullptr;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 3, l8 = lane & 7;
    for (long l0 = (long)blockIdx.x * 32; l0 < total; l0 += (long)gridDim.x * 32) {
        const long l = l0 + warp * 8 + l8;
        const bool act = l < total;
        const long k = act ? l / lines_per_blk : 0, ll = act ? l - k * lines_per_blk : 0;
        float* blk = p.scratch + k * p.s.B;
        const long base = axis == 1 ? (ll >> p.s.lbx) * ((long)p.s.bx * p.s.by) + (ll & (p.s.bx - 1)) : ll;
        float* vdst = nullptr;
        int nclip = 0;
        long vstride = 0;
        if (scatter && act) {
            int x0, y0, z0;
            big_block_origin(p, k, x0, y0, z0);
            const int gx = x0 + (int)(ll & (p.s.bx - 1)), go = (int)(ll >> p.s.lbx);
            if (axis == 2) {
                if (gx < p.g.nx && y0 + go < p.g.ny) {
                    vdst = p.vol_out + ((long)z0 * p.g.ny + y0 + go) * p.g.nx + gx;
                    nclip = min(N, p.g.nz - z0);
                    vstride = (long)p.g.nx * p.g.ny;
                }
            } else if (gx < p.g.nx && z0 + go < p.g.nz) {
                vdst = p.vol_out + ((long)(z0 + go) * p.g.ny + y0) * p.g.nx + gx;
                nclip = min(N, p.g.ny - y0);
                vstride = p.g.nx;
            }
        }
        float v[C];
#pragma unroll
        for (int i = 0; i < C; ++i) {
            const int pos = g * C + i;
            v[i] = act ? blk[base + (long)(FWD ? pos : lift_pos_to_idx(pos, N)) * stride] : 0.f;
        }
        lc_line<N, FWD>(v, g, l8, 8);
        if (act) {
            if (scatter) {
                if (vdst) {
#pragma unroll
                    for (int i = 0; i < C; ++i) if (g * C + i < nclip) vdst[(long)(g * C + i) * vstride] = v[i];
                }
            } else {
#pragma unroll
                for (int i = 0; i < C; ++i) {
                    const int pos = g * C + i;
                    blk[base + (long)(FWD ? lift_pos_to_idx(pos, N) : pos) * stride] = v[i];
                }
            }
        }
    }
}

template <bool EXACT, bool FWD>
static cudaError_t launch_big_axis(const BigParams& p, int axis, int sm_count, cudaStream_t st)
{
    const int n = axis == 0 ? p.s.bx : axis == 1 ? p.s.by : p.s.bz;
    if (n < 2) return cudaSuccess;
    const long lines = (long)p.nb * (p.s.B / n);
#define CVX_BIG_X(NN)                                                                                              \
    case NN: {                                                                                                     \
        constexpr int RT = NN >= 256 ? 64 : 128;                                                                   \
        const size_t smem = (size_t)RT * (NN + 1) * 4 * ((EXACT && NN > 64) ? 2 : 1);                              \
        cudaError_t e = cudaFuncSetAttribute(k_big_x<NN, EXACT, FWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        if (e != cudaSuccess) return e;                                                                            \
        long grid = (lines + RT - 1) / RT;                                                                         \
        if (grid > (long)sm_count * 16) grid = (long)sm_count * 16;                                                \
        k_big_x<NN, EXACT, FWD><<<(unsigned)grid, 128, smem, st>>>(p);                                             \
        break;                                                                                                     \
    }
#define CVX_BIG_S(NN)                                                                                              \
    case NN: {                                                                                                     \
        constexpr int T = NN <= 64 ? 128 : 64;                                                                     \
        const size_t smem = NN <= 64 ? 0 : (size_t)NN * T * 4 * (EXACT ? 2 : 1);                                   \
        cudaError_t e = cudaFuncSetAttribute(k_big_strided<NN, EXACT, FWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        if (e != cudaSuccess) return e;                                                                            \
        long grid = (lines + T - 1) / T;                                                                           \
        if (grid > (long)sm_count * 32) grid = (long)sm_count * 32;                                                \
        k_big_strided<NN, EXACT, FWD><<<(unsigned)grid, T, smem, st>>>(p, axis);                                   \
        break;                                                                                                     \
    }
    if (!EXACT && n >= 128 && !getenv("CVX_GPU_NO_COOP")) {
        // long lines, LIFTING: four lanes per line, quarters in registers (k_big_x_coop / k_big_strided_coop)
        long grid = (lines + 31) / 32;
        const long cap = (long)sm_count * (axis == 0 ? 16 : 32);
        if (grid > cap) grid = cap;
        cudaError_t e = cudaSuccess;
        if (axis == 0) {
            const size_t smem = (size_t)32 * 4 * (n / 4 + 1) * 4;
            if (n == 128) {
                e = cudaFuncSetAttribute(k_big_x_coop<128, FWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                if (e != cudaSuccess) return e;
                k_big_x_coop<128, FWD><<<(unsigned)grid, 128, smem, st>>>(p);
            } else {
                e = cudaFuncSetAttribute(k_big_x_coop<256, FWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                if (e != cudaSuccess) return e;
                k_big_x_coop<256, FWD><<<(unsigned)grid, 128, smem, st>>>(p);
            }
        } else if (n == 128) {
            k_big_strided_coop<128, FWD><<<(unsigned)grid, 128, 0, st>>>(p, axis);
        } else {
            k_big_strided_coop<256, FWD><<<(unsigned)grid, 128, 0, st>>>(p, axis);
        }
        return cudaGetLastError();
    }
    if (axis == 0) {
        switch (n) { CVX_BIG_X(8) CVX_BIG_X(16) CVX_BIG_X(32) CVX_BIG_X(64) CVX_BIG_X(128) CVX_BIG_X(256) default: return cudaErrorInvalidValue; }
    } else {
        switch (n) { CVX_BIG_S(8) CVX_BIG_S(16) CVX_BIG_S(32) CVX_BIG_S(64) CVX_BIG_S(128) CVX_BIG_S(256) default: return cudaErrorInvalidValue; }
    }
#undef CVX_BIG_X
#undef CVX_BIG_S
    return cudaGetLastError();
}

// all three axes of a batch; forward gathers from p.vol_in (when set), inverse scatters to p.vol_out (when set)
static cudaError_t launch_big_transform(const BigParams& p, bool fwd, bool exact, int sm_count, cudaStream_t st)
{
    cudaError_t e = cudaSuccess;
    for (int axis = 0; axis < 3 && e == cudaSuccess; ++axis) {
        BigParams q = p;
        if (axis != 0) q.vol_in = nullptr;
        if (axis != 2 && !(axis == 1 && p.s.bz < 2)) q.vol_out = nullptr;
        if (fwd) e = exact ? launch_big_axis<true, true>(q, axis, sm_count, st) : launch_big_axis<false, true>(q, axis, sm_count, st);
        else e = exact ? launch_big_axis<true, false>(q, axis, sm_count, st) : launch_big_axis<false, false>(q, axis, sm_count, st);
    }
    return e;
}
