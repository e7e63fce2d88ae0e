// cvx_encode32ws.inl -- k_encode32ws: warp-specialised 32^3 encoder (LIFTING arithmetic), included after
// cvx_encode32.inl.
//
// k_encode32 runs transform (FMA pipe) and coder (integer pipe, half rate on this part) one after the other.  Here a
// CTA has 24 warps: warps 0..15 (the transform group) take block t+1 from the TMA load through the three line passes
// to the compact (raster index, coefficient) list of its non-zeros, while warps 16..23 (the coder group) turn block t's
// list into tokens: group-of-8 marking, token sizes, payload offset by the decoupled look-back, emission into a
// staging buffer and the deferred copy-out -- exactly the k_encode32 routines, on half the threads.
//   hand-over: two list buffers; FULL[b] (transform -> coder: list b and its message are ready) and EMPTY[b]
//   (coder -> transform: list b has been read) are named barriers over all 768 threads, one side arriving, the other
//   waiting.
//   a dense block (list overflow) or a multiplier outside the sparse path's range is coded by ALL 768 threads with the
//   generic CTA-wide coder (encode_tail) on the coefficients in the block buffer: the transform group posts a DENSE
//   message and both groups meet in encode_tail's CTA barriers.
// Registers: launched at 80 per thread, then 96 for the transform warpgroups and 48 for the coder's (setmaxnreg).

constexpr int E32_TT = 512;                          // transform threads (warps 0..15)
constexpr int E32_CT = 256;                          // coder threads (warps 16..23)
constexpr int E32_T = E32_TT + E32_CT;
constexpr int E32_STAGE = 11264;                     // payload staging bytes, two buffers (deferred copy-out)
constexpr int E32_OFF_LIST = F32_OFF_BASE + F32_ROWS * 4;                  // two lists of F32_LCAP entries
constexpr int E32_OFF_STAGE = E32_OFF_LIST + 2 * F32_LCAP * 8;
constexpr int E32_SMEM = E32_OFF_STAGE + 2 * E32_STAGE;                    // 227 328 bytes

typedef NamedBar<1, E32_TT> ETBar;                   // transform group
typedef NamedBar<2, E32_CT> ECBar;                   // coder group
constexpr int E32_BAR_FULL = 3, E32_BAR_EMPTY = 5;   // + list buffer index
__device__ __forceinline__ void e32_arrive(int id) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "n"(E32_T) : "memory"); }
__device__ __forceinline__ void e32_wait(int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(E32_T) : "memory"); }

struct E32Msg {
    long t;          // ticket (block index); -1: no more blocks
    int n;           // list entries; -1: dense block, code it together
    float mulfac;
};

#ifdef CVX_PHASE_TIMING
#define E32_PH_DECL long long cph_last = clock64(); unsigned long long cph_acc[16] = {0}
#define E32_PH(i) do { if (ctid == 0) { long long c_ = clock64(); cph_acc[i] += (unsigned long long)(c_ - cph_last); cph_last = c_; } } while (0)
#define E32_PH_FLUSH do { if (ctid == 0) for (int i_ = 6; i_ < 12; ++i_) atomicAdd(&g_phase[0][i_], cph_acc[i_]); } while (0)
#else
#define E32_PH_DECL
#define E32_PH(i)
#define E32_PH_FLUSH
#endif

__global__ void __launch_bounds__(E32_T, 1) k_encode32ws(const __grid_constant__ CUtensorMap tmap, const EncParams p)
{
    constexpr bool EXACT = false;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    float* sm = reinterpret_cast<float*>(smem_raw);
    uint2* s_lists = reinterpret_cast<uint2*>(smem_raw + E32_OFF_LIST);
    unsigned char* s_stage = smem_raw + E32_OFF_STAGE;
    __shared__ uint64_t s_bar[1];                    // the block's four TMA boxes complete one barrier phase
    __shared__ int scratch[72];
    __shared__ float s_red[32];
    __shared__ unsigned int s_ticket[2];
    __shared__ unsigned long long s_off;
    __shared__ unsigned long long s_lb[E32_CT / 32 + 2];
    __shared__ int s_chunk[F32_LCAP / 32];
    __shared__ E32Msg s_msg[2];

    const int tid = threadIdx.x, lane = tid & 31;
    const float gmul = p.use_local ? 1.0f : *p.glob_mulfac;

    auto issue_load = [&](unsigned int t) {      // one thread: TMA of block t into the block buffer, 4 boxes of 32x32x8
        const int iix = (int)(t % p.nbx);
        const unsigned int q = t / p.nbx;
        const int iiy = (int)(q % p.nby), iiz = (int)(q / p.nby);
        mbar_expect_tx(&s_bar[0], F32_B * 4);
#pragma unroll
        for (int b = 0; b < 4; ++b) tma_load_3d(sm + b * 8192, &tmap, &s_bar[0], iix * 32, iiy * 32, iiz * 32 + b * 8);
    };
    // the generic coder on the block buffer, all 768 threads (dense block / raw fallback, CvxCompress.cpp:350-360)
    auto dense_block = [&](long t, float mulfac, uint2* list) {
        auto get8 = [&](int g, float* c) {
            const int i = g * 8;
            float4 a = *reinterpret_cast<const float4*>(&sm[physr(i, 5)]);
            float4 b = *reinterpret_cast<const float4*>(&sm[physr(i + 4, 5)]);
            c[0] = a.x; c[1] = a.y; c[2] = a.z; c[3] = a.w; c[4] = b.x; c[5] = b.y; c[6] = b.z; c[7] = b.w;
        };
        encode_tail<E32_T>(p, t, sm, mulfac, get8, [](long i) { return (long)physr((int)i, 5); },
                           reinterpret_cast<int4*>(list), scratch, &s_off);
    };

    if (tid == 0) {
        mbar_init(&s_bar[0], 1);
        const unsigned int t0 = atomicAdd(p.ticket, 1u);
        s_ticket[0] = t0;
        if (t0 < p.nblk) { fence_proxy_async(); issue_load(t0); }
    }
    __syncthreads();

    if (tid >= E32_TT) {
#include "cvx_encode32ws_coder.inl"
        return;
    }
#include "cvx_encode32ws_transform.inl"
}
