// ulk_front.cu -- tile-resident front end of the segmentation path (sm_100a: TMA + shared memory).
//
// After the threshold pass has produced the foreground bit words and their rank prefix, ONE
// kernel turns the dense frame into the rank-space arrays every later kernel works on AND
// evaluates the steepest incident edge of every foreground voxel (K_MinEdge in component space:
// every 4/6-connected foreground neighbour belongs to the same component, so no labels are
// needed yet).  A CTA owns a [TZ, TY, 128] brick of the frame:
//   * the contour brick plus a one-voxel halo is staged in shared memory by a single 3-D TMA
//     load (cp.async.bulk.tensor.3d, out-of-frame elements zero-filled, completion on an
//     mbarrier) while the CTA's threads stage the foreground words / rank offsets of the brick's
//     rows and of the halo rows;
//   * a warp then takes one 128-voxel row chunk at a time, compacts its foreground voxels
//     (their ranks are consecutive), and lane k handles the k-th of them: six bit tests and six
//     shared-memory loads replace the 14 dependent global loads and ~400 instructions per
//     voxel of the list form;
//   * outputs per foreground voxel, all coalesced in rank order: voxel index, contour value,
//     component-forest seed (x-run head), steepest edge (weight, edge key, other endpoint).
// Edge weights and the (weight, key) order are those of edges_from_mask
// (core/segmentation/vendored/graph.py:9-59); see K_MinEdge in ulk_voxel.h for the functor form
// that the host emulation and the non-TMA fallback run.
#include "ulk_seg.h"
#include "ulk_voxel.h"
#include "ulk_front.h"

#ifndef ULK_EMU
#include <cuda.h>

#define FT_TZ 16
#define FT_TY 4
#define FT_TX 128
// x halo of the staged brick: 16 bytes of elements, so that the box start and the row pitch in
// shared memory stay 16-byte aligned for every dtype (f32: 4, f16: 8, u8: 16 elements)
#define FT_HX(CT) ((int)(16 / sizeof(CT)))
#define FT_TXP(CT) (FT_TX + 2 * FT_HX(CT))
#define FT_ROWS ((FT_TZ + 2) * (FT_TY + 2))
#define FT_CAP 4096                    // brick-local voxel list entries per round
#define FT_WPR 6                       // staged words per row: one halo word each side of the brick's four

template <class CT> struct FtConv;
template <> struct FtConv<float> { static __device__ __forceinline__ float f(float v) { return v; } };
template <> struct FtConv<uint16_t> { static __device__ __forceinline__ float f(uint16_t v) { return ulk_half_to_float(v); } };
template <> struct FtConv<uint8_t> { static __device__ __forceinline__ float f(uint8_t v) { return (float)v; } };

static __device__ __forceinline__ u32 ft_smem(const void *p) { return (u32)__cvta_generic_to_shared(p); }

struct FrontArgs {
    const u32 *words; const i32 *off; UlkDims dm; i32 S;
    UlkOW *ow; i32 *fgl; float *ctf; i32 *P; u32 *mw, *pk; i32 *po; u32 *cmax;
};

template <class CT> __global__ void __launch_bounds__(256)
ulk_front_tile_kernel(const __grid_constant__ CUtensorMap tmap, const FrontArgs a)
{
    extern __shared__ __align__(128) unsigned char ft_raw[];
    constexpr int HX = FT_HX(CT), TXP = FT_TXP(CT);
    CT *sct = (CT *)ft_raw;                                                        // [FT_ROWS][TXP]
    uint2 *sow = (uint2 *)(ft_raw + ((sizeof(CT) * FT_ROWS * TXP + 127) / 128) * 128);      // [FT_ROWS][FT_WPR]
    unsigned char *xs = (unsigned char *)(sow + FT_ROWS * FT_WPR);                // brick-local voxel list, u16 per voxel
    __shared__ __align__(8) unsigned long long bar;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const i32 x0 = blockIdx.x * FT_TX, y0 = blockIdx.y * FT_TY, z0 = blockIdx.z * FT_TZ;
    const i32 D = a.dm.D, H = a.dm.H, W = a.dm.W, S = a.S;
    const u32 barp = ft_smem(&bar);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(barp), "r"(1));
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        const u32 bytes = (u32)(sizeof(CT) * FT_ROWS * TXP);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(barp), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(ft_smem(sct)), "l"(&tmap), "r"(x0 - HX), "r"(y0 - 1), "r"(z0 - 1), "r"(barp)
                     : "memory");
    }
    // foreground words + rank offsets of the brick's rows and the halo rows; the brick's own
    // records also go out as the interleaved foreground map
    const i32 wx0 = x0 >> 5;
    {
        constexpr int NIT = (FT_ROWS * FT_WPR + 255) / 256;
        i64 tt[NIT]; u32 ww[NIT]; i32 oo[NIT]; bool own[NIT];
        // all loads of a thread are issued before the first is used
#pragma unroll
        for (int u = 0; u < NIT; ++u) {
            const int i = tid + u * 256;
            const int row = i / FT_WPR, k = i - row * FT_WPR;
            const int lz = row / (FT_TY + 2), ly = row - lz * (FT_TY + 2);
            const i32 z = z0 - 1 + lz, y = y0 - 1 + ly, wx = wx0 - 1 + k;
            const bool ok = i < FT_ROWS * FT_WPR && z >= 0 && z < D && y >= 0 && y < H && wx >= 0 && wx < S;
            tt[u] = ok ? ((i64)z * H + y) * S + wx : -1;
            own[u] = ok && lz >= 1 && lz <= FT_TZ && ly >= 1 && ly <= FT_TY && k >= 1 && k <= 4;
            ww[u] = 0; oo[u] = 0;
            if (ok) { ww[u] = a.words[tt[u]]; oo[u] = a.off[tt[u]]; }
        }
#pragma unroll
        for (int u = 0; u < NIT; ++u) {
            const int i = tid + u * 256;
            if (i < FT_ROWS * FT_WPR) sow[i] = make_uint2(ww[u], (u32)oo[u]);
            if (own[u]) {
                UlkOW r;
                r.w = ww[u]; r.o = oo[u];
                a.ow[tt[u]] = r;
            }
        }
    }
    __syncthreads();
    {   // every thread waits for the TMA bytes (phase 0)
        u32 done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(barp), "r"(0) : "memory");
    }
    // Edge weights: w = float32(double(a + b) / 2 + 1e-8).  For a + b == 0 or 2^-24 <= |a + b| < 2^40
    // the float expression (a + b) * 0.5f + 1e-8f is bit-identical (ulk_wsum_fast), and when every
    // FOREGROUND contour value is 0 or in [2^-23, 2^39) -- true for any sane contour map -- every
    // sum this kernel uses is in that range (only edges between foreground voxels count).  The
    // kernel always takes the float expression; every voxel tests its own value and a single odd
    // one raises a.cmax[1], upon which the caller recomputes the edges with K_MinEdge.
    __shared__ int s_wsum[8];
    // ---- CTA-wide compaction of the brick's foreground voxels: thread t owns word t & 3 of row
    //      chunk t >> 2 (FT_TZ * FT_TY * 4 == 256); a block scan of the population counts gives the
    //      slot of its first voxel in the brick-local list
    uint16_t *list = (uint16_t *)xs;                          // [FT_CAP] entries per round
    const int myrc = tid >> 2, myj = tid & 3;
    const int mlz = myrc / FT_TY, mly = myrc - mlz * FT_TY;
    const int msrow = (mlz + 1) * (FT_TY + 2) + (mly + 1);
    u32 myw = 0;
    if (z0 + mlz < D && y0 + mly < H) myw = sow[msrow * FT_WPR + 1 + myj].x;
    const int mycnt = __popc(myw);
    int incl = mycnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    int base = incl - mycnt, total = 0;
#pragma unroll
    for (int wq = 0; wq < 8; ++wq) { const int t = s_wsum[wq]; if (wq < warp) base += t; total += t; }
    bool odd = false;
    const i32 sy = W, sz = W * H;
    u32 mx = 0;
    // bricks denser than FT_CAP voxels (> 50 % foreground) take more than one round
    for (int round0 = 0; round0 < total; round0 += FT_CAP) {
    if (round0) __syncthreads();
    {
        int slot = base - round0;
        for (u32 m = myw; m; m &= m - 1, ++slot)
            if (slot >= 0 && slot < FT_CAP) list[slot] = (uint16_t)((myrc << 7) | (myj << 5) | (__ffs(m) - 1));
    }
    __syncthreads();
    const int nround = min(total - round0, FT_CAP);
    for (int e = tid; e < nround; e += 256) {
        const int ent = list[e];
        const int rc = ent >> 7, lx = ent & 127;
        const int lz = rc / FT_TY, ly = rc - lz * FT_TY;
        const int srow = (lz + 1) * (FT_TY + 2) + (ly + 1);
        const uint2 *rw = sow + srow * FT_WPR;               // rw[1..4]: the chunk's words
        const int j = lx >> 5, b = lx & 31;
        const u32 below = (1u << b) - 1u;
        const uint2 me = rw[1 + j];
        const u32 wj = me.x;
        const i32 r = (i32)me.y + __popc(wj & below);
        const i32 v = ((z0 + lz) * H + (y0 + ly)) * W + x0 + lx;
        const CT *pc = sct + srow * TXP + HX + lx;
        const float c = FtConv<CT>::f(pc[0]);
        // neighbours in increasing edge-key order: -z, -y, -x, +x, +y, +z
        const uint2 ez0 = rw[1 + j - (FT_TY + 2) * FT_WPR], ey0 = rw[1 + j - FT_WPR];
        const uint2 ey1 = rw[1 + j + FT_WPR], ez1 = rw[1 + j + (FT_TY + 2) * FT_WPR];
        const bool h0 = (ez0.x >> b) & 1u, h1 = (ey0.x >> b) & 1u, h4 = (ey1.x >> b) & 1u, h5 = (ez1.x >> b) & 1u;
        const bool h2 = b > 0 ? ((wj >> (b - 1)) & 1u) : (rw[j].x >> 31);
        const bool h3 = b < 31 ? ((wj >> (b + 1)) & 1u) : (rw[2 + j].x & 1u);
        const float s0 = FtConv<CT>::f(pc[-(FT_TY + 2) * TXP]) + c, s1 = FtConv<CT>::f(pc[-TXP]) + c;
        const float s2 = FtConv<CT>::f(pc[-1]) + c, s3 = c + FtConv<CT>::f(pc[1]);
        const float s4 = c + FtConv<CT>::f(pc[TXP]), s5 = c + FtConv<CT>::f(pc[(FT_TY + 2) * TXP]);
        const float q0 = s0 * 0.5f + 1e-8f, q1 = s1 * 0.5f + 1e-8f, q2 = s2 * 0.5f + 1e-8f;
        const float q3 = s3 * 0.5f + 1e-8f, q4 = s4 * 0.5f + 1e-8f, q5 = s5 * 0.5f + 1e-8f;
        {
            u32 cb;
            memcpy(&cb, &c, 4);
            odd |= !(cb == 0u || (cb >= 0x34000000u && cb < 0x53000000u));      // 0 or [2^-23, 2^39), sign clear
        }
        // first strictly smaller weight in key order wins: (weight, key) minimum.  Float order
        // equals the order of the encoded bits K_MinEdge compares (in-range weights are finite)
        int best = -1;
        u32 bw = ULK_NONE32;
        float bf = __int_as_float(0x7f800000);
        if (h0 && q0 < bf) { bf = q0; best = 0; }
        if (h1 && q1 < bf) { bf = q1; best = 1; }
        if (h2 && q2 < bf) { bf = q2; best = 2; }
        if (h3 && q3 < bf) { bf = q3; best = 3; }
        if (h4 && q4 < bf) { bf = q4; best = 4; }
        if (h5 && q5 < bf) { bf = q5; best = 5; }
        if (best >= 0) bw = ulk_f2ord(bf);
        // edge key = 3 * source voxel + direction; other endpoint = rank of the neighbour
        u32 bk = ULK_NONE32;
        i32 bo = ULK_PO_NONE;
        if (best >= 0) {
            const uint2 er = best == 0 ? ez0 : (best == 1 ? ey0 : (best == 4 ? ey1 : ez1));
            const i32 ro = (i32)er.y + __popc(er.x & below);
            bo = best == 2 ? r - 1 : (best == 3 ? r + 1 : ro);
            const i32 src = best == 0 ? v - sz : (best == 1 ? v - sy : (best == 2 ? v - 1 : v));
            const u32 dir = (best == 0 || best == 5) ? 2u : ((best == 1 || best == 4) ? 1u : 0u);
            bk = 3u * (u32)src + dir;
        }
        // x-run head inside the 32-voxel segment: the component forest starts with every run hooked
        const u32 linked = wj & (wj << 1);
        const u32 stop = ~linked & (0xffffffffu >> (31 - b));
        const int st = 31 - __clz(stop);
        __stcs(a.fgl + r, v);
        __stcs(a.ctf + r, c);
        __stcs(a.P + r, (i32)me.y + __popc(wj & ((1u << st) - 1u)));
        __stcs(a.mw + r, bw);
        __stcs(a.pk + r, bk);
        __stcs(a.po + r, bo);
        const float ac = fabsf(c);
        if (ac <= 3.0e38f) { u32 bits; memcpy(&bits, &ac, 4); mx = bits > mx ? bits : mx; }
    }
    }
    mx = __reduce_max_sync(0xffffffffu, mx);
    if (lane == 0 && mx > *a.cmax) atomicMax(a.cmax, mx);
    if (__any_sync(0xffffffffu, odd) && lane == 0) a.cmax[1] = 1u;
}

typedef CUresult (*ulk_encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                         const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                         CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static ulk_encode_tiled_fn ulk_get_encode()
{
    static ulk_encode_tiled_fn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (ulk_encode_tiled_fn)p;
        else
            cudaGetLastError();
    }
    return fn;
}

bool ulk_front_tile_supported(const void *ct, int ct_dtype, const UlkDims &dm)
{
    if (getenv("ULK_NO_TMA")) return false;
    if (ct_dtype != ULK_F32 && ct_dtype != ULK_F16 && ct_dtype != ULK_U8) return false;
    const i64 es = ulk_dtype_size(ct_dtype);
    if (((uintptr_t)ct & 15u) || ((i64)dm.W * es) % 16 != 0) return false;     // TMA: 16-byte base and strides
    return ulk_get_encode() != nullptr;
}

int ulk_front_tile(ulk_stream_t st, const void *ct, int ct_dtype, const UlkDims &dm, i32 S, const u32 *words, const i32 *off,
                   UlkOW *ow, i32 *fgl, float *ctf, i32 *P, u32 *mw, u32 *pk, i32 *po, u32 *cmax)
{
    const i64 es = ulk_dtype_size(ct_dtype);
    CUtensorMap tm;
    const cuuint64_t gdim[3] = {(cuuint64_t)dm.W, (cuuint64_t)dm.H, (cuuint64_t)dm.D};
    const cuuint64_t gstr[2] = {(cuuint64_t)dm.W * (cuuint64_t)es, (cuuint64_t)dm.W * (cuuint64_t)dm.H * (cuuint64_t)es};
    const cuuint32_t txp = (cuuint32_t)(FT_TX + 2 * (16 / es));
    const cuuint32_t box[3] = {txp, FT_TY + 2, FT_TZ + 2};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUtensorMapDataType dt = ct_dtype == ULK_F32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                   : (ct_dtype == ULK_F16 ? CU_TENSOR_MAP_DATA_TYPE_UINT16 : CU_TENSOR_MAP_DATA_TYPE_UINT8);
    CUresult rc = ulk_get_encode()(&tm, dt, 3, (void *)ct, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) ULK_FAIL(-2, "cuTensorMapEncodeTiled failed (%d)", (int)rc);
    FrontArgs a{words, off, dm, S, ow, fgl, ctf, P, mw, pk, po, cmax};
    dim3 grid((unsigned)((dm.W + FT_TX - 1) / FT_TX), (unsigned)((dm.H + FT_TY - 1) / FT_TY), (unsigned)((dm.D + FT_TZ - 1) / FT_TZ));
    const size_t smem = ((size_t)es * FT_ROWS * txp + 127) / 128 * 128 + sizeof(uint2) * FT_ROWS * FT_WPR + 2 * FT_CAP;
    if (g_ulk_prof_on) ulk_prof_begin("K_FrontTile", st);
#define ULK_FT_LAUNCH(CT)                                                                                                 \
    do {                                                                                                                  \
        static bool attr = false;                                                                                         \
        if (!attr) { cudaFuncSetAttribute(ulk_front_tile_kernel<CT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024); attr = true; } \
        ulk_front_tile_kernel<CT><<<grid, 256, smem, st>>>(tm, a);                                                        \
    } while (0)
    if (ct_dtype == ULK_F32) ULK_FT_LAUNCH(float);
    else if (ct_dtype == ULK_F16) ULK_FT_LAUNCH(uint16_t);
    else ULK_FT_LAUNCH(uint8_t);
#undef ULK_FT_LAUNCH
    if (g_ulk_prof_on) ulk_prof_end(st);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch("K_FrontTile", st);
    ULK_CUDA(cudaGetLastError());
    return 0;
}
#else
bool ulk_front_tile_supported(const void *, int, const UlkDims &) { return false; }
int ulk_front_tile(ulk_stream_t, const void *, int, const UlkDims &, i32, const u32 *, const i32 *, UlkOW *, i32 *, float *,
                   i32 *, u32 *, u32 *, i32 *, u32 *) { return -2; }
#endif
