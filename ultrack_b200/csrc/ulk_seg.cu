// ulk_seg.cu -- per-frame segmentation pipeline (host orchestration + kernel launches).
//
// Replaces, for one timepoint, ultrack's
//   core/segmentation/processing.py:160-340  (_process)
//   core/segmentation/hierarchy.py:33-90     (create_hierarchies)
//   core/segmentation/vendored/hierarchy.py  (Hierarchy, oversegment_components)
//   core/segmentation/vendored/node.py       (bbox / mask)
// with voxel-scale kernels (ulk_voxel.h) that collapse the frame into watershed basins
// and a basin-scale hierarchy builder (ulk_hier.h).
#include "ulk_seg.h"
#include "ulk_hier.h"
#include "ulk_voxel.h"
#include "ulk_front.h"
#include <cmath>

#define CHK(x) do { int rc__ = (x); if (rc__) return rc__; } while (0)
#define CHKA(a) do { if ((a).failed) ULK_FAIL(-3, "device allocation failed"); } while (0)

// ------------------------------------------------------------------ hierarchy kernels
struct K_HierStage1 {
    const i32 *bs, *es; UlkBasins B; UlkEdges E; UlkHierScratch S;
    ULK_D void operator()(i64 h) const { ulk_hier_stage1((i32)h, bs, es, B, E, S); }
};
struct K_HierCut {
    const i32 *bs, *es; UlkBasins B; UlkEdges E; UlkHierScratch S; double max_area;
    i32 *region_of; u64 *rkey_w, *rkey_k;
    ULK_D void operator()(i64 h) const { ulk_hier_cut((i32)h, bs, es, B, E, S, max_area, region_of, rkey_w, rkey_k); }
};
struct K_HierFull {
    const i32 *bs, *es; UlkBasins B; UlkEdges E; UlkHierScratch S; UlkHierParams P; UlkNodeOut O;
    ULK_D void operator()(i64 h) const { ulk_hier_full((i32)h, bs, es, B, E, S, P, O); }
};

// ------------------------------------------------------------------ node finalisation kernels
struct K_NodeGather {
    const i32 *bs; const i32 *noff; UlkNodeOut O;
    i32 *bbox, *centroid, *area, *hier_id, *parent, *hidx; double *frontier, *height; i64 *mwords;
    ULK_D void operator()(i64 hh) const
    {
        i32 h = (i32)hh;
        i32 o = 2 * bs[h];
        i32 cnt = O.count[h];
        i32 base = noff[h];
        for (i32 s = 0; s < cnt; ++s) {
            i32 g = o + s, n = base + s;
            i32 a = O.area[g];
            area[n] = a;
            frontier[n] = O.frontier[g];
            height[n] = O.height[g];
            hier_id[n] = h + 1;
            parent[n] = O.parent[g] >= 0 ? base + O.parent[g] : -1;
            hidx[n] = O.rsize[h] + O.hidx[g];
            i64 ext[3];
            for (int d = 0; d < 3; ++d) {
                i32 lo = O.bbmin[3 * g + d], hi = O.bbmax[3 * g + d] + 1;
                bbox[6 * n + d] = lo;
                bbox[6 * n + 3 + d] = hi;
                ext[d] = hi - lo;
                // core/segmentation/node.py:229-233: mean of the bbox-relative coordinates,
                // plus the bbox origin, rounded half to even
                i64 srel = O.csum[3 * g + d] - (i64)a * lo;
                double m = (double)srel / (double)a;
                centroid[3 * n + d] = (i32)rint(m + (double)lo);
            }
            mwords[n] = ext[0] * ext[1] * ((ext[2] + 31) / 32);
        }
    }
};
// vendored/hierarchy.py:161-162: a region whose bounding box holds fewer than 8 elements
// raises RuntimeError in the reference; report the first one back to the host.
struct K_RegionSizeCheck {
    const i32 *bs; UlkHierScratch S; u64 *err;   // min over failing regions of (region index, size)
    ULK_D void operator()(i64 hh) const
    {
        i32 h = (i32)hh;
        i32 o = 2 * bs[h], b = bs[h + 1] - bs[h];
        i32 root = o + 2 * b - 2;
        i64 vol = 1;
        for (int d = 0; d < 3; ++d) vol *= (i64)(S.nbbmax[3 * root + d] - S.nbbmin[3 * root + d] + 1);
        if (vol < 8) ulk_atomic_min(err, ((u64)(u32)h << 8) | (u64)vol);   // the reference raises at the first one
    }
};
struct K_BasinNodeGlobal {
    const i32 *g_lab; const i32 *noff; const i32 *basin_node; i32 *bnode;
    ULK_D void operator()(i64 b) const
    {
        i32 s = basin_node[b];
        bnode[b] = s < 0 ? -1 : noff[g_lab[b] - 1] + s;
    }
};
// one 32-byte record per node for the painter: bbox origin, row geometry, mask offset, parent
struct UlkPaintInfo { i32 z0, y0, x0, bh; i32 wpr, parent; i64 moff; };
struct K_PaintInfo {
    const i32 *parent; const i32 *bbox; const i64 *moff; UlkPaintInfo *info;
    ULK_D void operator()(i64 n) const
    {
        const i32 *bb = bbox + 6 * n;
        UlkPaintInfo r;
        r.z0 = bb[0]; r.y0 = bb[1]; r.x0 = bb[2]; r.bh = bb[4] - bb[1];
        r.wpr = (bb[5] - bb[2] + 31) / 32; r.parent = parent[n]; r.moff = moff[n];
        info[n] = r;
    }
};
struct K_MaskPaint {
    const i32 *bas; const i32 *bnode; const UlkPaintInfo *info; UlkDims dm;
    u32 *mask;
    ULK_D ULK_INLINE UlkPaintInfo load(i32 n) const
    {
#ifdef ULK_EMU
        return info[n];
#else
        const int4 a = ((const int4 *)info)[2 * n], c = ((const int4 *)info)[2 * n + 1];
        UlkPaintInfo r;
        r.z0 = a.x; r.y0 = a.y; r.x0 = a.z; r.bh = a.w; r.wpr = c.x; r.parent = c.y;
        r.moff = (i64)(((u64)(u32)c.w << 32) | (u32)c.z);
        return r;
#endif
    }
    struct S1 { i32 z, y, x, b; };
    struct S2 { i32 z, y, x, b, n; };
    ULK_D ULK_INLINE S1 load1(i32 r, i32, i32 z, i32 y, i32 x, bool ok) const { S1 s; s.z = z; s.y = y; s.x = x; s.b = ok ? bas[r] : -1; return s; }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const { S2 s; s.z = a.z; s.y = a.y; s.x = a.x; s.b = a.b; s.n = a.b >= 0 ? bnode[a.b] : -1; return s; }
    // lanes of one basin climb the same ancestor chain in lockstep; those that fall into the
    // same mask word (consecutive x of one row) merge their bits and issue one atomic
    ULK_D ULK_INLINE void run(const S2 &s, bool) const
    {
#ifdef ULK_EMU
        for (i32 n = s.n; n >= 0;) {
            const UlkPaintInfo r = load(n);
            i32 rx = s.x - r.x0;
            i64 word = r.moff + ((i64)(s.z - r.z0) * r.bh + (s.y - r.y0)) * r.wpr + (rx >> 5);
            ulk_atomic_or(&mask[word], 1u << (rx & 31));
            n = r.parent;
        }
#else
        // all lanes stay in the loop until the deepest chain of the warp ends (full-mask
        // collectives); adjacent lanes that hit the same mask word merge their bits with a
        // segmented shuffle OR and the run's first lane issues the atomic
        const int lane = threadIdx.x & 31;
        i32 n = s.b >= 0 ? s.n : -1;
        while (__any_sync(0xffffffffu, n >= 0)) {
            const bool act = n >= 0;
            i64 word = -1 - lane;
            u32 bits = 0;
            i32 parent = -1;
            if (act) {
                const UlkPaintInfo r = load(n);
                const i32 rx = s.x - r.x0;
                word = r.moff + ((i64)(s.z - r.z0) * r.bh + (s.y - r.y0)) * r.wpr + (rx >> 5);
                bits = 1u << (rx & 31);
                parent = r.parent;
            }
            const i64 wprev = __shfl_up_sync(0xffffffffu, word, 1);
            const u32 run = ulk_run_mask(__ballot_sync(0xffffffffu, lane == 0 || wprev != word), lane);
            const int first = __ffs(run) - 1, last = 31 - __clz(run);
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const u32 b2 = __shfl_down_sync(0xffffffffu, bits, off);
                if (lane + off <= last) bits |= b2;
            }
            if (act && lane == first) atomicOr(&mask[word], bits);
            n = parent;
        }
#endif
    }
};
struct K_AncestorCount {
    const i32 *parent; i32 *depth;
    ULK_D void operator()(i64 n) const
    {
        i32 d = 0;
        for (i32 a = parent[n]; a >= 0; a = parent[a]) ++d;
        depth[n] = d;
    }
};
struct K_OverlapWrite {
    const i32 *parent; const i32 *ooff; i32 *ov;
    ULK_D void operator()(i64 n) const
    {
        i32 k = ooff[n];
        for (i32 a = parent[n]; a >= 0; a = parent[a]) { ov[2 * k] = (i32)n; ov[2 * k + 1] = a; ++k; }
    }
};
struct K_Iota { u32 *p; ULK_D void operator()(i64 i) const { p[i] = (u32)i; } };
struct K_Valid { const u32 *c; i32 *f; ULK_D void operator()(i64 i) const { f[i] = c[i] != ULK_NONE32; } };

// ------------------------------------------------------------------ threshold + foreground map
// The dense front end streams the foreground volume ONCE and writes one bit per voxel (a 32-bit
// word per 32-voxel row segment); nothing V-sized is written.  A prefix sum of the words'
// population counts gives every segment the rank of its first foreground voxel, and one
// expansion pass builds the rank-space arrays every later kernel works on (K_Expand).
#ifndef ULK_EMU
// generic dtype / ragged width: a warp takes U consecutive row segments, lane l tests voxel l
template <int U> __global__ void __launch_bounds__(256)
ulk_threshold_kernel(K_Threshold k, u32 ntask)
{
    const int lane = threadIdx.x & 31;
    const u32 t0 = (u32)((((i64)blockIdx.x * 256 + threadIdx.x) >> 5) * U);
    if (t0 >= ntask) return;                  // the whole warp leaves together
    const u32 S = (u32)k.S;
    u32 r = t0 / S, sg = t0 - r * S;
    double f[U]; bool valid[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const i32 x = (i32)sg * 32 + lane;
        valid[u] = t0 + u < ntask && x < k.dm.W;
        f[u] = valid[u] ? ulk_load_as_double(k.fg, k.fgt, (i64)r * k.dm.W + x) : 0.0;
        if (++sg == S) { sg = 0; ++r; }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const u32 word = __ballot_sync(0xffffffffu, valid[u] && f[u] > k.thr);
        if (lane == 0 && t0 + u < ntask) k.words[t0 + u] = word;
    }
}
// W % 4 == 0 and aligned volumes: a lane loads four consecutive voxels with one vector load, a warp
// covers four 32-voxel row segments per task; the foreground word of a segment is the OR of its
// eight lanes' nibbles.  All U loads of a thread are issued before the first is used.
// The comparison `double(v) > thr` of the generic kernel is done in float32 against thr rounded
// DOWN to float32 (ulk_thr_floor_f32): every v here is exactly a float, and no float lies strictly
// between that bound and thr, so the two comparisons agree for every v and every thr.
static float ulk_thr_floor_f32(double thr)
{
    if (thr != thr) return (float)thr;                       // NaN: both comparisons are false
    float t = (float)thr;
    if ((double)t > thr) t = nextafterf(t, -INFINITY);
    return t;
}
struct UlkLoadF32 {
    typedef float4 Raw;
    static __device__ __forceinline__ Raw load(const void *p, i64 v) { return __ldcs((const float4 *)((const float *)p + v)); }
    static __device__ __forceinline__ u32 nib(const Raw &f, float thr)
    {
        return (f.x > thr ? 1u : 0u) | (f.y > thr ? 2u : 0u) | (f.z > thr ? 4u : 0u) | (f.w > thr ? 8u : 0u);
    }
};
struct UlkLoadU8 {
    typedef u32 Raw;
    static __device__ __forceinline__ Raw load(const void *p, i64 v) { return __ldcs((const u32 *)((const u8 *)p + v)); }
    static __device__ __forceinline__ u32 nib(const Raw &q, float thr)
    {
        return ((float)(q & 0xffu) > thr ? 1u : 0u) | ((float)((q >> 8) & 0xffu) > thr ? 2u : 0u) |
               ((float)((q >> 16) & 0xffu) > thr ? 4u : 0u) | ((float)(q >> 24) > thr ? 8u : 0u);
    }
};
struct UlkLoadF16 {
    typedef uint2 Raw;
    static __device__ __forceinline__ Raw load(const void *p, i64 v) { return __ldcs((const uint2 *)((const uint16_t *)p + v)); }
    static __device__ __forceinline__ u32 nib(const Raw &q, float thr)
    {
        const float a = ulk_half_to_float((uint16_t)(q.x & 0xffffu)), b = ulk_half_to_float((uint16_t)(q.x >> 16));
        const float c = ulk_half_to_float((uint16_t)(q.y & 0xffffu)), d = ulk_half_to_float((uint16_t)(q.y >> 16));
        return (a > thr ? 1u : 0u) | (b > thr ? 2u : 0u) | (c > thr ? 4u : 0u) | (d > thr ? 8u : 0u);
    }
};
template <int U, class L> __global__ void __launch_bounds__(256)
ulk_threshold_kernel_v4(const void *fg, float thr, i32 W, u32 S, u32 T, u32 ntask4, u32 *words)
{
    const int lane = threadIdx.x & 31;
    const u32 t0 = (u32)((((i64)blockIdx.x * 256 + threadIdx.x) >> 5) * U);
    if (t0 >= ntask4) return;                 // the whole warp leaves together
    u32 r = t0 / T, q = t0 - r * T;           // row, 128-voxel chunk of the row
    typename L::Raw f[U]; bool valid[U]; u32 wbase[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const i32 x = (i32)q * 128 + lane * 4;
        valid[u] = t0 + u < ntask4 && x < W;
        wbase[u] = r * S + q * 4u;
        if (valid[u]) f[u] = L::load(fg, (i64)r * W + x);
        if (++q == T) { q = 0; ++r; }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const u32 nib = valid[u] ? L::nib(f[u], thr) : 0u;
        const int sub = lane >> 3, pos = (lane & 7) * 4;
        const u32 word = __reduce_or_sync(0xffu << (sub * 8), nib << pos);
        if (valid[u] && (lane & 7) == 0) words[wbase[u] + sub] = word;
    }
}
// expansion, warp cooperative: a thread owns one segment record; the warp then walks its
// non-empty segments and lane l emits voxel l of the segment
__global__ void __launch_bounds__(256)
ulk_expand_kernel(K_Expand k, u32 ntask, UlkFastDiv fS)
{
    const int lane = threadIdx.x & 31;
    const u32 t = blockIdx.x * 256u + threadIdx.x;
    u32 w = 0; i32 o = 0, base = 0;
    if (t < ntask) {
        w = k.words[t]; o = k.off[t];
        const u32 r = fS.div(t);
        base = (i32)(r * (u32)k.dm.W + (t - r * (u32)k.S) * 32u);
        UlkOW e;
        e.w = w; e.o = o;
        k.ow[t] = e;
    }
    u32 m = __ballot_sync(0xffffffffu, w != 0);
    u32 mx = 0;
    while (m) {
        const int j = __ffs(m) - 1;
        m &= m - 1;
        const u32 wj = __shfl_sync(0xffffffffu, w, j);
        const i32 oj = __shfl_sync(0xffffffffu, o, j), bj = __shfl_sync(0xffffffffu, base, j);
        if ((wj >> lane) & 1u) {
            const int st = ulk_run_start(wj & (wj << 1), lane);
            const u32 bits = k.emit(oj + __popc(wj & ((1u << lane) - 1u)), bj + lane, oj + __popc(wj & ((1u << st) - 1u)));
            mx = bits > mx ? bits : mx;
        }
    }
    mx = __reduce_max_sync(0xffffffffu, mx);
    if (lane == 0 && mx > *k.cmax) atomicMax(k.cmax, mx);
}
#endif
// foreground words of the frame (one per 32-voxel row segment)
static int threshold_words(ulk_stream_t st, const UlkDims &dm, const void *fg, int fgt, double thr, i32 S, i64 nt, u32 *words)
{
    K_Threshold k{fg, fgt, thr, dm, S, words};
#ifdef ULK_EMU
    ulk_launch(nt, st, k);
#else
    if (g_ulk_prof_on) ulk_prof_begin(typeid(K_Threshold).name(), st);
    const int esz = ulk_dtype_size(fgt);
    const bool vec4 = (fgt == ULK_F32 || fgt == ULK_U8 || fgt == ULK_F16) && dm.W % 4 == 0 && !getenv("ULK_THR_SCALAR") &&
                      ((uintptr_t)fg & (uintptr_t)(4 * esz - 1)) == 0;
    if (vec4) {
        const u32 T = (u32)((dm.W + 127) / 128);
        const i64 nt4 = (i64)dm.D * dm.H * T;
        const float thr_f = ulk_thr_floor_f32(thr);
        static const int U4 = getenv("ULK_THR_U") ? atoi(getenv("ULK_THR_U")) : 2;     // measured: 1.06 ms (U=2), 1.32 (4), 1.91 (8) at C5
        const i64 warps = (nt4 + U4 - 1) / U4;
        const unsigned grid = (unsigned)((warps + 7) / 8);
#define ULK_THR_LAUNCH(UU, LL) ulk_threshold_kernel_v4<UU, LL><<<grid, 256, 0, st>>>(fg, thr_f, dm.W, (u32)S, T, (u32)nt4, words)
        if (fgt == ULK_F32) { if (U4 == 8) ULK_THR_LAUNCH(8, UlkLoadF32); else if (U4 == 2) ULK_THR_LAUNCH(2, UlkLoadF32); else ULK_THR_LAUNCH(4, UlkLoadF32); }
        else if (fgt == ULK_U8) { if (U4 == 8) ULK_THR_LAUNCH(8, UlkLoadU8); else if (U4 == 2) ULK_THR_LAUNCH(2, UlkLoadU8); else ULK_THR_LAUNCH(4, UlkLoadU8); }
        else { if (U4 == 8) ULK_THR_LAUNCH(8, UlkLoadF16); else if (U4 == 2) ULK_THR_LAUNCH(2, UlkLoadF16); else ULK_THR_LAUNCH(4, UlkLoadF16); }
#undef ULK_THR_LAUNCH
    } else {
        const i64 warps = (nt + 3) / 4;
        ulk_threshold_kernel<4><<<(unsigned)((warps + 7) / 8), 256, 0, st>>>(k, (u32)nt);
    }
    if (g_ulk_prof_on) ulk_prof_end(st);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch("K_Threshold", st);
#endif
    return 0;
}
static void expand_words(ulk_stream_t st, const K_Expand &k, i64 nt)
{
#ifdef ULK_EMU
    ulk_launch(nt, st, k);
#else
    UlkFastDiv fS;
    fS.init((u32)k.S);
    if (g_ulk_prof_on) ulk_prof_begin(typeid(K_Expand).name(), st);
    ulk_expand_kernel<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(k, (u32)nt, fS);
    if (g_ulk_prof_on) ulk_prof_end(st);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch("K_Expand", st);
#endif
}

// fixed-point position for the frontier sums: 2^e2 > max |contour| + 1, sums scaled by 2^(62-e2)
static int ulk_fix_shift_for(u32 cmax_bits)
{
    float cmaxf;
    memcpy(&cmaxf, &cmax_bits, 4);
    int e2 = 0;
    while (ldexp(1.0, e2) <= (double)cmaxf + 1.0) ++e2;
    return 62 - e2;
}

// K_Attach on the WHOLE foreground (ranks 0..n-1 are contiguous): a thread takes four consecutive
// ranks with one 128-bit load per array; the climb itself is the functor's (area attribute only --
// the volume attribute needs the warp reduction of the list form)
#ifndef ULK_EMU
__global__ void __launch_bounds__(256) ulk_attach_vec4_kernel(K_Attach k, i64 n)
{
    const i64 g = (i64)blockIdx.x * 256 + threadIdx.x;
    const i64 r0 = g * 4;                      // lanes past the end stay for the warp collectives below
    i32 b[4]; u32 j[4], m[4];
    if (r0 + 4 <= n) {
        const int4 bq = __ldcs((const int4 *)(k.bas + r0));
        const uint4 jq = __ldcs((const uint4 *)(k.jk + r0)), mq = __ldcs((const uint4 *)(k.mw + r0));
        b[0] = bq.x; b[1] = bq.y; b[2] = bq.z; b[3] = bq.w;
        j[0] = jq.x; j[1] = jq.y; j[2] = jq.z; j[3] = jq.w;
        m[0] = mq.x; m[1] = mq.y; m[2] = mq.z; m[3] = mq.w;
    } else {
        for (int e = 0; e < 4; ++e) {
            const bool ok = r0 + e < n;
            b[e] = ok ? k.bas[r0 + e] : -1; j[e] = ok ? k.jk[r0 + e] : ULK_JK_CORE; m[e] = ok ? k.mw[r0 + e] : 0u;
        }
    }
    K_Attach::S2 s[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        K_Attach::S1 a;
        a.b = b[e]; a.j = j[e]; a.m = m[e];
        s[e] = k.load2(a, true);
    }
    // the four voxels usually sit in one basin and attach to the same node: their counter updates
    // are merged before they become atomics (the per-basin counters are heavily contended)
    i32 pf = -1, pt = -1, pc = 0;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        if (!s[e].climb) continue;
        const i32 node = k.climb(s[e]);
        if (node == s[e].node) continue;
        const i32 from = s[e].o + s[e].node, to = s[e].o + node;
        if (from == pf && to == pt) { ++pc; continue; }
        if (pc) { atomicAdd(&k.attach[pf], -pc); atomicAdd(&k.attach[pt], pc); }
        pf = from; pt = to; pc = 1;
    }
    // ... and the last pending update of neighbouring lanes (consecutive ranks) likewise: segmented
    // sum over runs of equal (from, to)
    const int lane = threadIdx.x & 31;
    const i32 fprev = __shfl_up_sync(0xffffffffu, pf, 1), tprev = __shfl_up_sync(0xffffffffu, pt, 1);
    const u32 run = ulk_run_mask(__ballot_sync(0xffffffffu, lane == 0 || fprev != pf || tprev != pt), lane);
    const int first = __ffs(run) - 1, last = 31 - __clz(run);
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const i32 c2 = __shfl_down_sync(0xffffffffu, pc, off);
        if (lane + off <= last) pc += c2;
    }
    if (pc && lane == first) { atomicAdd(&k.attach[pf], -pc); atomicAdd(&k.attach[pt], pc); }
}
#endif

// ------------------------------------------------------------------ voxel stage
struct VoxelStage {
    i32 *bas = nullptr; i32 B = 0; i32 Eg = 0;
    i32 *g_root = nullptr, *g_lab = nullptr; u32 *g_L = nullptr;
    i32 *b_size = nullptr; u64 *b_top = nullptr; u32 *b_k3 = nullptr;
    i32 *bbmin = nullptr, *bbmax = nullptr; i64 *csum = nullptr;
    i32 *bs = nullptr, *es = nullptr;
    i32 *ea = nullptr, *eb = nullptr; u64 *wk = nullptr; u32 *k3 = nullptr, *cnt = nullptr; u64 *shi = nullptr, *slo = nullptr;
    u32 *mw = nullptr, *jk = nullptr;
};

// One watershed-collapse pass over a list of foreground voxels: `rl` holds the ranks of the
// list (nullptr: every foreground voxel), m its length; every array is indexed by rank.
// The steepest edges (mw, pk, po) come in precomputed in COMPONENT space by the front end; `redo`
// lists the ranks whose region is smaller than their component (voxels of oversegmented
// components in pass B): K_MinEdge runs again on those only, with the region labels.
static int voxel_stage(ulk_stream_t st, UlkArena &A, const UlkFg &fg, const i32 *rl, i64 m, const UlkCt &ct, const i32 *lab,
                       i32 R, bool need_join, bool need_k3, bool need_stats, int fix_shift,
                       u32 *mw, u32 *pk, i32 *po, const i32 *redo, i64 nredo,
                       VoxelStage &out, ulk_seg_stats *stats)
{
    const double fixscale = ldexp(1.0, fix_shift);
    const UlkDims &dm = fg.dm;
    const i64 n = fg.n;
    i32 *bp = A.get<i32>(n);
    u8 *state = A.get<u8>(n + 4);
    u32 *Lr = A.get<u32>(n), *best = A.get<u32>(n);
    u8 *fmark = A.get<u8>(n + 8);
    i32 *cand = A.get<i32>(m), *cand2 = A.get<i32>(m);
    i32 *groots = A.get<i32>(m);
    i32 *sroots = A.get<i32>(m);     // every supervertex root after the hook round
    i32 *ctr = A.get<i32>(8);        // [0] candidates [1] growing roots [2] scratch flag [4] basins [5] flow edges [6] supervertex roots
    CHKA(A);
    CHK(ulk_memset(fmark, 0, (size_t)n + 8, st));
    CHK(ulk_memset(ctr, 0, 8 * sizeof(i32), st));
    if (nredo > 0) ulk_launch_rkx<1>("MinEdge", redo, fg.fgl, nredo, dm, st, K_MinEdge{lab, ct, fg, mw, pk, po});
    ulk_launch_rkn<4>(rl, m, st, K_Hook1{mw, po, bp, state, Lr, best, ctr + 4});
    ulk_launch_rk_multi(rl, m, st, K_CompressCand{bp, mw, Lr, state, cand, ctr + 0, groots, ctr + 1, sroots, ctr + 6});
    i32 hc[8];
    CHK(ulk_read(hc, ctr, 8, st));
    i32 ncand = hc[0], ng = hc[1];
    const i32 nsroots = hc[6];
    // every growing root hooks at most once: that bounds the flow-edge list
    const i32 flow_cap = ng;
    i32 *groots2 = A.get<i32>(ng), *tgt = A.get<i32>(ng), *srcr = A.get<i32>(ng);
    i32 *flow_root = A.get<i32>(flow_cap), *flow_src = A.get<i32>(flow_cap);
    u32 *flow_key = A.get<u32>(flow_cap);
    CHKA(A);
    int rounds = 0;
    while (ng > 0) {
        CHK(ulk_memset(ctr, 0, 2 * sizeof(i32), st));
        ulk_launch(ncand, st, K_FindFlow{cand, lab, bp, mw, Lr, state, ct, fg, best, cand2, ctr + 0});
        ulk_launch(ng, st, K_Decide1{groots, bp, best, fg, tgt, srcr});
        ulk_launch(ng, st, K_Decide2{groots, tgt, srcr, bp, state, best, fmark, groots2, ctr + 1,
                                     flow_root, flow_key, flow_src, ctr + 5, ctr + 4});
        CHK(ulk_read(hc, ctr, 8, st));
        ncand = hc[0]; ng = hc[1];
        i32 *t = cand; cand = cand2; cand2 = t;
        t = groots; groots = groots2; groots2 = t;
        if (++rounds > 100000) ULK_FAIL(-4, "flow rounds did not converge");
    }
    if (stats) stats->flow_rounds += rounds;
    const i32 B = hc[4], nflow = hc[5];
    out.B = B;
    A.release(cand); A.release(cand2); A.release(groots); A.release(groots2); A.release(tgt); A.release(srcr);

    // ---- basin ids, grouped by region label: sort of the (label, root rank) keys
    u64 *rootkeys = A.get<u64>(B), *rootsorted = A.get<u64>(B);
    out.g_root = A.get<i32>(B); out.g_lab = A.get<i32>(B); out.g_L = A.get<u32>(B);
    out.bs = A.get<i32>(R + 1);
    i32 *lab_hist = A.get<i32>(R + 1);
    out.bas = A.get<i32>(n);
    CHKA(A);
    CHK(ulk_memset(ctr, 0, sizeof(i32), st));
    CHK(ulk_memset(lab_hist, 0, (size_t)(R + 1) * sizeof(i32), st));
    // the flow rounds re-hooked roots only: flattening the root list is enough
    ulk_launch_rk_multi(sroots, nsroots, st, K_CompressRoots{bp, lab, rootkeys, ctr + 0});
    A.release(sroots);
    CHK(ulk_sort_keys_u64(rootkeys, rootsorted, B, st));
    u32 *gid_at_root = best;         // free after the flow rounds
    ulk_launch(B, st, K_BasinRegroup{rootsorted, Lr, gid_at_root, out.g_root, out.g_lab, out.g_L, lab_hist});
    CHK(ulk_exclusive_scan_i32(lab_hist, out.bs, R + 1, st));
    // every list voxel gets its basin id (-1 outside the regions); a stencil kernel never reads the
    // basin of a foreground voxel outside the list (a foreground neighbour is in the same component)
    ulk_launch_rkn<4>(rl, m, st, K_AssignBasin{bp, gid_at_root, out.bas});
    A.release(rootkeys); A.release(rootsorted);

    // ---- per-basin accumulators, join keys (area attribute), bottleneck lower bound: one pass
    out.b_size = A.get<i32>(B); out.b_top = A.get<u64>(B);
    u64 *ncross = A.get<u64>(1);
    if (need_stats) { out.bbmin = A.get<i32>(3 * (i64)B); out.bbmax = A.get<i32>(3 * (i64)B); out.csum = A.get<i64>(3 * (i64)B); }
    i32 *hp = nullptr;
    out.mw = mw;
    if (need_join) { out.jk = A.get<u32>(n); hp = A.get<i32>(m); }
    i32 *xl = A.get<i32>(m);         // voxels on a boundary between basins of one region
    u32 *Tk = nullptr;
    if (need_k3) { out.b_k3 = A.get<u32>(B); Tk = A.get<u32>(B); }
    CHKA(A);
    CHK(ulk_memset(out.b_size, 0, (size_t)B * sizeof(i32), st));
    CHK(ulk_memset(out.b_top, 0, (size_t)B * sizeof(u64), st));
    CHK(ulk_memset(ncross, 0, sizeof(u64), st));
    CHK(ulk_memset(ctr, 0, 4 * sizeof(i32), st));
    if (need_k3) { CHK(ulk_memset(out.b_k3, 0, (size_t)B * sizeof(u32), st)); CHK(ulk_memset(Tk, 0, (size_t)B * sizeof(u32), st)); }
    if (need_stats) {
        ulk_fill<i32>(out.bbmin, 0x7fffffff, 3 * (i64)B, st);
        ulk_fill<i32>(out.bbmax, -0x7fffffff, 3 * (i64)B, st);
        CHK(ulk_memset(out.csum, 0, (size_t)B * 3 * sizeof(i64), st));
    }
    ulk_launch_rkx<1>("BasinAccum", rl, fg.fgl, m, dm, st, K_BasinAccum{out.bas, mw, pk, po, fg, out.b_size, out.b_top, ncross,
                                        need_stats ? 1 : 0, out.bbmin, out.bbmax, out.csum,
                                        need_join ? 1 : 0, out.g_L, out.jk, hp, ctr + 0,
                                        need_k3 ? 1 : 0, out.b_k3, Tk, xl, ctr + 1, out.g_lab});
    ulk_launch(nflow, st, K_FlowTop{flow_root, flow_key, flow_src, out.bas, ct, fg, out.b_top, Tk});
    A.release(flow_root); A.release(flow_key); A.release(flow_src);
    u64 h_ncross = 0;
    CHK(ulk_read2(&h_ncross, ncross, 1, hc, ctr, 4, st));
    const i32 nxl = hc[1];
    if (need_join) {
        // plateau voxels above their basin level: minimax relaxation along spanning-tree edges
        i32 nhp = hc[0];
        if (stats) stats->plateau_voxels += nhp;
        int it = 0;
        while (nhp > 0) {
            CHK(ulk_memset(ctr + 2, 0, sizeof(i32), st));
            ulk_launch(nhp, st, K_PlateauRelax{hp, lab, mw, pk, fmark, ct, fg, out.jk, ctr + 2});
            CHK(ulk_read(hc, ctr, 4, st));
            if (!hc[2]) break;
            if (++it > 1000000) ULK_FAIL(-4, "plateau relaxation did not converge");
        }
        A.release(hp);
    }

    // ---- region adjacency graph (hash of basin pairs)
    u64 cap = 1024;
    while (cap < 2 * h_ncross + 16) cap <<= 1;
    UlkRagTable T;
    T.key = A.get<u64>((i64)cap); T.wk = A.get<u64>((i64)cap); T.k3 = A.get<u32>((i64)cap);
    T.cnt = A.get<u32>((i64)cap); T.shi = A.get<u64>((i64)cap); T.slo = A.get<u64>((i64)cap);
    T.mask = cap - 1;
    CHKA(A);
    CHK(ulk_memset(T.key, 0xff, cap * 8, st));
    CHK(ulk_memset(T.wk, 0xff, cap * 8, st));
    CHK(ulk_memset(T.k3, 0xff, cap * 4, st));
    CHK(ulk_memset(T.cnt, 0, cap * 4, st));
    CHK(ulk_memset(T.shi, 0, cap * 8, st));
    CHK(ulk_memset(T.slo, 0, cap * 8, st));
    ulk_launch_rkx<1>("RagInsert", xl, fg.fgl, nxl, dm, st, K_RagInsert{out.bas, out.g_lab, ct, fg, T, fixscale});
    A.release(xl);
    i32 *eflag = A.get<i32>((i64)cap + 1), *epre = A.get<i32>((i64)cap + 1);
    CHKA(A);
    ulk_launch((i64)cap, st, K_RagFlag{T.key, eflag});
    CHK(ulk_memset(eflag + cap, 0, sizeof(i32), st));
    CHK(ulk_exclusive_scan_i32(eflag, epre, (i64)cap + 1, st));
    i32 Eg = 0;
    CHK(ulk_read(&Eg, epre + cap, 1, st));
    out.Eg = Eg;
    i32 *ea0 = A.get<i32>(Eg), *eb0 = A.get<i32>(Eg);
    u64 *wk0 = A.get<u64>(Eg), *shi0 = A.get<u64>(Eg), *slo0 = A.get<u64>(Eg);
    u32 *k30 = A.get<u32>(Eg), *cnt0 = A.get<u32>(Eg), *elab0 = A.get<u32>(Eg), *eidx0 = A.get<u32>(Eg);
    CHKA(A);
    ulk_launch((i64)cap, st, K_RagGather{T, epre, out.g_lab, ea0, eb0, wk0, k30, cnt0, shi0, slo0, elab0, eidx0});
    A.release(T.key); A.release(T.wk); A.release(T.k3); A.release(T.cnt); A.release(T.shi); A.release(T.slo);
    A.release(eflag); A.release(epre);
    // sort by (label, (w,key)): LSD, two stable passes
    u64 *wk1 = A.get<u64>(Eg);
    u32 *perm1 = A.get<u32>(Eg), *elab1 = A.get<u32>(Eg), *elab2 = A.get<u32>(Eg), *perm2 = A.get<u32>(Eg);
    CHKA(A);
    CHK(ulk_sort_pairs_u64(wk0, wk1, eidx0, perm1, Eg, st));
    ulk_launch(Eg, st, K_Gather<u32>{elab0, perm1, elab1});
    CHK(ulk_sort_pairs_u32(elab1, elab2, perm1, perm2, Eg, st));
    out.ea = A.get<i32>(Eg); out.eb = A.get<i32>(Eg); out.wk = A.get<u64>(Eg); out.k3 = A.get<u32>(Eg);
    out.cnt = A.get<u32>(Eg); out.shi = A.get<u64>(Eg); out.slo = A.get<u64>(Eg);
    out.es = A.get<i32>(R + 1);
    i32 *ehist = A.get<i32>(R + 1);
    CHKA(A);
    ulk_launch(Eg, st, K_Gather<i32>{ea0, perm2, out.ea});
    ulk_launch(Eg, st, K_Gather<i32>{eb0, perm2, out.eb});
    ulk_launch(Eg, st, K_Gather<u64>{wk0, perm2, out.wk});
    ulk_launch(Eg, st, K_Gather<u32>{k30, perm2, out.k3});
    ulk_launch(Eg, st, K_Gather<u32>{cnt0, perm2, out.cnt});
    ulk_launch(Eg, st, K_Gather<u64>{shi0, perm2, out.shi});
    ulk_launch(Eg, st, K_Gather<u64>{slo0, perm2, out.slo});
    CHK(ulk_memset(ehist, 0, (size_t)(R + 1) * sizeof(i32), st));
    ulk_launch(Eg, st, K_Hist{elab2, ehist});
    CHK(ulk_exclusive_scan_i32(ehist, out.es, R + 1, st));
    A.release(ea0); A.release(eb0); A.release(wk0); A.release(shi0); A.release(slo0);
    A.release(k30); A.release(cnt0); A.release(elab0); A.release(eidx0);
    A.release(wk1); A.release(perm1); A.release(elab1); A.release(elab2); A.release(perm2);

    // ---- raster-key bottleneck per basin
    if (need_k3) {
        i32 *P3 = bp;          // basin forest no longer needed
        i32 *nroot = A.get<i32>(B);
        CHKA(A);
        // only basins with a spanning-tree edge above K0 need the connectivity check; their voxels
        // form the list the check runs on (K_K3Need writes its marks into nroot)
        CHK(ulk_memset(ctr, 0, sizeof(i32), st));
        i32 *needlist = A.get<i32>(B);
        CHKA(A);
        ulk_launch(B, st, K_K3Need{Tk, out.b_k3, nroot, ctr + 0, needlist});
        CHK(ulk_read(hc, ctr, 1, st));
        const i32 *kl = rl;
        i64 nkl = hc[0] > 0 ? m : 0;
        i32 *klist = nullptr, *nkl_dev = nullptr;
        if (hc[0] > 0 && hc[0] < B && !getenv("ULK_K3_ALL")) {
            klist = A.get<i32>(m); nkl_dev = A.get<i32>(1);
            CHKA(A);
#ifdef ULK_EMU
            CHK(ulk_select_list_split_i32(rl, m, out.bas, nroot, klist, nkl_dev, st));
#else
            CHK(ulk_memset(nkl_dev, 0, sizeof(i32), st));
            ulk_launch_warp((i64)hc[0] * 32 * ULK_K3_WARPS, st, K_K3Collect{needlist, hc[0], out.bas, out.bbmin, out.bbmax, fg, klist, nkl_dev});
#endif
            i32 nk = 0;
            CHK(ulk_read(&nk, nkl_dev, 1, st));
            kl = klist; nkl = nk;
        }
        if (stats) stats->k3_voxels += nkl;
        ulk_launch_rk<4>(kl, fg.fgl, nkl, dm, st, K_K3Init{out.bas, out.b_k3, fg.fgl, P3});
        ulk_launch_rkx<4>("K3Union", kl, fg.fgl, nkl, dm, st, K_K3Union{out.bas, fg, out.b_k3, P3});
        // basins that edges <= K0 leave split (rare) go on with Boruvka rounds on THEIR voxels only
        const i32 *cur = kl;
        i64 ncur = nkl;
        i32 *sub = nullptr, *nsub_dev = nullptr;
        for (int it = 0; nkl > 0; ++it) {
            CHK(ulk_memset(nroot, 0, (size_t)B * sizeof(i32), st));
            CHK(ulk_memset(ctr, 0, sizeof(i32), st));
            ulk_launch_rk_multi(cur, ncur, st, K_K3FlattenCount{out.bas, P3, nroot, ctr + 0});
            CHK(ulk_read(hc, ctr, 1, st));
            if (hc[0] == 0) break;
            if (stats) stats->k3_rounds += 1;
            if (it == 0) {
                sub = A.get<i32>(nkl); nsub_dev = A.get<i32>(1);
                CHKA(A);
                CHK(ulk_select_list_split_i32(kl, nkl, out.bas, nroot, sub, nsub_dev, st));
                i32 nsub = 0;
                CHK(ulk_read(&nsub, nsub_dev, 1, st));
                cur = sub; ncur = nsub;
                ulk_launch(ncur, st, K_FillAt<u32>{cur, best, ULK_NONE32});     // K_K3Hook resets what it used
            }
            ulk_launch_rk1(cur, fg.fgl, ncur, dm, st, K_K3FindBest{out.bas, fg, nroot, P3, best});
            ulk_launch_rk1(cur, fg.fgl, ncur, dm, st, K_K3Hook{out.bas, fg, P3, best, out.b_k3});
            if (it > 100000) ULK_FAIL(-4, "bottleneck rounds did not converge");
        }
        if (sub) { A.release(sub); A.release(nsub_dev); }
        if (klist) { A.release(klist); A.release(nkl_dev); }
        A.release(needlist);
        A.release(Tk);
        A.release(nroot);
    }
    A.release(state); A.release(Lr); A.release(best); A.release(fmark);
    A.release(bp);
    if (stats) { stats->n_basins += B; stats->n_rag_edges += Eg; }
    return 0;
}

static int alloc_scratch(UlkArena &A, i32 B, i32 Eg, UlkHierScratch &S)
{
    i64 N2 = 2 * (i64)B + 2;
    i64 M = (Eg > B ? Eg : B) + 2;
    S.Ipar = A.get<i32>(N2); S.Ic0 = A.get<i32>(N2); S.Ic1 = A.get<i32>(N2);
    S.Iwk = A.get<u64>(N2); S.IminL = A.get<u32>(N2); S.attach = A.get<i32>(N2);
    S.awhi = A.get<u64>(N2); S.awlo = A.get<u64>(N2);
    S.acc1 = A.get<i32>(B + 2); S.wprime = A.get<double>(N2);
    S.par = A.get<i32>(N2); S.c0 = A.get<i32>(N2); S.c1 = A.get<i32>(N2);
    S.alt = A.get<double>(N2); S.alt1 = A.get<double>(N2); S.par1 = A.get<i32>(N2);
    S.acc = A.get<i32>(B + 2); S.uf = A.get<i32>(B + 2); S.rootnode = A.get<i32>(B + 2);
    S.ord = A.get<i32>(M); S.sal = A.get<double>(M);
    S.area = A.get<i32>(N2); S.Xhi = A.get<u64>(N2); S.Xlo = A.get<u64>(N2); S.Xc = A.get<i64>(N2);
    S.fp = A.get<i32>(N2); S.surv = A.get<u8>(N2); S.fin = A.get<i32>(N2);
    S.fr = A.get<double>(N2); S.low = A.get<double>(N2); S.emit_idx = A.get<i32>(N2);
    S.nbbmin = A.get<i32>(3 * N2); S.nbbmax = A.get<i32>(3 * N2); S.ncsum = A.get<i64>(3 * N2);
    return A.failed ? -3 : 0;
}

void ulk_frame_nodes_free_impl(ulk_frame_nodes *f)
{
    if (!f) return;
    ulk_dev_free(f->bbox); ulk_dev_free(f->centroid); ulk_dev_free(f->area); ulk_dev_free(f->frontier);
    ulk_dev_free(f->height); ulk_dev_free(f->hier_id); ulk_dev_free(f->parent); ulk_dev_free(f->hidx);
    ulk_dev_free(f->mask_off); ulk_dev_free(f->mask); ulk_dev_free(f->overlaps); ulk_dev_free(f->labels);
    delete f;
}

template <class T> static T *keep_alloc(i64 n)
{
    void *p = nullptr;
    if (ulk_dev_alloc(&p, (size_t)(n > 0 ? n : 1) * sizeof(T))) return nullptr;
    return (T *)p;
}

// the caller asked for the region-label volume: all zero when no region exists
static int empty_labels(ulk_stream_t st, ulk_frame_nodes *F, i64 V, int keep_labels)
{
    if (!keep_labels) return 0;
    F->labels = keep_alloc<i32>(V);
    if (!F->labels) ULK_FAIL(-3, "device allocation failed");
    CHK(ulk_memset(F->labels, 0, (size_t)V * sizeof(i32), st));
    return ulk_sync(st);
}

int ulk_segment_frame_impl(ulk_stream_t st, const void *fgp, int fg_dtype, const void *ctp, int ct_dtype,
                           const int64_t shape[3], const ulk_seg_params *prm, int keep_labels,
                           ulk_frame_nodes **result)
{
    UlkDims dm;
    dm.init(shape[0], shape[1], shape[2]);
    if (dm.V <= 0) ULK_FAIL(-1, "empty frame");
    if (dm.D > 65535 || dm.H > 65535) ULK_FAIL(-1, "frame too large: depth and height are limited to 65535");
    if (dm.V > 1431655000ll) ULK_FAIL(-1, "frame too large: %lld voxels (edge keys are 32-bit)", (long long)dm.V);
    if (prm->ws_hierarchy < 0 || prm->ws_hierarchy > 2)
        ULK_FAIL(-1, "ws_hierarchy=%d is not supported (0 area, 1 dynamics, 2 volume)", prm->ws_hierarchy);
    const i64 V = dm.V;
    ulk_seg_stats stats_local;
    memset(&stats_local, 0, sizeof(stats_local));
    UlkArena A;
    ulk_frame_nodes *F = new ulk_frame_nodes();
    memset(F, 0, sizeof(*F));
    F->shape[0] = dm.D; F->shape[1] = dm.H; F->shape[2] = dm.W;
    *result = F;

    // ---- K1 threshold -> foreground words -> ranks
    const i32 SG = (dm.W + 31) / 32;
    const i64 nt = (i64)dm.D * dm.H * SG;
    u32 *words = A.get<u32>(nt);
    i32 *off = A.get<i32>(nt);
    UlkOW *ow = A.get<UlkOW>(nt);
    u32 *cmax = A.get<u32>(2);           // [0] largest |contour| bits  [1] the tile front end met a value outside its fast range
    CHKA(A);
    CHK(ulk_memset(cmax, 0, 8, st));
    CHK(threshold_words(st, dm, fgp, fg_dtype, prm->threshold, SG, nt, words));
    CHK(ulk_exclusive_scan_popc_u32(words, off, nt, st));
    i32 h_last_off = 0; u32 h_last_w = 0;
    CHK(ulk_read2(&h_last_off, off + (nt - 1), 1, &h_last_w, words + (nt - 1), 1, st));
    const i32 nfg = h_last_off + __builtin_popcount(h_last_w);
    stats_local.n_foreground = nfg;
    F->fix_shift = ulk_fix_shift_for(0);
    if (nfg == 0) { F->n_nodes = 0; F->stats = stats_local; return empty_labels(st, F, V, keep_labels); }
    // ---- rank space: the foreground list, its contour values, the component forest, and the
    //      steepest edge of every voxel in component space (front end, ulk_front.cu)
    i32 *fgl = A.get<i32>(nfg);
    i32 *P = A.get<i32>(nfg);
    i32 *cnt = A.get<i32>(nfg);          // sizes at the roots, later the component code at the roots
    u32 *mw = A.get<u32>(nfg), *pk = A.get<u32>(nfg);
    i32 *po = A.get<i32>(nfg);
    UlkCt ct;
    ct.f = nullptr; ct.d = nullptr;
    float *ctf = nullptr; double *ctd = nullptr;
    if (ulk_ct_is_f32(ct_dtype)) { ctf = A.get<float>(nfg); ct.f = ctf; }
    else { ctd = A.get<double>(nfg); ct.d = ctd; }
    CHKA(A);
    UlkFg fg;
    fg.ow = ow; fg.fgl = fgl; fg.n = nfg; fg.S = SG; fg.SH = SG * dm.H; fg.dm = dm;
    if (ulk_front_tile_supported(ctp, ct_dtype, dm)) {
        CHK(ulk_front_tile(st, ctp, ct_dtype, dm, SG, words, off, ow, fgl, ctf, P, mw, pk, po, cmax));
    } else {
        expand_words(st, K_Expand{words, off, dm, SG, ctp, ct_dtype, ow, fgl, ctf, ctd, P, cmax}, nt);
        ulk_launch_rkx<1>("MinEdge", nullptr, fgl, nfg, dm, st, K_MinEdge{nullptr, ct, fg, mw, pk, po});
    }
    A.release(words); A.release(off);
    // ---- connected components (4/6-connected): union of runs, flatten + sizes + root list
    UlkFastDiv fS, fH;
    fS.init((u32)SG); fH.init((u32)dm.H);
    ulk_launch(nt, st, K_CclWords{fg, fS, fH, P, cnt});
    i32 *lab2 = P;                       // region labels are written in place over the component forest
    i32 *ctr = A.get<i32>(4);
    u32 *roots0 = A.get<u32>(nfg);
    CHKA(A);
    CHK(ulk_memset(ctr, 0, 4 * sizeof(i32), st));
    ulk_launch_rk_multi(nullptr, nfg, st, K_FlattenCount{P, cnt, (i32 *)roots0, ctr});
    i32 hc[4];
    u32 h_cm[2] = {0, 0};
    CHK(ulk_read2(hc, ctr, 4, h_cm, cmax, 2, st));
    const u32 h_cmax = h_cm[0];
    // a foreground contour value outside the tile kernel's exact float range (negative, NaN / inf,
    // below 2^-23 or above 2^39): the steepest edges are recomputed in the general form
    if (h_cm[1]) ulk_launch_rkx<1>("MinEdge", nullptr, fgl, nfg, dm, st, K_MinEdge{nullptr, ct, fg, mw, pk, po});
    const i32 n_roots = hc[0];
    const i32 max_area_i = prm->max_area > 2147483646.0 ? 2147483646 : (i32)prm->max_area;
    const int fix_shift = ulk_fix_shift_for(h_cmax);
    F->fix_shift = fix_shift;
    // ---- per-component decisions on the (sorted) root list: small-object removal
    //      (hierarchy.py:64-73), components above max_area, consecutive numbering
    u32 *roots = A.get<u32>(n_roots);
    i32 *keep = A.get<i32>(n_roots + 1), *keeppre = A.get<i32>(n_roots + 1);
    i32 *big = A.get<i32>(n_roots + 1), *bigpre = A.get<i32>(n_roots + 1);
    i32 *rsize = A.get<i32>(n_roots), *cc_big = A.get<i32>(n_roots), *big_cc = A.get<i32>(n_roots);
    CHKA(A);
    CHK(ulk_sort_keys_u32(roots0, roots, n_roots, st));
    A.release(roots0);
    i32 max_size = (i32)(prm->min_area / prm->min_area_factor) - 1;
    int do_remove = n_roots > 1 && max_size >= 1;
    ulk_launch(n_roots, st, K_RootFlags{roots, cnt, max_size, do_remove, max_area_i, keep, big, rsize});
    CHK(ulk_memset(keep + n_roots, 0, sizeof(i32), st));
    CHK(ulk_memset(big + n_roots, 0, sizeof(i32), st));
    CHK(ulk_exclusive_scan_i32(keep, keeppre, (i64)n_roots + 1, st));
    CHK(ulk_exclusive_scan_i32(big, bigpre, (i64)n_roots + 1, st));
    i32 n_cc = 0, n_big = 0;
    CHK(ulk_read2(&n_cc, keeppre + n_roots, 1, &n_big, bigpre + n_roots, 1, st));
    if (n_cc == 0) { F->n_nodes = 0; F->stats = stats_local; return empty_labels(st, F, V, keep_labels); }
    i32 *code = cnt;
    ulk_launch(n_roots, st, K_RootCode{roots, keep, keeppre, big, bigpre, code, cc_big, big_cc, n_big > 0 ? 1 : 0});
    A.release(roots); A.release(keep); A.release(keeppre); A.release(big); A.release(bigpre); A.release(rsize);

    // ---- region labels: components, oversegmented where area > max_area
    i32 R = n_cc;
    i32 *rlA = nullptr;
    i32 nA = 0;
    if (n_big == 0) {
        ulk_launch_rkn<4>(nullptr, nfg, st, K_LabelCc{code, 0, lab2, nullptr, nullptr, po});
    } else {
        UlkArena AA;
        // the labels of the oversegmentation pass live in place too: > 0 on the components above
        // max_area, negative everywhere else
        i32 *labA = lab2;
        // pass A only ever touches the voxels of the components above max_area: its kernels run
        // on that sub-list of ranks, collected while the labels are written (chunks of up to 32
        // raster-ordered voxels in arbitrary order: order only matters for coalescing)
        rlA = A.get<i32>(nfg);           // kept for pass B: its steepest edges are recomputed with the region labels
        i32 *nA_dev = AA.get<i32>(1);
        CHKA(AA); CHKA(A);
        CHK(ulk_memset(nA_dev, 0, sizeof(i32), st));
#ifdef ULK_EMU
        ulk_launch_rkn<4>(nullptr, nfg, st, K_LabelCc{code, 1, labA, rlA, nA_dev, po});
#else
        ulk_launch_rk_multi(nullptr, nfg, st, K_LabelCc{code, 1, labA, rlA, nA_dev, po});
#endif
        CHK(ulk_read(&nA, nA_dev, 1, st));
        stats_local.pass_a_voxels = nA;
        VoxelStage va;
        CHK(voxel_stage(st, AA, fg, rlA, nA, ct, labA, n_big, true, false, false, fix_shift, mw, pk, po, nullptr, 0, va, &stats_local));
        UlkHierScratch S;
        CHK(alloc_scratch(AA, va.B, va.Eg, S));
        UlkBasins Bv{va.b_size, va.g_L, va.b_top, nullptr, nullptr, nullptr, nullptr};
        UlkEdges Ev{va.ea, va.eb, va.wk, va.k3, va.cnt, va.shi, va.slo};
        ulk_launch(n_big, st, K_HierStage1{va.bs, va.es, Bv, Ev, S});
        UlkAttachInfo *ainfoA = AA.get<UlkAttachInfo>(va.B);
        CHKA(AA);
        ulk_launch(va.B, st, K_AttachInfo{va.g_lab, va.bs, S.Ipar, S.Iwk, ainfoA});
        ulk_launch_rkn<4>(rlA, nA, st, K_Attach{va.bas, va.mw, va.jk, ainfoA, S.Ipar, S.Iwk, S.attach, nullptr, nullptr, 0.0});
        i32 *region_of = AA.get<i32>(va.B);
        u64 *rkey_w = AA.get<u64>(va.B), *rkey_k = AA.get<u64>(va.B);
        CHKA(AA);
        ulk_launch(n_big, st, K_HierCut{va.bs, va.es, Bv, Ev, S, prm->max_area, region_of, rkey_w, rkey_k});
        i64 nrec = (i64)n_cc + va.B;
        u32 *rec_cc = AA.get<u32>(nrec), *perm0 = AA.get<u32>(nrec), *perm1 = AA.get<u32>(nrec);
        u64 *rec_w = AA.get<u64>(nrec), *rec_k = AA.get<u64>(nrec), *k0 = AA.get<u64>(nrec), *k1 = AA.get<u64>(nrec);
        u32 *c0 = AA.get<u32>(nrec), *c1 = AA.get<u32>(nrec);
        i32 *label_of_rec = AA.get<i32>(nrec);
        CHKA(AA);
        ulk_launch(n_cc, st, K_SmallCcRecords{cc_big, rec_cc, rec_w, rec_k});
        ulk_launch(va.B, st, K_RegionRecords{region_of, rkey_w, rkey_k, va.g_lab, big_cc, n_cc, rec_cc, rec_w, rec_k});
        ulk_launch(nrec, st, K_Iota{perm0});
        CHK(ulk_sort_pairs_u64(rec_k, k1, perm0, perm1, nrec, st));           // by tie key
        ulk_launch(nrec, st, K_Gather<u64>{rec_w, perm1, k0});
        CHK(ulk_sort_pairs_u64(k0, k1, perm1, perm0, nrec, st));              // then by altitude
        ulk_launch(nrec, st, K_Gather<u32>{rec_cc, perm0, c0});
        CHK(ulk_sort_pairs_u32(c0, c1, perm0, perm1, nrec, st));              // then by component
        ulk_launch(nrec, st, K_RankToLabel{perm1, c1, label_of_rec});
        ulk_launch_rkn<4>(nullptr, nfg, st, K_WriteLabels{n_cc, va.bas, region_of, label_of_rec, lab2, po});
        // number of valid records = regions
        i32 *nvalid = AA.get<i32>(nrec + 1), *nvpre = AA.get<i32>(nrec + 1);
        CHKA(AA);
        ulk_launch(nrec, st, K_Valid{rec_cc, nvalid});
        CHK(ulk_memset(nvalid + nrec, 0, sizeof(i32), st));
        CHK(ulk_exclusive_scan_i32(nvalid, nvpre, nrec + 1, st));
        CHK(ulk_read(&R, nvpre + nrec, 1, st));      // synchronises: pass A's scratch can go
    }
    A.release(cc_big); A.release(big_cc); A.release(cnt);
    if (keep_labels) {
        F->labels = keep_alloc<i32>(V);
        if (!F->labels) return -3;
        CHK(ulk_memset(F->labels, 0, (size_t)V * sizeof(i32), st));
        ulk_launch(nfg, st, K_ScatterLabels{lab2, fgl, F->labels});
    }
    F->n_hier = R;

    // ---- pass B: the configured hierarchy on every region
    const bool area_mode = prm->ws_hierarchy == 0 || prm->ws_hierarchy == 2;    // join keys: area and volume attributes
    const bool vol_mode = prm->ws_hierarchy == 2;
    VoxelStage vb;
    CHK(voxel_stage(st, A, fg, nullptr, nfg, ct, lab2, R, area_mode, true, true, fix_shift, mw, pk, po, rlA, nA, vb, &stats_local));
    if (rlA) A.release(rlA);
    UlkHierScratch S;
    CHK(alloc_scratch(A, vb.B, vb.Eg, S));
    UlkBasins Bv{vb.b_size, vb.g_L, vb.b_top, vb.b_k3, vb.bbmin, vb.bbmax, vb.csum};
    UlkEdges Ev{vb.ea, vb.eb, vb.wk, vb.k3, vb.cnt, vb.shi, vb.slo};
    ulk_launch(R, st, K_HierStage1{vb.bs, vb.es, Bv, Ev, S});
    if (area_mode) {
        UlkAttachInfo *ainfo = A.get<UlkAttachInfo>(vb.B);
        CHKA(A);
        ulk_launch(vb.B, st, K_AttachInfo{vb.g_lab, vb.bs, S.Ipar, S.Iwk, ainfo});
        K_Attach ka{vb.bas, vb.mw, vb.jk, ainfo, S.Ipar, S.Iwk, S.attach,
                    vol_mode ? S.awhi : nullptr, vol_mode ? S.awlo : nullptr, ldexp(1.0, fix_shift)};
#ifndef ULK_EMU
        if (!vol_mode) {
            if (g_ulk_prof_on) ulk_prof_begin(typeid(K_Attach).name(), st);
            ulk_attach_vec4_kernel<<<(unsigned)(((i64)nfg + 1023) / 1024), 256, 0, st>>>(ka, (i64)nfg);
            if (g_ulk_prof_on) ulk_prof_end(st);
            ++g_ulk_launches;
            if (g_ulk_debug_sync) ulk_debug_after_launch("K_Attach(vec4)", st);
        } else
#endif
            ulk_launch_rkn<4>(nullptr, nfg, st, ka);
        A.release(ainfo);
    }
    i64 N2 = 2 * (i64)vb.B + 2;
    UlkNodeOut O;
    O.count = A.get<i32>(R + 1); O.nsurv = A.get<i32>(R + 1); O.rsize = A.get<i32>(R + 1);
    O.area = A.get<i32>(N2); O.frontier = A.get<double>(N2); O.height = A.get<double>(N2);
    O.bbmin = A.get<i32>(3 * N2); O.bbmax = A.get<i32>(3 * N2); O.csum = A.get<i64>(3 * N2);
    O.parent = A.get<i32>(N2); O.hidx = A.get<i32>(N2); O.basin_node = A.get<i32>(vb.B);
    u64 *err = A.get<u64>(1);
    CHKA(A);
    CHK(ulk_memset(err, 0xff, sizeof(u64), st));
    CHK(ulk_memset(O.count + R, 0, sizeof(i32), st));
    UlkHierParams HP;
    HP.kind = prm->ws_hierarchy; HP.min_area = prm->min_area; HP.max_area = prm->max_area;
    HP.min_frontier = prm->min_frontier; HP.fix_shift = fix_shift;
    ulk_launch(R, st, K_HierFull{vb.bs, vb.es, Bv, Ev, S, HP, O});
    ulk_launch(R, st, K_RegionSizeCheck{vb.bs, S, err});
    // ---- node table
    i32 *noff = A.get<i32>(R + 1);
    CHKA(A);
    CHK(ulk_exclusive_scan_i32(O.count, noff, R + 1, st));
    u64 h_err = 0;
    i32 N = 0;
    CHK(ulk_read2(&h_err, err, 1, &N, noff + R, 1, st));
    if (h_err != ULK_NONE64) ULK_FAIL(-10, "Region too small. Size of %d found.", (int)(h_err & 0xff));
    F->n_nodes = N;
    F->bbox = keep_alloc<i32>(6 * (i64)N); F->centroid = keep_alloc<i32>(3 * (i64)N);
    F->area = keep_alloc<i32>(N); F->frontier = keep_alloc<double>(N); F->height = keep_alloc<double>(N);
    F->hier_id = keep_alloc<i32>(N); F->parent = keep_alloc<i32>(N); F->hidx = keep_alloc<i32>(N);
    F->mask_off = keep_alloc<i64>(N + 1);
    i64 *mwords = A.get<i64>(N + 1);
    i32 *bnode = A.get<i32>(vb.B);
    CHKA(A);
    if (!F->bbox || !F->centroid || !F->area || !F->frontier || !F->height || !F->hier_id || !F->parent || !F->hidx || !F->mask_off)
        ULK_FAIL(-3, "device allocation failed");
    ulk_launch(R, st, K_NodeGather{vb.bs, noff, O, F->bbox, F->centroid, F->area, F->hier_id, F->parent, F->hidx,
                                   F->frontier, F->height, mwords});
    CHK(ulk_memset(mwords + N, 0, sizeof(i64), st));
    CHK(ulk_exclusive_scan_i64(mwords, F->mask_off, (i64)N + 1, st));
    // overlaps = (node, emitted proper ancestor) pairs, processing.py:289-296: their count comes
    // back together with the mask size (one host round trip)
    i32 *depth = A.get<i32>(N + 1), *ooff = A.get<i32>(N + 1);
    CHKA(A);
    ulk_launch(N, st, K_AncestorCount{F->parent, depth});
    CHK(ulk_memset(depth + N, 0, sizeof(i32), st));
    CHK(ulk_exclusive_scan_i32(depth, ooff, (i64)N + 1, st));
    i64 total_words = 0;
    i32 M = 0;
    CHK(ulk_read2(&total_words, F->mask_off + N, 1, &M, ooff + N, 1, st));
    F->mask_words = total_words;
    F->mask = keep_alloc<u32>(total_words);
    if (!F->mask) ULK_FAIL(-3, "device allocation failed");
    CHK(ulk_memset(F->mask, 0, (size_t)total_words * 4, st));
    ulk_launch(vb.B, st, K_BasinNodeGlobal{vb.g_lab, noff, O.basin_node, bnode});
    UlkPaintInfo *pinfo = A.get<UlkPaintInfo>(N);
    CHKA(A);
    ulk_launch(N, st, K_PaintInfo{F->parent, F->bbox, F->mask_off, pinfo});
    ulk_launch_rkx<4>("MaskPaint", nullptr, fgl, nfg, dm, st, K_MaskPaint{vb.bas, bnode, pinfo, dm, F->mask});
    F->n_overlaps = M;
    F->overlaps = keep_alloc<i32>(2 * (i64)M);
    if (!F->overlaps) ULK_FAIL(-3, "device allocation failed");
    ulk_launch(N, st, K_OverlapWrite{F->parent, ooff, F->overlaps});
    CHK(ulk_sync(st));
    F->stats = stats_local;
    return 0;
}
