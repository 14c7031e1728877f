// ulk_common.h -- shared geometry / dtype helpers for the voxel-scale kernels.
#pragma once
#include "ulk_rt.h"
#include <vector>

#define ULK_NONE32 0xFFFFFFFFu
#define ULK_NONE64 ((u64)0xFFFFFFFFFFFFFFFFull)

// dtype codes of the C ABI (include/ultrack_b200.h)
enum { ULK_U8 = 0, ULK_I8 = 1, ULK_U16 = 2, ULK_I16 = 3, ULK_F16 = 4, ULK_F32 = 5, ULK_F64 = 6,
       ULK_I32 = 7, ULK_U32 = 8, ULK_I64 = 9, ULK_U64 = 10 };

ULK_HD ULK_INLINE float ulk_half_to_float(uint16_t h)
{
    u32 sign = (u32)(h & 0x8000u) << 16;
    u32 exp = (h >> 10) & 0x1f, man = h & 0x3ffu, bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else {
            int e = -1;
            do { ++e; man <<= 1; } while (!(man & 0x400u));
            bits = sign | ((u32)(127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
        }
    } else if (exp == 31) bits = sign | 0x7f800000u | (man << 13);
    else bits = sign | ((exp + 112) << 23) | (man << 13);
    float f;
    memcpy(&f, &bits, 4);
    return f;
}

ULK_HD ULK_INLINE double ulk_load_as_double(const void *p, int dt, i64 i)
{
    switch (dt) {
    case ULK_U8: return (double)((const uint8_t *)p)[i];
    case ULK_I8: return (double)((const int8_t *)p)[i];
    case ULK_U16: return (double)((const uint16_t *)p)[i];
    case ULK_I16: return (double)((const int16_t *)p)[i];
    case ULK_F16: return (double)ulk_half_to_float(((const uint16_t *)p)[i]);
    case ULK_F32: return (double)((const float *)p)[i];
    case ULK_F64: return ((const double *)p)[i];
    case ULK_I32: return (double)((const i32 *)p)[i];
    case ULK_U32: return (double)((const u32 *)p)[i];
    case ULK_I64: return (double)((const i64 *)p)[i];
    default: return (double)((const u64 *)p)[i];
    }
}
inline int ulk_dtype_size(int dt)
{
    switch (dt) {
    case ULK_U8: case ULK_I8: return 1;
    case ULK_U16: case ULK_I16: case ULK_F16: return 2;
    case ULK_F32: case ULK_I32: case ULK_U32: return 4;
    default: return 8;
    }
}
// contour dtypes whose pairwise sum numba evaluates in float32 (exactly or by its typing rules)
inline bool ulk_ct_is_f32(int dt) { return dt == ULK_U8 || dt == ULK_I8 || dt == ULK_U16 || dt == ULK_I16 || dt == ULK_F16 || dt == ULK_F32; }

// division of a u32 by a run-time constant without the divide unit (Granlund-Montgomery,
// branch-free form): q = (t + ((v - t) >> s1)) >> s2 with t = hi32(M * v)
struct UlkFastDiv {
    u32 M, s1, s2;
    void init(u32 d)
    {
        u32 l = 0;
        while (((u64)1 << l) < d) ++l;
        M = (u32)((((u64)1 << 32) * (((u64)1 << l) - d)) / d + 1);
        s1 = l < 1 ? l : 1;
        s2 = l > 1 ? l - 1 : 0;
    }
    ULK_HD ULK_INLINE u32 div(u32 v) const
    {
#if defined(__CUDA_ARCH__)
        u32 t = __umulhi(M, v);
#else
        u32 t = (u32)(((u64)M * v) >> 32);
#endif
        return (t + ((v - t) >> s1)) >> s2;
    }
};
struct UlkDims {
    i32 D, H, W;
    i64 V;
    UlkFastDiv fW, fH;
    void init(i64 d, i64 h, i64 w)
    {
        D = (i32)d; H = (i32)h; W = (i32)w;
        V = d * h * w;
        fW.init((u32)(W > 0 ? W : 1)); fH.init((u32)(H > 0 ? H : 1));
    }
    ULK_HD ULK_INLINE i32 stride(int d) const { return d == 0 ? 1 : (d == 1 ? W : W * H); }
    ULK_HD ULK_INLINE void coords(i32 v, i32 &z, i32 &y, i32 &x) const
    {
        u32 t = fW.div((u32)v);
        x = (i32)((u32)v - t * (u32)W);
        u32 q = fH.div(t);
        y = (i32)(t - q * (u32)H);
        z = (i32)q;
    }
    // can voxel with coords c step forward / backward along direction d (0=x,1=y,2=z)
    ULK_HD ULK_INLINE bool fwd(int d, i32 z, i32 y, i32 x) const { return d == 0 ? x + 1 < W : (d == 1 ? y + 1 < H : z + 1 < D); }
    ULK_HD ULK_INLINE bool bwd(int d, i32 z, i32 y, i32 x) const { return d == 0 ? x > 0 : (d == 1 ? y > 0 : z > 0); }
};

// weight of an edge from the float32 sum of its endpoints' contours: for s == 0 and
// 2^-24 <= |s| < 2^40 the float expression equals the double one bit for bit (checked
// exhaustively over every float of that range), which keeps float<->double conversions off the
// dense kernels
ULK_HD ULK_INLINE bool ulk_wsum_fast(float s)
{
    u32 b;
    memcpy(&b, &s, 4);
    u32 e = (b >> 23) & 0xffu;
    return s == 0.0f || (e - 103u) < 64u;
}
ULK_HD ULK_INLINE float ulk_wsum_slow(float s) { return (float)((double)s / 2.0 + 1e-8); }
ULK_HD ULK_INLINE float ulk_wsum(float s) { return ulk_wsum_fast(s) ? s * 0.5f + 1e-8f : ulk_wsum_slow(s); }
struct UlkCt {
    const float *f;
    const double *d;
    ULK_HD ULK_INLINE float w(i32 p, i32 q) const
    {
        if (f) return ulk_wsum(f[p] + f[q]);
        double s = d[p] + d[q];
        return (float)(s / 2.0 + 1e-8);
    }
};

// frame-scoped device allocations
struct UlkArena {
    std::vector<void *> ptrs;
    bool failed = false;
    template <class T> T *get(i64 n)
    {
        void *p = nullptr;
        if (ulk_ws_alloc(&p, (size_t)(n > 0 ? n : 1) * sizeof(T))) { failed = true; return nullptr; }
        ptrs.push_back(p);
        return (T *)p;
    }
    void release(void *p)
    {
        for (size_t i = 0; i < ptrs.size(); ++i)
            if (ptrs[i] == p) { ulk_ws_free(p); ptrs[i] = ptrs.back(); ptrs.pop_back(); return; }
    }
    ~UlkArena() { for (void *p : ptrs) ulk_ws_free(p); }
};

template <class T> struct K_Fill {
    T *p; T v;
    ULK_HD void operator()(i64 i) const { p[i] = v; }
};
template <class T> int ulk_fill(T *p, T v, i64 n, ulk_stream_t s) { ulk_launch(n, s, K_Fill<T>{p, v}); return 0; }

template <class T> int ulk_read(T *host, const T *dev, i64 n, ulk_stream_t s) { return ulk_d2h(host, dev, (size_t)n * sizeof(T), s); }
template <class T, class U> int ulk_read2(T *h1, const T *d1, i64 n1, U *h2, const U *d2, i64 n2, ulk_stream_t s)
{
    return ulk_d2h2(h1, d1, (size_t)n1 * sizeof(T), h2, d2, (size_t)n2 * sizeof(U), s);
}
