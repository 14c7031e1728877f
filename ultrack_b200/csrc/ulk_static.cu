// ulk_static.cu -- ulk_smooth_contours_static: the contour-smoothing step for callers that
// cannot allocate (code under CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc).  Same arithmetic as ulk_smooth_contours (ulk_aux.cu: separable Gaussian with
// scipy's "reflect" boundary, float64 taps, float32 per axis, then division by the global
// maximum), but
//   * the ping-pong volume and the maximum live in ONE file-scope device array of fixed size
//     (ULK_SMOOTH_STATIC_MAX_VOXELS float32 + a 256-byte header), resident on every GPU that loads
//     the library until the process exits;
//   * the Gaussian taps travel in the kernel parameters (radius <= ULK_SMOOTH_STATIC_MAX_RADIUS),
//     so nothing is copied from the host;
//   * the maximum is never read back: the division kernel takes it from the scratch header;
//   * the call only enqueues work on `stream` and does not synchronise.
// No cudaMalloc / cudaFree / pool / workspace call is made on this path.  Every call on a device
// shares the array: calls that overlap on the GPU (two streams, two host threads) race.
// This entry point has no counterpart in the
// reference.
#include "ulk_seg.h"
#include "ulk_common.h"
#include "../../include/ultrack_b200.h"
#include <cmath>
#include <mutex>

#define ULK_STATIC_HEADER_BYTES 256
#define ULK_STATIC_SCRATCH_BYTES (ULK_STATIC_HEADER_BYTES + 4ull * ULK_SMOOTH_STATIC_MAX_VOXELS)

#ifndef ULK_EMU
static __device__ __align__(256) unsigned char g_ulk_static_scratch[ULK_STATIC_SCRATCH_BYTES];
#else
alignas(256) static unsigned char g_ulk_static_scratch[ULK_STATIC_SCRATCH_BYTES];
#endif

// same taps, same accumulation order as K_Gauss1D (ulk_aux.cu); hw[j] = normalised weight at
// offset +-j
struct K_StaticGauss1D {
    const float *in; float *out; i32 r; UlkDims dm; int axis;   // axis: 0 = x, 1 = y, 2 = z
    double hw[ULK_SMOOTH_STATIC_MAX_RADIUS + 1];
    ULK_D void operator()(i64 vv) const
    {
        i32 z, y, x;
        dm.coords((i32)vv, z, y, x);
        const i32 n = axis == 0 ? dm.W : (axis == 1 ? dm.H : dm.D);
        const i32 c = axis == 0 ? x : (axis == 1 ? y : z);
        const i64 st = dm.stride(axis);
        const float *base = in + (vv - (i64)c * st);
        auto at = [&](i32 k) {
            i32 p = k % (2 * n);
            if (p < 0) p += 2 * n;
            if (p >= n) p = 2 * n - 1 - p;
            return (double)base[(i64)p * st];
        };
        double acc = at(c) * hw[0];
        for (i32 j = r; j >= 1; --j) acc += (at(c - j) + at(c + j)) * hw[j];
        out[vv] = (float)acc;
    }
};
struct K_StaticClear { u32 *mx; ULK_D void operator()(i64) const { *mx = 0u; } };
struct K_StaticMax {
    const float *p; u32 *mx;
    ULK_D void operator()(i64 i) const
    {
        const float v = p[i];
        if (v > 0.0f) { u32 b; memcpy(&b, &v, 4); if (b > *mx) ulk_atomic_max(mx, b); }
    }
};
// out = in / max (out may equal in); the maximum comes from the scratch header
struct K_StaticDiv {
    const float *in; float *out; const u32 *mx;
    ULK_D void operator()(i64 i) const
    {
        const u32 b = *mx;
        float d;
        memcpy(&d, &b, 4);
        out[i] = in[i] / d;
    }
};

// plain launches: no profiling events, no debug synchronise
template <class F> static int static_launch(i64 n, ulk_stream_t st, const F &f)
{
#ifndef ULK_EMU
    ulk_kernel<F><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, f);
    ++g_ulk_launches;
    ULK_CUDA(cudaGetLastError());
#else
    (void)st;
    for (i64 i = 0; i < n; ++i) f(i);
#endif
    return 0;
}

// Device address of the scratch array, looked up once per device with that device current.
// In host code `g_ulk_static_scratch` is only a handle for the runtime, never a device pointer.
#define ULK_STATIC_MAX_DEVICES 64
static unsigned char *g_static_ptr[ULK_STATIC_MAX_DEVICES];
static std::mutex g_static_mu;
static int static_scratch(unsigned char **out)
{
#ifndef ULK_EMU
    int dev = -1;
    ULK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= ULK_STATIC_MAX_DEVICES) ULK_FAIL(-1, "ulk_smooth_contours_static: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[dev]) {
        void *p = nullptr;
        ULK_CUDA(cudaGetSymbolAddress(&p, g_ulk_static_scratch));
        if (!p) ULK_FAIL(-2, "ulk_smooth_contours_static: no device address for the static scratch");
        g_static_ptr[dev] = (unsigned char *)p;
    }
    *out = g_static_ptr[dev];
#else
    *out = g_ulk_static_scratch;
#endif
    return 0;
}

extern "C" int64_t ulk_smooth_contours_static_max_voxels(void) { return (int64_t)ULK_SMOOTH_STATIC_MAX_VOXELS; }

extern "C" int ulk_smooth_contours_static(float *contours, const int64_t shape[3], const double sigma[3], void *stream)
{
    if (!contours || !shape || !sigma) ULK_FAIL(-1, "null argument");
    if (shape[0] <= 0 || shape[1] <= 0 || shape[2] <= 0) ULK_FAIL(-1, "ulk_smooth_contours_static: empty shape");
    const double nvox = (double)shape[0] * (double)shape[1] * (double)shape[2];
    if (nvox > (double)ULK_SMOOTH_STATIC_MAX_VOXELS)
        ULK_FAIL(-12, "ulk_smooth_contours_static: %lld x %lld x %lld voxels exceed the static scratch limit of %lld voxels "
                      "(use ulk_smooth_contours)", (long long)shape[0], (long long)shape[1], (long long)shape[2],
                 (long long)ULK_SMOOTH_STATIC_MAX_VOXELS);
    UlkDims dm;
    dm.init(shape[0], shape[1], shape[2]);
    for (int ax = 2; ax >= 0; --ax) {
        const double sg = sigma[2 - ax];
        const i64 n = ax == 0 ? dm.W : (ax == 1 ? dm.H : dm.D);
        if (!(sg > 0.0) || n <= 1) continue;
        if (!(4.0 * sg + 0.5 < (double)(ULK_SMOOTH_STATIC_MAX_RADIUS + 1)))
            ULK_FAIL(-12, "ulk_smooth_contours_static: sigma %g needs a window radius above the limit of %d taps "
                          "(use ulk_smooth_contours)", sg, ULK_SMOOTH_STATIC_MAX_RADIUS);
    }
    ulk_stream_t st = (ulk_stream_t)stream;
    unsigned char *scratch = nullptr;
    int rc = static_scratch(&scratch);
    if (rc) return rc;
    u32 *mx = (u32 *)scratch;
    float *tmp = (float *)(scratch + ULK_STATIC_HEADER_BYTES);
    // the array keeps the previous call's contents: the maximum must start at zero; the volume
    // part is fully overwritten by a pass before it is read
    rc = static_launch(1, st, K_StaticClear{mx});
    if (rc) return rc;
    const float *src = contours;
    float *dst = tmp;
    for (int ax = 2; ax >= 0; --ax) {          // scipy filters axis 0 (z) first
        const double sg = sigma[2 - ax];
        const i64 n = ax == 0 ? dm.W : (ax == 1 ? dm.H : dm.D);
        if (!(sg > 0.0) || n <= 1) continue;
        const i32 r = (i32)(4.0 * sg + 0.5);
        double w[2 * ULK_SMOOTH_STATIC_MAX_RADIUS + 1];
        double sum = 0.0;
        for (i32 k = -r; k <= r; ++k) { w[k + r] = exp(-0.5 / (sg * sg) * (double)k * (double)k); sum += w[k + r]; }
        K_StaticGauss1D g;
        g.in = src; g.out = dst; g.r = r; g.dm = dm; g.axis = ax;
        for (i32 j = 0; j <= ULK_SMOOTH_STATIC_MAX_RADIUS; ++j) g.hw[j] = j <= r ? w[r - j] / sum : 0.0;
        rc = static_launch(dm.V, st, g);
        if (rc) return rc;
        float *t = (float *)src; src = dst; dst = t;
    }
    rc = static_launch(dm.V, st, K_StaticMax{src, mx});
    if (rc) return rc;
    return static_launch(dm.V, st, K_StaticDiv{src, contours, mx});
}
