// ulk_rt.h -- thin runtime layer under the ultrack_b200 kernels.
//
// Product build: nvcc, sm_100a; kernels are functor structs launched through
// ulk_launch<> (named kernels, 256-thread CTAs), device-wide sort / scan come
// from CUB (CUDA toolkit headers), memory from the stream-ordered allocator.
//
// ULK_EMU build (tests/emu only, never loaded by the package): the same functor
// bodies are compiled by g++ and executed thread by thread on the host so the
// kernel LOGIC can be checked against the oracle in the CPU-only test tier.
// It is a test harness, not a fallback: ultrack_b200/_lib.py only ever loads the
// CUDA library and raises when it is missing.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#ifdef ULK_EMU
#include <algorithm>
#include <numeric>
#include <vector>
#include <cmath>
#include <type_traits>
#define ULK_HD
#define ULK_D
#define ULK_INLINE inline
typedef void *ulk_stream_t;
#else
#include <cuda_runtime.h>
#include <cub/cub.cuh>
#include <typeinfo>
#define ULK_HD __host__ __device__
#define ULK_D __device__
#define ULK_INLINE __forceinline__
typedef cudaStream_t ulk_stream_t;
#endif

typedef int64_t i64;
typedef uint64_t u64;
typedef uint32_t u32;
typedef int32_t i32;
typedef uint8_t u8;

// a phased rank-space functor may declare S0 load0(r, ok): the loads that need the rank only
// (see ulk_kernel_rk); load1 then takes the S0 as its first argument
template <class F, class = void> struct ulk_has_load0 : std::false_type {};
template <class F> struct ulk_has_load0<F, std::void_t<typename F::S0>> : std::true_type {};

// ---------------------------------------------------------------- errors
void ulk_set_error(const char *fmt, ...);
#define ULK_FAIL(code, ...) do { ulk_set_error(__VA_ARGS__); return (code); } while (0)

#ifndef ULK_EMU
#define ULK_CUDA(call)                                                              \
    do {                                                                            \
        cudaError_t e__ = (call);                                                   \
        if (e__ != cudaSuccess) {                                                   \
            ulk_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call,              \
                          cudaGetErrorString(e__));                                 \
            return -2;                                                              \
        }                                                                           \
    } while (0)
#else
#define ULK_CUDA(call) do { (void)0; } while (0)
#endif

// ---------------------------------------------------------------- atomics
#ifdef ULK_EMU
template <class T> ULK_INLINE T ulk_atomic_min(T *p, T v) { T o = *p; if (v < o) *p = v; return o; }
template <class T> ULK_INLINE T ulk_atomic_max(T *p, T v) { T o = *p; if (v > o) *p = v; return o; }
template <class T> ULK_INLINE T ulk_atomic_add(T *p, T v) { T o = *p; *p = o + v; return o; }
template <class T> ULK_INLINE T ulk_atomic_or(T *p, T v) { T o = *p; *p = o | v; return o; }
template <class T> ULK_INLINE T ulk_atomic_cas(T *p, T c, T v) { T o = *p; if (o == c) *p = v; return o; }
template <class T> ULK_INLINE T ulk_atomic_exch(T *p, T v) { T o = *p; *p = v; return o; }
ULK_INLINE int ulk_popc(u32 x) { return __builtin_popcount(x); }
ULK_INLINE u32 ulk_funnel_r(u32 lo, u32 hi, u32 s) { return s ? (lo >> s) | (hi << (32 - s)) : lo; }
#else
ULK_D ULK_INLINE i32 ulk_atomic_min(i32 *p, i32 v) { return atomicMin(p, v); }
ULK_D ULK_INLINE u32 ulk_atomic_min(u32 *p, u32 v) { return atomicMin(p, v); }
ULK_D ULK_INLINE u64 ulk_atomic_min(u64 *p, u64 v) { return atomicMin((unsigned long long *)p, (unsigned long long)v); }
ULK_D ULK_INLINE i32 ulk_atomic_max(i32 *p, i32 v) { return atomicMax(p, v); }
ULK_D ULK_INLINE u32 ulk_atomic_max(u32 *p, u32 v) { return atomicMax(p, v); }
ULK_D ULK_INLINE u64 ulk_atomic_max(u64 *p, u64 v) { return atomicMax((unsigned long long *)p, (unsigned long long)v); }
ULK_D ULK_INLINE i32 ulk_atomic_add(i32 *p, i32 v) { return atomicAdd(p, v); }
ULK_D ULK_INLINE u32 ulk_atomic_add(u32 *p, u32 v) { return atomicAdd(p, v); }
ULK_D ULK_INLINE u64 ulk_atomic_add(u64 *p, u64 v) { return atomicAdd((unsigned long long *)p, (unsigned long long)v); }
ULK_D ULK_INLINE i64 ulk_atomic_add(i64 *p, i64 v) { return (i64)atomicAdd((unsigned long long *)p, (unsigned long long)v); }
ULK_D ULK_INLINE u32 ulk_atomic_or(u32 *p, u32 v) { return atomicOr(p, v); }
ULK_D ULK_INLINE i32 ulk_atomic_cas(i32 *p, i32 c, i32 v) { return atomicCAS(p, c, v); }
ULK_D ULK_INLINE u32 ulk_atomic_cas(u32 *p, u32 c, u32 v) { return atomicCAS(p, c, v); }
ULK_D ULK_INLINE u64 ulk_atomic_cas(u64 *p, u64 c, u64 v) { return atomicCAS((unsigned long long *)p, (unsigned long long)c, (unsigned long long)v); }
ULK_D ULK_INLINE i32 ulk_atomic_exch(i32 *p, i32 v) { return atomicExch(p, v); }
ULK_D ULK_INLINE int ulk_popc(u32 x) { return __popc(x); }
ULK_D ULK_INLINE u32 ulk_funnel_r(u32 lo, u32 hi, u32 s) { return __funnelshift_r(lo, hi, s); }
#endif

// ---------------------------------------------------------------- launch
#ifdef ULK_EMU
template <class F> void ulk_launch(i64 n, ulk_stream_t, const F &f) { for (i64 i = 0; i < n; ++i) f(i); }
// warp-cooperative kernels: the emulation runs the functor's scalar statement `operator()`
template <class F> void ulk_launch_warp(i64 n, ulk_stream_t, const F &f) { for (i64 i = 0; i < n; ++i) f(i); }
// voxel kernels that need coordinates: f(v, z, y, x) for every voxel of a [D,H,W] frame
template <class F> void ulk_launch_vox(i32 D, i32 H, i32 W, ulk_stream_t, const F &f)
{
    i64 v = 0;
    for (i32 z = 0; z < D; ++z) for (i32 y = 0; y < H; ++y) for (i32 x = 0; x < W; ++x) f(v++, z, y, x);
}
template <class F> void ulk_launch_vox_warp(i32 D, i32 H, i32 W, ulk_stream_t s, const F &f) { ulk_launch_vox(D, H, W, s, f); }
// list kernels: f(v, z, y, x) for every voxel index of a list (the foreground voxels of a frame)
template <class F, class DM> void ulk_launch_list(const i32 *list, i64 n, const DM &dm, ulk_stream_t, const F &f)
{
    for (i64 i = 0; i < n; ++i) { i32 z, y, x; dm.coords(list[i], z, y, x); f((i64)list[i], z, y, x); }
}
template <class F, class DM> void ulk_launch_list_warp(const i32 *list, i64 n, const DM &dm, ulk_stream_t s, const F &f) { ulk_launch_list(list, n, dm, s, f); }
// phased list kernels: S1 load1(v,z,y,x,ok) -> S2 load2(S1,ok) -> run(S2,ok)
template <int U, class F, class DM> void ulk_launch_list_ph(const i32 *list, i64 n, const DM &dm, ulk_stream_t, const F &f)
{
    for (i64 i = 0; i < n; ++i) { i32 z, y, x; dm.coords(list[i], z, y, x); f.run(f.load2(f.load1(list[i], z, y, x, true), true), true); }
}
template <int U, class F, class DM> void ulk_launch_list_phx(const char *, const i32 *list, i64 n, const DM &dm, ulk_stream_t s, const F &f) { ulk_launch_list_ph<U>(list, n, dm, s, f); }
// pointer-chasing list kernels (no coordinates): the emulation runs the scalar statement
template <class F> void ulk_launch_list_multi(const i32 *list, i64 n, ulk_stream_t, const F &f)
{
    for (i64 i = 0; i < n; ++i) f((i64)list[i], 0, 0, 0);
}
// rank-space list kernels (see the CUDA forms below): entry i of a list is the foreground rank
// r = rl ? rl[i] : i, its voxel is fgl[r]
template <int U, class F, class DM> void ulk_launch_rk(const i32 *rl, const i32 *fgl, i64 m, const DM &dm, ulk_stream_t, const F &f)
{
    for (i64 i = 0; i < m; ++i) {
        i32 r = rl ? rl[i] : (i32)i, v = fgl[r], z, y, x;
        dm.coords(v, z, y, x);
        if constexpr (ulk_has_load0<F>::value) f.run(f.load2(f.load1(f.load0(r, true), r, v, z, y, x, true), true), true);
        else f.run(f.load2(f.load1(r, v, z, y, x, true), true), true);
    }
}
template <int UDEF, class F, class DM> void ulk_launch_rkx(const char *, const i32 *rl, const i32 *fgl, i64 m, const DM &dm, ulk_stream_t s, const F &f) { ulk_launch_rk<UDEF>(rl, fgl, m, dm, s, f); }
template <int U, class F> void ulk_launch_rkn(const i32 *rl, i64 m, ulk_stream_t, const F &f)
{
    for (i64 i = 0; i < m; ++i) { i32 r = rl ? rl[i] : (i32)i; f.run(f.load2(f.load1(r, true), true), true); }
}
template <class F> void ulk_launch_rk_multi(const i32 *rl, i64 m, ulk_stream_t, const F &f)
{
    for (i64 i = 0; i < m; ++i) f(rl ? rl[i] : (i32)i);
}
template <class F, class DM> void ulk_launch_rk1(const i32 *rl, const i32 *fgl, i64 m, const DM &dm, ulk_stream_t, const F &f)
{
    for (i64 i = 0; i < m; ++i) {
        i32 r = rl ? rl[i] : (i32)i, v = fgl[r], z, y, x;
        dm.coords(v, z, y, x);
        f(r, v, z, y, x);
    }
}
#else
template <class F> __global__ void __launch_bounds__(256) ulk_kernel(i64 n, F f)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) f(i);
}
extern long long g_ulk_launches;
// ULK_DEBUG_SYNC=1: every launch is followed by a stream synchronize; the first failing kernel
// is named on stderr (debugging aid, off by default)
extern int g_ulk_debug_sync;
void ulk_debug_after_launch(const char *name, ulk_stream_t s);
// optional per-kernel timing (CUDA events on the launching stream), see ulk_profile_* in the C ABI
extern bool g_ulk_prof_on;
void ulk_prof_begin(const char *name, ulk_stream_t s);
void ulk_prof_end(ulk_stream_t s);
template <class F> void ulk_launch(i64 n, ulk_stream_t s, const F &f)
{
    if (n <= 0) return;
    i64 grid = (n + 255) / 256;
    if (g_ulk_prof_on) ulk_prof_begin(typeid(F).name(), s);
    ulk_kernel<F><<<(unsigned)grid, 256, 0, s>>>(n, f);
    if (g_ulk_prof_on) ulk_prof_end(s);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch(typeid(F).name(), s);
}
// warp-cooperative kernels: every thread of the grid calls F::warp(i, i < n), so the body may
// use full-mask warp collectives; lane l of a warp holds element 32*k + l.
template <class F> __global__ void __launch_bounds__(256) ulk_kernel_warp(i64 n, F f)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    f.warp(i, i < n);
}
template <class F> void ulk_launch_warp(i64 n, ulk_stream_t s, const F &f)
{
    if (n <= 0) return;
    i64 grid = (n + 255) / 256;
    if (g_ulk_prof_on) ulk_prof_begin(typeid(F).name(), s);
    ulk_kernel_warp<F><<<(unsigned)grid, 256, 0, s>>>(n, f);
    if (g_ulk_prof_on) ulk_prof_end(s);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch(typeid(F).name(), s);
}
// voxel kernels that need coordinates: 3-D grid (x blocks, y, z), no index division; lane l of a
// warp holds x = 32*k + l of one row.  The _warp form calls F::warp(v, z, y, x, valid) from
// every thread (valid = x < W) so the body may use full-mask warp collectives.
#define ULK_VOX_BX 128
template <class F> __global__ void __launch_bounds__(ULK_VOX_BX) ulk_kernel_vox(i32 H, i32 W, F f)
{
    i32 x = blockIdx.x * ULK_VOX_BX + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
    if (x < W) f(((i64)z * H + y) * W + x, z, y, x);
}
template <class F> __global__ void __launch_bounds__(ULK_VOX_BX) ulk_kernel_vox_warp(i32 H, i32 W, F f)
{
    i32 x = blockIdx.x * ULK_VOX_BX + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
    f.warp(((i64)z * H + y) * W + x, z, y, x, x < W);
}
template <class F> void ulk_launch_vox(i32 D, i32 H, i32 W, ulk_stream_t s, const F &f)
{
    if (D <= 0 || H <= 0 || W <= 0) return;
    dim3 grid((unsigned)((W + ULK_VOX_BX - 1) / ULK_VOX_BX), (unsigned)H, (unsigned)D);
    if (g_ulk_prof_on) ulk_prof_begin(typeid(F).name(), s);
    ulk_kernel_vox<F><<<grid, ULK_VOX_BX, 0, s>>>(H, W, f);
    if (g_ulk_prof_on) ulk_prof_end(s);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch(typeid(F).name(), s);
}
template <class F> void ulk_launch_vox_warp(i32 D, i32 H, i32 W, ulk_stream_t s, const F &f)
{
    if (D <= 0 || H <= 0 || W <= 0) return;
    dim3 grid((unsigned)((W + ULK_VOX_BX - 1) / ULK_VOX_BX), (unsigned)H, (unsigned)D);
    if (g_ulk_prof_on) ulk_prof_begin(typeid(F).name(), s);
    ulk_kernel_vox_warp<F><<<grid, ULK_VOX_BX, 0, s>>>(H, W, f);
    if (g_ulk_prof_on) ulk_prof_end(s);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch(typeid(F).name(), s);
}
// list kernels: one thread per entry of a voxel list (the foreground voxels of a frame, in
// raster order): every lane does useful work, and consecutive entries are mostly consecutive
// voxels of a row, so accesses stay coalesced.  Coordinates come from UlkDims::coords (multiply-
// shift division).  The _warp form calls F::warp(v, z, y, x, valid) from every thread.
template <class F, class DM> __global__ void __launch_bounds__(256) ulk_kernel_list(const i32 *list, i64 n, DM dm, F f)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    i32 v = list[i], z, y, x;
    dm.coords(v, z, y, x);
    f((i64)v, z, y, x);
}
template <class F, class DM> __global__ void __launch_bounds__(256) ulk_kernel_list_warp(const i32 *list, i64 n, DM dm, F f)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    i32 v = 0, z = 0, y = 0, x = 0;
    if (i < n) { v = list[i]; dm.coords(v, z, y, x); }
    f.warp((i64)v, z, y, x, i < n);
}
template <class F, class DM> void ulk_launch_list(const i32 *list, i64 n, const DM &dm, ulk_stream_t s, const F &f)
{
    if (n <= 0) return;
    if (g_ulk_prof_on) ulk_prof_begin(typeid(F).name(), s);
    ulk_kernel_list<F, DM><<<(unsigned)((n + 255) / 256), 256, 0, s>>>(list, n, dm, f);
    if (g_ulk_prof_on) ulk_prof_end(s);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch(typeid(F).name(), s);
}
template <class F, class DM> void ulk_launch_list_warp(const i32 *list, i64 n, const DM &dm, ulk_stream_t s, const F &f)
{
    if (n <= 0) return;
    if (g_ulk_prof_on) ulk_prof_begin(typeid(F).name(), s);
    ulk_kernel_list_warp<F, DM><<<(unsigned)((n + 255) / 256), 256, 0, s>>>(list, n, dm, f);
    if (g_ulk_prof_on) ulk_prof_end(s);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch(typeid(F).name(), s);
}
// pointer-chasing list kernels: a thread takes ULK_MULTI entries (256 apart, so every load of
// the list stays coalesced) and calls F::multi(v[], valid[]) from every thread of the block; the
// body walks its ULK_MULTI chains in lockstep so that many loads are in flight per thread
#define ULK_MULTI 4
template <class F> __global__ void __launch_bounds__(256) ulk_kernel_list_multi(const i32 *list, i64 n, F f)
{
    const i64 base = (i64)blockIdx.x * (256 * ULK_MULTI) + threadIdx.x;
    i32 v[ULK_MULTI];
    bool ok[ULK_MULTI];
#pragma unroll
    for (int u = 0; u < ULK_MULTI; ++u) {
        const i64 i = base + u * 256;
        ok[u] = i < n;
        v[u] = ok[u] ? list[i] : -1;
    }
    f.multi(v, ok);
}
// phased list kernels: a thread takes U entries (256 apart: lane l of a warp still holds list
// entry 32*k + l, so warp collectives see neighbours of the raster-ordered list) and runs the
// functor's phases entry-interleaved -- S1 load1(v,z,y,x,ok), S2 load2(S1,ok), run(S2,ok) -- so
// that the U independent loads of every dependency level are in flight together.  Every thread
// calls every phase (ok = false past the end of the list): phases may use full-mask collectives.
template <int U, class F, class DM> __global__ void __launch_bounds__(256) ulk_kernel_list_ph(const i32 *list, i64 n, DM dm, F f)
{
    const i64 base = (i64)blockIdx.x * (256 * U) + threadIdx.x;
    typename F::S1 a[U];
    typename F::S2 b[U];
    bool ok[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const i64 i = base + u * 256;
        ok[u] = i < n;
        i32 v = 0, z = 0, y = 0, x = 0;
        if (ok[u]) { v = list[i]; dm.coords(v, z, y, x); }
        a[u] = f.load1(v, z, y, x, ok[u]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) b[u] = f.load2(a[u], ok[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) f.run(b[u], ok[u]);
}
template <int U, class F, class DM> void ulk_launch_list_ph(const i32 *list, i64 n, const DM &dm, ulk_stream_t s, const F &f)
{
    if (n <= 0) return;
    if (g_ulk_prof_on) ulk_prof_begin(typeid(F).name(), s);
    ulk_kernel_list_ph<U, F, DM><<<(unsigned)((n + 256 * U - 1) / (256 * U)), 256, 0, s>>>(list, n, dm, f);
    if (g_ulk_prof_on) ulk_prof_end(s);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch(typeid(F).name(), s);
}
// tuning aid: entries per thread from the environment (ULK_U_<tag> = 1, 2 or 4), else the default
template <int UDEF, class F, class DM> void ulk_launch_list_phx(const char *tag, const i32 *list, i64 n, const DM &dm, ulk_stream_t s, const F &f)
{
    char name[64];
    snprintf(name, sizeof(name), "ULK_U_%s", tag);
    const char *e = getenv(name);
    int U = e ? atoi(e) : UDEF;
    if (U == 1) ulk_launch_list_ph<1>(list, n, dm, s, f);
    else if (U == 2) ulk_launch_list_ph<2>(list, n, dm, s, f);
    else ulk_launch_list_ph<4>(list, n, dm, s, f);
}
template <class F> void ulk_launch_list_multi(const i32 *list, i64 n, ulk_stream_t s, const F &f)
{
    if (n <= 0) return;
    if (g_ulk_prof_on) ulk_prof_begin(typeid(F).name(), s);
    ulk_kernel_list_multi<F><<<(unsigned)((n + 256 * ULK_MULTI - 1) / (256 * ULK_MULTI)), 256, 0, s>>>(list, n, f);
    if (g_ulk_prof_on) ulk_prof_end(s);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch(typeid(F).name(), s);
}
// ---- rank-space list kernels.  Every per-voxel array of the voxel stage is indexed by the
// foreground RANK of the voxel (its position in the raster-ordered foreground list fgl), so a
// kernel's own accesses are dense whatever the foreground fraction.  A list entry i is the rank
// r = rl ? rl[i] : i (rl == nullptr: the whole foreground); its voxel is fgl[r].  Lane l of a warp
// holds entry 32k + l: x-neighbours in the frame are neighbours in the warp.
// phased, with coordinates: S1 load1(r, v, z, y, x, ok) -> S2 load2(S1, ok) -> run(S2, ok)
template <int U, class F, class DM> __global__ void __launch_bounds__(256) ulk_kernel_rk(const i32 *rl, const i32 *fgl, i64 m, DM dm, F f)
{
    const i64 base = (i64)blockIdx.x * (256 * U) + threadIdx.x;
    typename F::S1 a[U];
    typename F::S2 b[U];
    bool ok[U];
    if constexpr (ulk_has_load0<F>::value) {
        // the entries that depend on the rank alone are requested together with the voxel index,
        // before the coordinate arithmetic waits for it (no branch in between: rank 0 stands in
        // for the entries past the end of the list)
        typename F::S0 p[U];
        i32 rr[U], vv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const i64 i = base + u * 256;
            ok[u] = i < m;
            rr[u] = ok[u] ? (rl ? rl[i] : (i32)i) : 0;
            vv[u] = fgl[rr[u]];
            p[u] = f.load0(rr[u], ok[u]);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            i32 z, y, x;
            dm.coords(vv[u], z, y, x);
            a[u] = f.load1(p[u], rr[u], vv[u], z, y, x, ok[u]);
        }
    } else {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const i64 i = base + u * 256;
            ok[u] = i < m;
            i32 r = 0, v = 0, z = 0, y = 0, x = 0;
            if (ok[u]) { r = rl ? rl[i] : (i32)i; v = fgl[r]; dm.coords(v, z, y, x); }
            a[u] = f.load1(r, v, z, y, x, ok[u]);
        }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) b[u] = f.load2(a[u], ok[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) f.run(b[u], ok[u]);
}
template <int U, class F, class DM> void ulk_launch_rk(const i32 *rl, const i32 *fgl, i64 m, const DM &dm, ulk_stream_t s, const F &f)
{
    if (m <= 0) return;
    if (g_ulk_prof_on) ulk_prof_begin(typeid(F).name(), s);
    ulk_kernel_rk<U, F, DM><<<(unsigned)((m + 256 * U - 1) / (256 * U)), 256, 0, s>>>(rl, fgl, m, dm, f);
    if (g_ulk_prof_on) ulk_prof_end(s);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch(typeid(F).name(), s);
}
template <int UDEF, class F, class DM> void ulk_launch_rkx(const char *tag, const i32 *rl, const i32 *fgl, i64 m, const DM &dm, ulk_stream_t s, const F &f)
{
    char name[64];
    snprintf(name, sizeof(name), "ULK_U_%s", tag);
    const char *e = getenv(name);
    int U = e ? atoi(e) : UDEF;
    if (U == 1) ulk_launch_rk<1>(rl, fgl, m, dm, s, f);
    else if (U == 2) ulk_launch_rk<2>(rl, fgl, m, dm, s, f);
    else ulk_launch_rk<4>(rl, fgl, m, dm, s, f);
}
// phased, rank only (no voxel index, no coordinates): S1 load1(r, ok) -> S2 load2(S1, ok) -> run(S2, ok)
template <int U, class F> __global__ void __launch_bounds__(256) ulk_kernel_rkn(const i32 *rl, i64 m, F f)
{
    const i64 base = (i64)blockIdx.x * (256 * U) + threadIdx.x;
    typename F::S1 a[U];
    typename F::S2 b[U];
    bool ok[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const i64 i = base + u * 256;
        ok[u] = i < m;
        const i32 r = ok[u] ? (rl ? rl[i] : (i32)i) : 0;
        a[u] = f.load1(r, ok[u]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) b[u] = f.load2(a[u], ok[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) f.run(b[u], ok[u]);
}
template <int U, class F> void ulk_launch_rkn(const i32 *rl, i64 m, ulk_stream_t s, const F &f)
{
    if (m <= 0) return;
    if (g_ulk_prof_on) ulk_prof_begin(typeid(F).name(), s);
    ulk_kernel_rkn<U, F><<<(unsigned)((m + 256 * U - 1) / (256 * U)), 256, 0, s>>>(rl, m, f);
    if (g_ulk_prof_on) ulk_prof_end(s);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch(typeid(F).name(), s);
}
// pointer chasing: ULK_MULTI ranks per thread, F::multi(r[], ok[]) from every thread of the block
template <class F> __global__ void __launch_bounds__(256) ulk_kernel_rk_multi(const i32 *rl, i64 m, F f)
{
    const i64 base = (i64)blockIdx.x * (256 * ULK_MULTI) + threadIdx.x;
    i32 r[ULK_MULTI];
    bool ok[ULK_MULTI];
#pragma unroll
    for (int u = 0; u < ULK_MULTI; ++u) {
        const i64 i = base + u * 256;
        ok[u] = i < m;
        r[u] = ok[u] ? (rl ? rl[i] : (i32)i) : -1;
    }
    f.multi(r, ok);
}
template <class F> void ulk_launch_rk_multi(const i32 *rl, i64 m, ulk_stream_t s, const F &f)
{
    if (m <= 0) return;
    if (g_ulk_prof_on) ulk_prof_begin(typeid(F).name(), s);
    ulk_kernel_rk_multi<F><<<(unsigned)((m + 256 * ULK_MULTI - 1) / (256 * ULK_MULTI)), 256, 0, s>>>(rl, m, f);
    if (g_ulk_prof_on) ulk_prof_end(s);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch(typeid(F).name(), s);
}
// one thread per entry, with coordinates: f(r, v, z, y, x)   (short lists)
template <class F, class DM> __global__ void __launch_bounds__(256) ulk_kernel_rk1(const i32 *rl, const i32 *fgl, i64 m, DM dm, F f)
{
    const i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const i32 r = rl ? rl[i] : (i32)i, v = fgl[r];
    i32 z, y, x;
    dm.coords(v, z, y, x);
    f(r, v, z, y, x);
}
template <class F, class DM> void ulk_launch_rk1(const i32 *rl, const i32 *fgl, i64 m, const DM &dm, ulk_stream_t s, const F &f)
{
    if (m <= 0) return;
    if (g_ulk_prof_on) ulk_prof_begin(typeid(F).name(), s);
    ulk_kernel_rk1<F, DM><<<(unsigned)((m + 255) / 256), 256, 0, s>>>(rl, fgl, m, dm, f);
    if (g_ulk_prof_on) ulk_prof_end(s);
    ++g_ulk_launches;
    if (g_ulk_debug_sync) ulk_debug_after_launch(typeid(F).name(), s);
}
#endif
int ulk_prof_report(char *buf, i64 cap);   // "name\tlaunches\ttotal_ms\n" per kernel; returns bytes needed
void ulk_prof_enable(int on);
void ulk_prof_reset();

// ---------------------------------------------------------------- memory
struct UlkArena;  // growable pool of device buffers, see ulk_rt.cu
void ulk_set_alloc_stream(ulk_stream_t s);
int ulk_dev_alloc(void **p, size_t bytes);
void ulk_dev_free(void *p);
// frame-scoped scratch: carved out of one retained device slab (host-side bookkeeping only, no
// driver call per buffer); everything that uses it is ordered on the call's stream.
// ulk_ws_reserve sizes the slab before a call starts (no-op when it is already large enough).
int ulk_ws_reserve(size_t bytes, ulk_stream_t s);
int ulk_ws_alloc(void **p, size_t bytes);
void ulk_ws_free(void *p);
// bytes: [0] slab capacity, [1] high-water mark of scratch in use, [2] bytes served outside the slab
void ulk_ws_stats(size_t out[3]);
int ulk_memset(void *p, int byte, size_t bytes, ulk_stream_t s);
int ulk_d2h(void *dst, const void *src, size_t bytes, ulk_stream_t s);  // synchronises
int ulk_d2h2(void *dst1, const void *src1, size_t n1, void *dst2, const void *src2, size_t n2, ulk_stream_t s);  // one sync
int ulk_h2d(void *dst, const void *src, size_t bytes, ulk_stream_t s);
int ulk_d2d(void *dst, const void *src, size_t bytes, ulk_stream_t s);
int ulk_sync(ulk_stream_t s);

// ---------------------------------------------------------------- primitives
// exclusive prefix sum of n int32 -> int32 (out may alias in); returns total via *total (host)
int ulk_exclusive_scan_i32(const i32 *in, i32 *out, i64 n, ulk_stream_t s);
int ulk_exclusive_scan_i64(const i64 *in, i64 *out, i64 n, ulk_stream_t s);
// stable LSD radix sort of (key, value) pairs; results land in keys_out / vals_out
int ulk_sort_pairs_u64(const u64 *keys_in, u64 *keys_out, const u32 *vals_in, u32 *vals_out,
                       i64 n, ulk_stream_t s);
int ulk_sort_pairs_u32(const u32 *keys_in, u32 *keys_out, const u32 *vals_in, u32 *vals_out,
                       i64 n, ulk_stream_t s);
int ulk_sort_keys_u32(const u32 *keys_in, u32 *keys_out, i64 n, ulk_stream_t s);
// indices i (ascending) with in[i] >= 0 -> out; *count_dev receives how many
int ulk_select_nonneg_i32(const i32 *in, i64 n, i32 *out, i32 *count_dev, ulk_stream_t s);
int ulk_sort_keys_u64(const u64 *keys_in, u64 *keys_out, i64 n, ulk_stream_t s);
// entries v of a voxel list (order kept) whose basin still has more than one root
int ulk_select_list_split_i32(const i32 *list, i64 n, const i32 *bas, const i32 *nroot, i32 *out, i32 *count_dev, ulk_stream_t s);
// exclusive prefix sum of the population counts of n 32-bit words
int ulk_exclusive_scan_popc_u32(const u32 *words, i32 *out, i64 n, ulk_stream_t s);

// ---------------------------------------------------------------- numerics
// order-preserving map float32 -> uint32 (handles negatives)
ULK_HD ULK_INLINE u32 ulk_f2ord(float f)
{
    u32 b;
    memcpy(&b, &f, 4);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
ULK_HD ULK_INLINE float ulk_ord2f(u32 o)
{
    u32 b = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
    float f;
    memcpy(&f, &b, 4);
    return f;
}
ULK_HD ULK_INLINE u64 ulk_d2ord(double d)
{
    u64 b;
    memcpy(&b, &d, 8);
    return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
