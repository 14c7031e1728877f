// ulk_rt.cu -- runtime layer implementation (see ulk_rt.h).
#include "ulk_rt.h"
#include <cstdarg>
#include <string>
#include <utility>

static thread_local std::string g_err;
long long g_ulk_launches = 0;

void ulk_set_error(const char *fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
}
const char *ulk_error_cstr() { return g_err.c_str(); }

#ifdef ULK_EMU
// ------------------------------------------------------------------ host emulation
void ulk_prof_enable(int) {}
void ulk_prof_reset() {}
int ulk_prof_report(char *buf, i64 cap) { if (buf && cap > 0) buf[0] = 0; return 1; }
void ulk_set_alloc_stream(ulk_stream_t) {}
int ulk_dev_alloc(void **p, size_t bytes) { *p = malloc(bytes ? bytes : 1); return *p ? 0 : -3; }
void ulk_dev_free(void *p) { free(p); }
int ulk_ws_reserve(size_t, ulk_stream_t) { return 0; }
int ulk_ws_alloc(void **p, size_t bytes) { return ulk_dev_alloc(p, bytes); }
void ulk_ws_free(void *p) { free(p); }
void ulk_ws_stats(size_t out[3]) { out[0] = out[1] = out[2] = 0; }
int ulk_memset(void *p, int byte, size_t bytes, ulk_stream_t) { memset(p, byte, bytes); return 0; }
int ulk_d2h(void *dst, const void *src, size_t bytes, ulk_stream_t) { memcpy(dst, src, bytes); return 0; }
int ulk_d2h2(void *d1, const void *s1, size_t n1, void *d2, const void *s2, size_t n2, ulk_stream_t)
{
    memcpy(d1, s1, n1); memcpy(d2, s2, n2);
    return 0;
}
int ulk_h2d(void *dst, const void *src, size_t bytes, ulk_stream_t) { memcpy(dst, src, bytes); return 0; }
int ulk_d2d(void *dst, const void *src, size_t bytes, ulk_stream_t) { memmove(dst, src, bytes); return 0; }
int ulk_sync(ulk_stream_t) { return 0; }

int ulk_exclusive_scan_i32(const i32 *in, i32 *out, i64 n, ulk_stream_t)
{
    i32 acc = 0;
    for (i64 i = 0; i < n; ++i) { i32 v = in[i]; out[i] = acc; acc += v; }
    return 0;
}
int ulk_exclusive_scan_i64(const i64 *in, i64 *out, i64 n, ulk_stream_t)
{
    i64 acc = 0;
    for (i64 i = 0; i < n; ++i) { i64 v = in[i]; out[i] = acc; acc += v; }
    return 0;
}
template <class K>
static int emu_sort(const K *ki, K *ko, const u32 *vi, u32 *vo, i64 n)
{
    std::vector<i64> idx((size_t)n);
    std::iota(idx.begin(), idx.end(), (i64)0);
    std::stable_sort(idx.begin(), idx.end(), [&](i64 a, i64 b) { return ki[a] < ki[b]; });
    std::vector<K> k2((size_t)n);
    std::vector<u32> v2((size_t)n);
    for (i64 i = 0; i < n; ++i) { k2[i] = ki[idx[i]]; v2[i] = vi[idx[i]]; }
    for (i64 i = 0; i < n; ++i) { ko[i] = k2[i]; vo[i] = v2[i]; }
    return 0;
}
int ulk_sort_pairs_u64(const u64 *ki, u64 *ko, const u32 *vi, u32 *vo, i64 n, ulk_stream_t) { return emu_sort(ki, ko, vi, vo, n); }
int ulk_sort_pairs_u32(const u32 *ki, u32 *ko, const u32 *vi, u32 *vo, i64 n, ulk_stream_t) { return emu_sort(ki, ko, vi, vo, n); }
int ulk_select_nonneg_i32(const i32 *in, i64 n, i32 *out, i32 *count, ulk_stream_t)
{
    i32 c = 0;
    for (i64 i = 0; i < n; ++i) if (in[i] >= 0) out[c++] = (i32)i;
    *count = c;
    return 0;
}
int ulk_select_list_split_i32(const i32 *list, i64 n, const i32 *bas, const i32 *nroot, i32 *out, i32 *count, ulk_stream_t)
{
    i32 c = 0;
    for (i64 i = 0; i < n; ++i) { i32 r = list ? list[i] : (i32)i; if (bas[r] >= 0 && nroot[bas[r]] > 1) out[c++] = r; }
    *count = c;
    return 0;
}
int ulk_exclusive_scan_popc_u32(const u32 *words, i32 *out, i64 n, ulk_stream_t)
{
    i32 acc = 0;
    for (i64 i = 0; i < n; ++i) { out[i] = acc; acc += __builtin_popcount(words[i]); }
    return 0;
}
int ulk_sort_keys_u32(const u32 *ki, u32 *ko, i64 n, ulk_stream_t)
{
    std::vector<u32> k(ki, ki + n);
    std::sort(k.begin(), k.end());
    for (i64 i = 0; i < n; ++i) ko[i] = k[i];
    return 0;
}
int ulk_sort_keys_u64(const u64 *ki, u64 *ko, i64 n, ulk_stream_t)
{
    std::vector<u64> k(ki, ki + n);
    std::sort(k.begin(), k.end());
    for (i64 i = 0; i < n; ++i) ko[i] = k[i];
    return 0;
}

#else
// ------------------------------------------------------------------ CUDA
// per-kernel timing: one event pair per launch, resolved when the report is read
#include <cxxabi.h>
#include <thrust/iterator/counting_iterator.h>
#include <thrust/iterator/transform_iterator.h>
#include <map>
#include <vector>
bool g_ulk_prof_on = false;
int g_ulk_debug_sync = getenv("ULK_DEBUG_SYNC") ? atoi(getenv("ULK_DEBUG_SYNC")) : 0;
void ulk_debug_after_launch(const char *name, ulk_stream_t s)
{
    cudaError_t e = cudaStreamSynchronize(s);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) { fprintf(stderr, "ULK_DEBUG_SYNC: %s failed: %s\n", name, cudaGetErrorString(e)); fflush(stderr); }
}
struct ProfRec { const char *name; cudaEvent_t a, b; };
static std::vector<ProfRec> g_prof_recs;
static std::vector<cudaEvent_t> g_prof_free;
static std::map<std::string, std::pair<long long, double>> g_prof_acc;
static cudaEvent_t prof_event()
{
    if (!g_prof_free.empty()) { cudaEvent_t e = g_prof_free.back(); g_prof_free.pop_back(); return e; }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}
static void prof_drain()
{
    for (auto &r : g_prof_recs) {
        cudaEventSynchronize(r.b);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, r.a, r.b);
        auto &acc = g_prof_acc[r.name];
        acc.first += 1;
        acc.second += ms;
        g_prof_free.push_back(r.a);
        g_prof_free.push_back(r.b);
    }
    g_prof_recs.clear();
}
void ulk_prof_begin(const char *name, ulk_stream_t s)
{
    ProfRec r{name, prof_event(), prof_event()};
    cudaEventRecord(r.a, s);
    g_prof_recs.push_back(r);
}
void ulk_prof_end(ulk_stream_t s)
{
    cudaEventRecord(g_prof_recs.back().b, s);
    if (g_prof_recs.size() >= 4096) prof_drain();
}
void ulk_prof_enable(int on)
{
    g_ulk_prof_on = on != 0;
    // the event pool is created here, not launch by launch inside a timed region
    if (g_ulk_prof_on)
        while (g_prof_free.size() < 2 * 4096 + 64) {
            cudaEvent_t e;
            if (cudaEventCreate(&e) != cudaSuccess) break;
            g_prof_free.push_back(e);
        }
}
void ulk_prof_reset() { prof_drain(); g_prof_acc.clear(); }
int ulk_prof_report(char *buf, i64 cap)
{
    prof_drain();
    std::string out;
    for (auto &kv : g_prof_acc) {
        int status = 0;
        char *dem = abi::__cxa_demangle(kv.first.c_str(), nullptr, nullptr, &status);
        char line[512];
        snprintf(line, sizeof(line), "%s\t%lld\t%.6f\n", status == 0 && dem ? dem : kv.first.c_str(), kv.second.first, kv.second.second);
        free(dem);
        out += line;
    }
    if (buf && cap > 0) { size_t n = out.size() < (size_t)cap - 1 ? out.size() : (size_t)cap - 1; memcpy(buf, out.data(), n); buf[n] = 0; }
    return (int)out.size() + 1;
}
struct ProfScope {
    ulk_stream_t s; bool on;
    ProfScope(const char *name, ulk_stream_t st) : s(st), on(g_ulk_prof_on) { if (on) ulk_prof_begin(name, s); }
    ~ProfScope() { if (on) ulk_prof_end(s); }
};
// stream-ordered allocator: frame-scoped buffers come from the default memory pool, whose
// release threshold is raised so that freed blocks stay cached between frames
static cudaStream_t g_alloc_stream = 0;
static bool g_pool_ready = false;
void ulk_set_alloc_stream(ulk_stream_t s)
{
    // scratch blocks are recycled in host order: work left on another stream must be done first
    if (s != g_alloc_stream) cudaStreamSynchronize(g_alloc_stream);
    g_alloc_stream = s;
    if (!g_pool_ready) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
            unsigned long long thr = ~0ull;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
        }
        g_pool_ready = true;
    }
}
int ulk_dev_alloc(void **p, size_t bytes)
{
    ULK_CUDA(cudaMallocAsync(p, bytes ? bytes : 1, g_alloc_stream));
    return 0;
}
void ulk_dev_free(void *p) { if (p) cudaFreeAsync(p, g_alloc_stream); }

// ---- workspace slab: best-fit free list over one retained allocation
static char *g_ws_base = nullptr;
static size_t g_ws_cap = 0, g_ws_used = 0, g_ws_peak = 0, g_ws_outside = 0, g_ws_outside_now = 0;
static std::map<size_t, size_t> g_ws_freelist;             // offset -> size
static std::map<void *, size_t> g_ws_live;                 // slab pointer -> size
static std::map<void *, size_t> g_ws_live_outside;         // cudaMallocAsync pointer -> size
static const size_t WS_ALIGN = 512;
int ulk_ws_reserve(size_t bytes, ulk_stream_t s)
{
    (void)s;
    // grow when the request, or what the previous call had to take from outside the slab, exceeds it
    size_t want = bytes;
    if (g_ws_peak > want) want = g_ws_peak + (g_ws_peak >> 4);
    if (want <= g_ws_cap || !g_ws_live.empty()) return 0;
    cudaDeviceSynchronize();
    if (g_ws_base) { cudaFree(g_ws_base); g_ws_base = nullptr; g_ws_cap = 0; }
    g_ws_freelist.clear();
    {   // hand cached blocks of the stream-ordered pool back before sizing the slab
        int dev = 0;
        cudaMemPool_t pool;
        cudaGetDevice(&dev);
        if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
    }
    size_t free_b = 0, total_b = 0;
    cudaMemGetInfo(&free_b, &total_b);
    size_t lim = free_b > (6ull << 30) ? free_b - (6ull << 30) : free_b / 2;   // leave room for results
    if (want > lim) want = lim;
    want = (want + WS_ALIGN - 1) / WS_ALIGN * WS_ALIGN;
    while (want >= (64ull << 20)) {
        void *p = nullptr;
        if (cudaMalloc(&p, want) == cudaSuccess) { g_ws_base = (char *)p; g_ws_cap = want; break; }
        cudaGetLastError();
        want = want / 4 * 3 / WS_ALIGN * WS_ALIGN;
    }
    if (g_ws_cap) g_ws_freelist[0] = g_ws_cap;
    // prime the stream-ordered pool the results come from (node tables, masks, links of the two
    // or three frames alive at a time): one block allocated and handed back stays cached (the
    // release threshold is unlimited), so steady-state calls never grow the pool
    {
        size_t prime = bytes / 16;        // ~2 B per frame voxel: the masks + tables of three frames at 20-30 % foreground
        if (prime < (256ull << 20)) prime = 256ull << 20;
        void *p = nullptr;
        if (cudaMallocAsync(&p, prime, g_alloc_stream) == cudaSuccess) cudaFreeAsync(p, g_alloc_stream);
        else cudaGetLastError();
    }
    return 0;
}
int ulk_ws_alloc(void **p, size_t bytes)
{
    size_t need = ((bytes ? bytes : 1) + WS_ALIGN - 1) / WS_ALIGN * WS_ALIGN;
    auto best = g_ws_freelist.end();
    for (auto it = g_ws_freelist.begin(); it != g_ws_freelist.end(); ++it)
        if (it->second >= need && (best == g_ws_freelist.end() || it->second < best->second)) best = it;
    g_ws_used += need;
    if (g_ws_used > g_ws_peak) g_ws_peak = g_ws_used;
    if (best == g_ws_freelist.end()) {
        g_ws_used -= need;
        int rc = ulk_dev_alloc(p, need);
        if (rc) return rc;
        g_ws_used += need;
        g_ws_live_outside[*p] = need;
        g_ws_outside_now += need;
        if (g_ws_outside_now > g_ws_outside) g_ws_outside = g_ws_outside_now;
        return 0;
    }
    size_t off = best->first, sz = best->second;
    g_ws_freelist.erase(best);
    if (sz > need) g_ws_freelist[off + need] = sz - need;
    *p = g_ws_base + off;
    g_ws_live[*p] = need;
    return 0;
}
void ulk_ws_free(void *p)
{
    if (!p) return;
    auto io = g_ws_live_outside.find(p);
    if (io != g_ws_live_outside.end()) {
        g_ws_used -= io->second; g_ws_outside_now -= io->second;
        g_ws_live_outside.erase(io);
        ulk_dev_free(p);
        return;
    }
    auto il = g_ws_live.find(p);
    if (il == g_ws_live.end()) return;
    size_t off = (size_t)((char *)p - g_ws_base), sz = il->second;
    g_ws_used -= sz;
    g_ws_live.erase(il);
    auto nx = g_ws_freelist.lower_bound(off);
    if (nx != g_ws_freelist.end() && off + sz == nx->first) { sz += nx->second; nx = g_ws_freelist.erase(nx); }
    if (nx != g_ws_freelist.begin()) {
        auto pv = std::prev(nx);
        if (pv->first + pv->second == off) { pv->second += sz; return; }
    }
    g_ws_freelist[off] = sz;
}
void ulk_ws_stats(size_t out[3]) { out[0] = g_ws_cap; out[1] = g_ws_peak; out[2] = g_ws_outside; }
int ulk_memset(void *p, int byte, size_t bytes, ulk_stream_t s)
{
    ProfScope ps("cudaMemsetAsync", s);
    ULK_CUDA(cudaMemsetAsync(p, byte, bytes, s));
    return 0;
}
// small counter read-backs go through one pinned staging page (a pageable destination makes
// the driver stage and pin on every call)
static void *g_pinned_page = nullptr;
static const size_t kPinnedPage = 4096;
int ulk_d2h(void *dst, const void *src, size_t bytes, ulk_stream_t s)
{
    if (bytes <= kPinnedPage) {
        if (!g_pinned_page && cudaHostAlloc(&g_pinned_page, kPinnedPage, cudaHostAllocDefault) != cudaSuccess) {
            g_pinned_page = nullptr;
            cudaGetLastError();
        }
        if (g_pinned_page) {
            ULK_CUDA(cudaMemcpyAsync(g_pinned_page, src, bytes, cudaMemcpyDeviceToHost, s));
            ULK_CUDA(cudaStreamSynchronize(s));
            memcpy(dst, g_pinned_page, bytes);
            return 0;
        }
    }
    ULK_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s));
    ULK_CUDA(cudaStreamSynchronize(s));
    return 0;
}
// two small read-backs behind ONE stream synchronisation (every host round trip costs 10-30 us)
int ulk_d2h2(void *dst1, const void *src1, size_t n1, void *dst2, const void *src2, size_t n2, ulk_stream_t s)
{
    const size_t o2 = (n1 + 15) / 16 * 16;
    if (o2 + n2 <= kPinnedPage) {
        if (!g_pinned_page && cudaHostAlloc(&g_pinned_page, kPinnedPage, cudaHostAllocDefault) != cudaSuccess) {
            g_pinned_page = nullptr;
            cudaGetLastError();
        }
        if (g_pinned_page) {
            ULK_CUDA(cudaMemcpyAsync(g_pinned_page, src1, n1, cudaMemcpyDeviceToHost, s));
            ULK_CUDA(cudaMemcpyAsync((char *)g_pinned_page + o2, src2, n2, cudaMemcpyDeviceToHost, s));
            ULK_CUDA(cudaStreamSynchronize(s));
            memcpy(dst1, g_pinned_page, n1);
            memcpy(dst2, (char *)g_pinned_page + o2, n2);
            return 0;
        }
    }
    int rc = ulk_d2h(dst1, src1, n1, s);
    return rc ? rc : ulk_d2h(dst2, src2, n2, s);
}
int ulk_h2d(void *dst, const void *src, size_t bytes, ulk_stream_t s) { ULK_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, s)); return 0; }
int ulk_d2d(void *dst, const void *src, size_t bytes, ulk_stream_t s)
{
    ProfScope ps("cudaMemcpyAsync(d2d)", s);
    ULK_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, s));
    return 0;
}
int ulk_sync(ulk_stream_t s) { ULK_CUDA(cudaStreamSynchronize(s)); return 0; }

// CUB temp storage: one growable scratch buffer per process (calls are stream-ordered
// and the library is used from one host thread per device).
static void *g_cub_tmp = nullptr;
static size_t g_cub_cap = 0;
static int cub_tmp(size_t need)
{
    if (need <= g_cub_cap) return 0;
    if (g_cub_tmp) { cudaDeviceSynchronize(); cudaFree(g_cub_tmp); }
    g_cub_cap = need + (need >> 2) + 256;
    ULK_CUDA(cudaMalloc(&g_cub_tmp, g_cub_cap));
    return 0;
}
int ulk_exclusive_scan_i32(const i32 *in, i32 *out, i64 n, ulk_stream_t s)
{
    if (n <= 0) return 0;
    size_t need = 0;
    ULK_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, need, in, out, (int)n, s));
    if (cub_tmp(need)) return -2;
    ProfScope ps("cub::DeviceScan::ExclusiveSum", s);
    ULK_CUDA(cub::DeviceScan::ExclusiveSum(g_cub_tmp, need, in, out, (int)n, s));
    ++g_ulk_launches;
    return 0;
}
int ulk_exclusive_scan_i64(const i64 *in, i64 *out, i64 n, ulk_stream_t s)
{
    if (n <= 0) return 0;
    size_t need = 0;
    ULK_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, need, in, out, (int)n, s));
    if (cub_tmp(need)) return -2;
    ProfScope ps("cub::DeviceScan::ExclusiveSum", s);
    ULK_CUDA(cub::DeviceScan::ExclusiveSum(g_cub_tmp, need, in, out, (int)n, s));
    ++g_ulk_launches;
    return 0;
}
int ulk_sort_pairs_u64(const u64 *ki, u64 *ko, const u32 *vi, u32 *vo, i64 n, ulk_stream_t s)
{
    if (n <= 0) return 0;
    size_t need = 0;
    ULK_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, need, ki, ko, vi, vo, (int)n, 0, 64, s));
    if (cub_tmp(need)) return -2;
    ProfScope ps("cub::DeviceRadixSort::SortPairs", s);
    ULK_CUDA(cub::DeviceRadixSort::SortPairs(g_cub_tmp, need, ki, ko, vi, vo, (int)n, 0, 64, s));
    ++g_ulk_launches;
    return 0;
}
int ulk_sort_pairs_u32(const u32 *ki, u32 *ko, const u32 *vi, u32 *vo, i64 n, ulk_stream_t s)
{
    if (n <= 0) return 0;
    size_t need = 0;
    ULK_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, need, ki, ko, vi, vo, (int)n, 0, 32, s));
    if (cub_tmp(need)) return -2;
    ProfScope ps("cub::DeviceRadixSort::SortPairs", s);
    ULK_CUDA(cub::DeviceRadixSort::SortPairs(g_cub_tmp, need, ki, ko, vi, vo, (int)n, 0, 32, s));
    ++g_ulk_launches;
    return 0;
}
int ulk_sort_keys_u32(const u32 *ki, u32 *ko, i64 n, ulk_stream_t s)
{
    if (n <= 0) return 0;
    size_t need = 0;
    ULK_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, need, ki, ko, (int)n, 0, 32, s));
    if (cub_tmp(need)) return -2;
    ProfScope ps("cub::DeviceRadixSort::SortKeys", s);
    ULK_CUDA(cub::DeviceRadixSort::SortKeys(g_cub_tmp, need, ki, ko, (int)n, 0, 32, s));
    ++g_ulk_launches;
    return 0;
}
int ulk_sort_keys_u64(const u64 *ki, u64 *ko, i64 n, ulk_stream_t s)
{
    if (n <= 0) return 0;
    size_t need = 0;
    ULK_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, need, ki, ko, (int)n, 0, 64, s));
    if (cub_tmp(need)) return -2;
    ProfScope ps("cub::DeviceRadixSort::SortKeys", s);
    ULK_CUDA(cub::DeviceRadixSort::SortKeys(g_cub_tmp, need, ki, ko, (int)n, 0, 64, s));
    ++g_ulk_launches;
    return 0;
}
struct UlkNonNegAt {
    const i32 *p;
    __device__ bool operator()(const i32 &i) const { return p[i] >= 0; }
};
int ulk_select_nonneg_i32(const i32 *in, i64 n, i32 *out, i32 *count_dev, ulk_stream_t s)
{
    if (n <= 0) return 0;
    thrust::counting_iterator<i32> it(0);
    UlkNonNegAt pred{in};
    size_t need = 0;
    ULK_CUDA(cub::DeviceSelect::If(nullptr, need, it, out, count_dev, (int)n, pred, s));
    if (cub_tmp(need)) return -2;
    ProfScope ps("cub::DeviceSelect::If", s);
    ULK_CUDA(cub::DeviceSelect::If(g_cub_tmp, need, it, out, count_dev, (int)n, pred, s));
    ++g_ulk_launches;
    return 0;
}
struct UlkPopc {
    __host__ __device__ i32 operator()(const u32 &w) const
    {
#ifdef __CUDA_ARCH__
        return __popc(w);
#else
        return __builtin_popcount(w);
#endif
    }
};
int ulk_exclusive_scan_popc_u32(const u32 *words, i32 *out, i64 n, ulk_stream_t s)
{
    if (n <= 0) return 0;
    auto it = thrust::make_transform_iterator(words, UlkPopc());
    size_t need = 0;
    ULK_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, need, it, out, (int)n, s));
    if (cub_tmp(need)) return -2;
    ProfScope ps("cub::DeviceScan::ExclusiveSum", s);
    ULK_CUDA(cub::DeviceScan::ExclusiveSum(g_cub_tmp, need, it, out, (int)n, s));
    ++g_ulk_launches;
    return 0;
}
struct UlkSplitAt {
    const i32 *bas, *nroot;
    __device__ bool operator()(const i32 &v) const { const i32 b = bas[v]; return b >= 0 && nroot[b] > 1; }
};
int ulk_select_list_split_i32(const i32 *list, i64 n, const i32 *bas, const i32 *nroot, i32 *out, i32 *count_dev, ulk_stream_t s)
{
    if (n <= 0) return ulk_memset(count_dev, 0, sizeof(i32), s);
    UlkSplitAt pred{bas, nroot};
    size_t need = 0;
    if (!list) {      // the identity list: every foreground rank
        thrust::counting_iterator<i32> it(0);
        ULK_CUDA(cub::DeviceSelect::If(nullptr, need, it, out, count_dev, (int)n, pred, s));
        if (cub_tmp(need)) return -2;
        ProfScope ps("cub::DeviceSelect::If", s);
        ULK_CUDA(cub::DeviceSelect::If(g_cub_tmp, need, it, out, count_dev, (int)n, pred, s));
        ++g_ulk_launches;
        return 0;
    }
    ULK_CUDA(cub::DeviceSelect::If(nullptr, need, list, out, count_dev, (int)n, pred, s));
    if (cub_tmp(need)) return -2;
    ProfScope ps("cub::DeviceSelect::If", s);
    ULK_CUDA(cub::DeviceSelect::If(g_cub_tmp, need, list, out, count_dev, (int)n, pred, s));
    ++g_ulk_launches;
    return 0;
}
#endif
