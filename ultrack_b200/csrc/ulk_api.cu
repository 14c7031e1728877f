// ulk_api.cu -- the extern "C" boundary (include/ultrack_b200.h).
#include "ulk_seg.h"
#include "ulk_common.h"

extern long long g_ulk_launches;

// scratch slab reserved per frame voxel: the rank-space voxel stage peaks at ~56 B per FOREGROUND
// voxel (12.5 GB at C5, 21 % foreground = 12 B per frame voxel); denser frames spill into the
// stream-ordered pool (ulk_ws_alloc falls back to it), they do not fail
#define ULK_WS_BYTES_PER_VOXEL 32
#define API_CHK(x) do { int rc__ = (x); if (rc__) return rc__; } while (0)

extern "C" {

const char *ulk_last_error(void) { return ulk_error_cstr(); }

int ulk_device_count(void)
{
#ifdef ULK_EMU
    return 0;
#else
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
#endif
}

int64_t ulk_launch_count(void) { return (int64_t)g_ulk_launches; }
void ulk_workspace_stats(int64_t out[3])
{
    size_t v[3];
    ulk_ws_stats(v);
    for (int i = 0; i < 3; ++i) out[i] = (int64_t)v[i];
}

void ulk_profile_enable(int on) { ulk_prof_enable(on); }
void ulk_profile_reset(void) { ulk_prof_reset(); }
int ulk_profile_report(char *buf, int64_t cap) { return ulk_prof_report(buf, cap); }

int ulk_set_device(int device)
{
#ifndef ULK_EMU
    ULK_CUDA(cudaSetDevice(device));
#else
    (void)device;
#endif
    return 0;
}

int ulk_segment_frame(const void *foreground, int fg_dtype, const void *contours, int ct_dtype,
                      const int64_t shape[3], const ulk_seg_params *params, int keep_labels,
                      void *stream, ulk_frame_nodes **out)
{
    if (!foreground || !contours || !shape || !params || !out) ULK_FAIL(-1, "null argument");
    ulk_stream_t st = (ulk_stream_t)stream;
    ulk_set_alloc_stream(st);
    *out = nullptr;
    ulk_ws_reserve((size_t)(shape[0] * shape[1] * shape[2]) * ULK_WS_BYTES_PER_VOXEL + (64u << 20), st);
    int rc = ulk_segment_frame_impl(st, foreground, fg_dtype, contours, ct_dtype, shape, params, keep_labels, out);
    if (rc) { ulk_frame_nodes_free_impl(*out); *out = nullptr; ulk_sync(st); }
    return rc;
}

int ulk_segment_frame_host(const void *foreground, int fg_dtype, const void *contours, int ct_dtype,
                           const int64_t shape[3], const ulk_seg_params *params, int keep_labels,
                           void *stream, ulk_frame_nodes **out)
{
    if (!foreground || !contours || !shape || !params || !out) ULK_FAIL(-1, "null argument");
    ulk_stream_t st = (ulk_stream_t)stream;
    ulk_set_alloc_stream(st);
    i64 V = shape[0] * shape[1] * shape[2];
    ulk_ws_reserve((size_t)V * (ULK_WS_BYTES_PER_VOXEL + ulk_dtype_size(fg_dtype) + ulk_dtype_size(ct_dtype)) + (64u << 20), st);
    UlkArena A;
    u8 *dfg = A.get<u8>(V * ulk_dtype_size(fg_dtype));
    u8 *dct = A.get<u8>(V * ulk_dtype_size(ct_dtype));
    if (A.failed) ULK_FAIL(-3, "device allocation failed");
    API_CHK(ulk_h2d(dfg, foreground, (size_t)V * ulk_dtype_size(fg_dtype), st));
    API_CHK(ulk_h2d(dct, contours, (size_t)V * ulk_dtype_size(ct_dtype), st));
    return ulk_segment_frame(dfg, fg_dtype, dct, ct_dtype, shape, params, keep_labels, stream, out);
}

// ---- pipelined ingest: the copy of frame t+1 runs on the library's copy stream while frame t
// is being segmented on the caller's stream.  Staging buffers are cached (two frames deep).
struct ulk_frame_upload {
    void *dfg, *dct;
    size_t bfg, bct;
    int fg_dtype, ct_dtype;
    int64_t shape[3];
#ifndef ULK_EMU
    cudaEvent_t done;
#endif
};
#ifndef ULK_EMU
static cudaStream_t g_copy_stream = nullptr;
struct StageBuf { void *p; size_t cap; bool busy; };
static StageBuf g_stage[4] = {{nullptr, 0, false}, {nullptr, 0, false}, {nullptr, 0, false}, {nullptr, 0, false}};
static void *stage_get(size_t bytes)
{
    for (auto &b : g_stage) if (!b.busy && b.p && b.cap >= bytes) { b.busy = true; return b.p; }
    for (auto &b : g_stage) if (!b.busy) {
        if (b.p) { cudaFree(b.p); b.p = nullptr; b.cap = 0; }
        if (cudaMalloc(&b.p, bytes) != cudaSuccess) { cudaGetLastError(); b.p = nullptr; return nullptr; }
        b.cap = bytes; b.busy = true;
        return b.p;
    }
    return nullptr;
}
static void stage_put(void *p) { for (auto &b : g_stage) if (b.p == p) b.busy = false; }
#endif

int ulk_upload_begin(const void *foreground, int fg_dtype, const void *contours, int ct_dtype,
                     const int64_t shape[3], ulk_frame_upload **out)
{
    if (!foreground || !contours || !shape || !out) ULK_FAIL(-1, "null argument");
    *out = nullptr;
    i64 V = shape[0] * shape[1] * shape[2];
    if (V <= 0) ULK_FAIL(-1, "empty frame");
    ulk_frame_upload *u = new ulk_frame_upload();
    memset(u, 0, sizeof(*u));
    u->fg_dtype = fg_dtype; u->ct_dtype = ct_dtype;
    for (int d = 0; d < 3; ++d) u->shape[d] = shape[d];
    u->bfg = (size_t)V * ulk_dtype_size(fg_dtype);
    u->bct = (size_t)V * ulk_dtype_size(ct_dtype);
#ifdef ULK_EMU
    u->dfg = malloc(u->bfg); u->dct = malloc(u->bct);
    memcpy(u->dfg, foreground, u->bfg); memcpy(u->dct, contours, u->bct);
#else
    if (!g_copy_stream) ULK_CUDA(cudaStreamCreateWithFlags(&g_copy_stream, cudaStreamNonBlocking));
    u->dfg = stage_get(u->bfg);
    u->dct = u->dfg ? stage_get(u->bct) : nullptr;
    if (!u->dfg || !u->dct) {
        if (u->dfg) stage_put(u->dfg);
        delete u;
        ULK_FAIL(-3, "device allocation failed (upload staging)");
    }
    ULK_CUDA(cudaEventCreateWithFlags(&u->done, cudaEventDisableTiming));
    ULK_CUDA(cudaMemcpyAsync(u->dfg, foreground, u->bfg, cudaMemcpyHostToDevice, g_copy_stream));
    ULK_CUDA(cudaMemcpyAsync(u->dct, contours, u->bct, cudaMemcpyHostToDevice, g_copy_stream));
    ULK_CUDA(cudaEventRecord(u->done, g_copy_stream));
#endif
    *out = u;
    return 0;
}

void ulk_upload_free(ulk_frame_upload *u)
{
    if (!u) return;
#ifdef ULK_EMU
    free(u->dfg); free(u->dct);
#else
    cudaEventSynchronize(u->done);       // the staging buffers go back only when the copy is over
    cudaEventDestroy(u->done);
    stage_put(u->dfg); stage_put(u->dct);
#endif
    delete u;
}

int ulk_segment_frame_uploaded(ulk_frame_upload *u, const ulk_seg_params *params, int keep_labels,
                               void *stream, ulk_frame_nodes **out)
{
    if (!u || !params || !out) ULK_FAIL(-1, "null argument");
#ifndef ULK_EMU
    ULK_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, u->done, 0));
#endif
    // ulk_segment_frame returns after the stream has drained: the staging can be reused then
    int rc = ulk_segment_frame(u->dfg, u->fg_dtype, u->dct, u->ct_dtype, u->shape, params, keep_labels, stream, out);
    ulk_upload_free(u);
    return rc;
}

void ulk_frame_nodes_free(ulk_frame_nodes *nodes) { ulk_frame_nodes_free_impl(nodes); }

// copies between a node table and caller buffers; `dev` selects device-resident caller
// buffers (d2d) over host buffers
static int copy_out(void *dst, const void *src, size_t bytes, ulk_stream_t st, int dev)
{
    return dev ? ulk_d2d(dst, src, bytes, st) : ulk_d2h(dst, src, bytes, st);
}
static int copy_in(void *dst, const void *src, size_t bytes, ulk_stream_t st, int dev)
{
    return dev ? ulk_d2d(dst, src, bytes, st) : ulk_h2d(dst, src, bytes, st);
}

static int frame_nodes_export(const ulk_frame_nodes *f, int32_t *bbox, int32_t *centroid, int32_t *area,
                              double *frontier, double *height, int32_t *hier_id, int32_t *parent,
                              int32_t *hidx, int64_t *mask_off, uint32_t *mask, int32_t *overlaps,
                              int32_t *labels, void *stream, int dev)
{
    if (!f) ULK_FAIL(-1, "null argument");
    ulk_stream_t st = (ulk_stream_t)stream;
    i64 n = f->n_nodes;
    if (n > 0) {
        if (bbox) API_CHK(copy_out(bbox, f->bbox, (size_t)n * 6 * 4, st, dev));
        if (centroid) API_CHK(copy_out(centroid, f->centroid, (size_t)n * 3 * 4, st, dev));
        if (area) API_CHK(copy_out(area, f->area, (size_t)n * 4, st, dev));
        if (frontier && f->frontier) API_CHK(copy_out(frontier, f->frontier, (size_t)n * 8, st, dev));
        if (height && f->height) API_CHK(copy_out(height, f->height, (size_t)n * 8, st, dev));
        if (hier_id && f->hier_id) API_CHK(copy_out(hier_id, f->hier_id, (size_t)n * 4, st, dev));
        if (parent && f->parent) API_CHK(copy_out(parent, f->parent, (size_t)n * 4, st, dev));
        if (hidx && f->hidx) API_CHK(copy_out(hidx, f->hidx, (size_t)n * 4, st, dev));
        if (mask_off) API_CHK(copy_out(mask_off, f->mask_off, (size_t)(n + 1) * 8, st, dev));
        if (mask && f->mask_words > 0) API_CHK(copy_out(mask, f->mask, (size_t)f->mask_words * 4, st, dev));
        if (overlaps && f->n_overlaps > 0) API_CHK(copy_out(overlaps, f->overlaps, (size_t)f->n_overlaps * 2 * 4, st, dev));
    }
    if (labels && f->labels)
        API_CHK(copy_out(labels, f->labels, (size_t)f->shape[0] * f->shape[1] * f->shape[2] * 4, st, dev));
    return ulk_sync(st);
}

int ulk_frame_nodes_to_host(const ulk_frame_nodes *f, int32_t *bbox, int32_t *centroid, int32_t *area,
                            double *frontier, double *height, int32_t *hier_id, int32_t *parent,
                            int32_t *hidx, int64_t *mask_off, uint32_t *mask, int32_t *overlaps,
                            int32_t *labels, void *stream)
{
    return frame_nodes_export(f, bbox, centroid, area, frontier, height, hier_id, parent, hidx, mask_off, mask,
                              overlaps, labels, stream, 0);
}

int ulk_frame_nodes_to_device(const ulk_frame_nodes *f, int32_t *bbox, int32_t *centroid, int32_t *area,
                              int64_t *mask_off, uint32_t *mask, void *stream)
{
    return frame_nodes_export(f, bbox, centroid, area, nullptr, nullptr, nullptr, nullptr, nullptr, mask_off, mask,
                              nullptr, nullptr, stream, 1);
}

static int frame_nodes_import(int32_t n_nodes, const int32_t shape[3], const int32_t *bbox,
                              const int32_t *centroid, const int32_t *area, const int64_t *mask_off,
                              int64_t words, const uint32_t *mask, void *stream, int dev, ulk_frame_nodes **out)
{
    if (!shape || !out || n_nodes < 0 || words < 0) ULK_FAIL(-1, "bad argument");
    if (n_nodes > 0 && (!bbox || !centroid || !area || !mask_off || (words > 0 && !mask))) ULK_FAIL(-1, "null argument");
    ulk_stream_t st = (ulk_stream_t)stream;
    ulk_set_alloc_stream(st);
    ulk_frame_nodes *F = new ulk_frame_nodes();
    memset(F, 0, sizeof(*F));
    *out = F;
    F->n_nodes = n_nodes;
    for (int d = 0; d < 3; ++d) F->shape[d] = shape[d];
    if (n_nodes == 0) return 0;
    i64 n = n_nodes;
    F->mask_words = words;
    void *p;
    int rc = 0;
    rc |= ulk_dev_alloc(&p, (size_t)n * 6 * 4); F->bbox = (i32 *)p;
    rc |= ulk_dev_alloc(&p, (size_t)n * 3 * 4); F->centroid = (i32 *)p;
    rc |= ulk_dev_alloc(&p, (size_t)n * 4); F->area = (i32 *)p;
    rc |= ulk_dev_alloc(&p, (size_t)(n + 1) * 8); F->mask_off = (i64 *)p;
    rc |= ulk_dev_alloc(&p, (size_t)(words > 0 ? words : 1) * 4); F->mask = (u32 *)p;
    if (rc) { ulk_frame_nodes_free_impl(F); *out = nullptr; ULK_FAIL(-3, "device allocation failed"); }
    API_CHK(copy_in(F->bbox, bbox, (size_t)n * 6 * 4, st, dev));
    API_CHK(copy_in(F->centroid, centroid, (size_t)n * 3 * 4, st, dev));
    API_CHK(copy_in(F->area, area, (size_t)n * 4, st, dev));
    API_CHK(copy_in(F->mask_off, mask_off, (size_t)(n + 1) * 8, st, dev));
    if (words > 0) API_CHK(copy_in(F->mask, mask, (size_t)words * 4, st, dev));
    API_CHK(ulk_sync(st));
    return 0;
}

int ulk_frame_nodes_from_host(int32_t n_nodes, const int32_t shape[3], const int32_t *bbox,
                              const int32_t *centroid, const int32_t *area, const int64_t *mask_off,
                              const uint32_t *mask, void *stream, ulk_frame_nodes **out)
{
    if (n_nodes > 0 && !mask_off) ULK_FAIL(-1, "null argument");
    return frame_nodes_import(n_nodes, shape, bbox, centroid, area, mask_off, n_nodes > 0 ? mask_off[n_nodes] : 0,
                              mask, stream, 0, out);
}

int ulk_frame_nodes_from_device(int32_t n_nodes, const int32_t shape[3], const int32_t *bbox,
                                const int32_t *centroid, const int32_t *area, const int64_t *mask_off,
                                int64_t mask_words, const uint32_t *mask, void *stream, ulk_frame_nodes **out)
{
    return frame_nodes_import(n_nodes, shape, bbox, centroid, area, mask_off, mask_words, mask, stream, 1, out);
}

int ulk_link_frames(const ulk_frame_nodes *src, const ulk_frame_nodes *dst, const double *dst_shift_host,
                    const ulk_link_params *params, void *stream, ulk_links **out)
{
    if (!src || !dst || !params || !out) ULK_FAIL(-1, "null argument");
    ulk_stream_t st = (ulk_stream_t)stream;
    ulk_set_alloc_stream(st);
    *out = nullptr;
    int rc = ulk_link_frames_impl(st, src, dst, dst_shift_host, params, out);
    if (rc && *out) { ulk_links_free(*out); *out = nullptr; }
    return rc;
}

int ulk_node_stats(const ulk_frame_nodes *nodes, const void *image, int image_dtype,
                   int32_t channel, int32_t n_channels, double *mean, double *std, double *vmin, double *vmax,
                   double *centroid, void *stream)
{
    if (!nodes) ULK_FAIL(-1, "null argument");
    if (n_channels < 1 || channel < 0 || channel >= n_channels) ULK_FAIL(-1, "bad channel %d of %d", channel, n_channels);
    if (nodes->n_nodes > 0 && !nodes->mask) ULK_FAIL(-1, "node table without masks");
    if (!image && (mean || std || vmin || vmax)) ULK_FAIL(-1, "intensity statistics need an image");
    ulk_stream_t st = (ulk_stream_t)stream;
    ulk_set_alloc_stream(st);
    return ulk_node_stats_impl(st, nodes, image, image_dtype, channel, n_channels, mean, std, vmin, vmax, centroid);
}

int ulk_links_to_host(const ulk_links *l, int32_t *source, int32_t *target, double *weight, void *stream)
{
    if (!l) ULK_FAIL(-1, "null argument");
    ulk_stream_t st = (ulk_stream_t)stream;
    if (l->n_links <= 0) return 0;
    if (source) API_CHK(ulk_d2h(source, l->source, (size_t)l->n_links * 4, st));
    if (target) API_CHK(ulk_d2h(target, l->target, (size_t)l->n_links * 4, st));
    if (weight) API_CHK(ulk_d2h(weight, l->weight, (size_t)l->n_links * 8, st));
    return 0;
}

void ulk_links_free(ulk_links *l)
{
    if (!l) return;
    ulk_dev_free(l->source); ulk_dev_free(l->target); ulk_dev_free(l->weight);
    delete l;
}

int ulk_frame_blobs_to_host(const ulk_frame_nodes *nodes, const uint8_t *head, const uint8_t *tail,
                            const ulk_blob_spec *spec, int64_t id_base, int32_t time,
                            const int64_t *blob_off_host, uint8_t *out_host, void *stream)
{
    if (!nodes || !head || !tail || !spec || !blob_off_host || !out_host) ULK_FAIL(-1, "null argument");
    if (spec->ndim != 2 && spec->ndim != 3) ULK_FAIL(-1, "blob spec: ndim must be 2 or 3");
    if (nodes->n_nodes > 0 && !nodes->mask) ULK_FAIL(-1, "node table without masks");
    ulk_stream_t st = (ulk_stream_t)stream;
    ulk_set_alloc_stream(st);
    return ulk_frame_blobs_impl(st, nodes, head, tail, spec, id_base, time, blob_off_host, out_host);
}

int ulk_labels_to_contours(const int32_t *labels, int32_t n_sets, const int64_t shape[3],
                           uint8_t *foreground, float *contours, void *stream)
{
    if (!labels || !shape || !foreground || !contours || n_sets <= 0) ULK_FAIL(-1, "bad argument");
    ulk_stream_t st = (ulk_stream_t)stream;
    ulk_set_alloc_stream(st);
    return ulk_labels_to_contours_impl(st, labels, n_sets, shape, foreground, contours);
}

int ulk_smooth_contours(float *contours, const int64_t shape[3], const double sigma[3], void *stream)
{
    if (!contours || !shape || !sigma) ULK_FAIL(-1, "null argument");
    ulk_stream_t st = (ulk_stream_t)stream;
    ulk_set_alloc_stream(st);
    return ulk_smooth_contours_impl(st, contours, shape, sigma);
}

int ulk_node_shifts(const ulk_frame_nodes *nodes, const void *field, int field_dtype, int32_t n_vec,
                    const int64_t field_shape[3], double *shift_host, void *stream)
{
    if (!nodes || !field || !field_shape || !shift_host) ULK_FAIL(-1, "null argument");
    if (n_vec < 1 || n_vec > 3) ULK_FAIL(-1, "n_vec must be 1, 2 or 3");
    ulk_stream_t st = (ulk_stream_t)stream;
    ulk_set_alloc_stream(st);
    return ulk_node_shifts_impl(st, nodes, field, field_dtype, n_vec, field_shape, shift_host);
}

int ulk_synth_frame(const int64_t shape[3], int32_t t, uint64_t seed, double radius, double fill,
                    double motion, double noise, float *foreground, float *contours, void *stream)
{
    if (!shape || !foreground || !contours) ULK_FAIL(-1, "null argument");
    ulk_stream_t st = (ulk_stream_t)stream;
    ulk_set_alloc_stream(st);
    return ulk_synth_frame_impl(st, shape, t, seed, radius, fill, motion, noise, foreground, contours);
}

}  // extern "C"
