/*
 * ultrack_b200.h -- C ABI of the B200-native candidate-generation path.
 *
 * ultrack (the reference) has no FFI layer on this path: segment() and link() are Python
 * functions whose arithmetic lives in numba / higra / scipy.  This library sits UNDER the
 * same two Python entry points (ultrack_b200.segment / ultrack_b200.link mirror
 * ultrack/core/segmentation/processing.py:393 and ultrack/core/linking/processing.py:294)
 * and replaces, per timepoint, the bodies of
 *
 *   ulk_segment_frame*   <- core/segmentation/processing.py:160-340   (_process: threshold,
 *                           create_hierarchies, Hierarchy.nodes, ids / overlaps)
 *                           core/segmentation/hierarchy.py:33-90, vendored/hierarchy.py:125-365,
 *                           vendored/node.py:27-141, node.py:229-233
 *   ulk_link_frames      <- core/linking/processing.py:172-291        (compute_spatial_neighbors:
 *                           KD-tree k-NN, IoU, top-k) and core/segmentation/node.py:23-118
 *   ulk_labels_to_contours <- utils/edge.py:18-96                      (labels_to_contours)
 *   ulk_node_shifts      <- imgproc/flow.py:639-751                   (add_flow, per timepoint)
 *
 * Conventions: every function returns 0 on success and a negative code on failure and never
 * throws; ulk_last_error() gives the message of the calling thread's last failure.  Code -10
 * is the reference's RuntimeError("Region too small...") (vendored/hierarchy.py:161-162).
 * Frames are C-contiguous [D,H,W] (2-D frames: D = 1).  Device pointers unless the name
 * says _host.  Results are owned by the library until the matching ulk_*_free.
 */
#ifndef ULTRACK_B200_H
#define ULTRACK_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ULK_API __attribute__((visibility("default")))

/* element types of foreground / contours / labels */
enum { ULK_DT_U8 = 0, ULK_DT_I8 = 1, ULK_DT_U16 = 2, ULK_DT_I16 = 3, ULK_DT_F16 = 4, ULK_DT_F32 = 5,
       ULK_DT_F64 = 6, ULK_DT_I32 = 7, ULK_DT_U32 = 8, ULK_DT_I64 = 9, ULK_DT_U64 = 10 };

/* SegmentationConfig fields _process reads (config/segmentationconfig.py:19-57).
 * `threshold` must already be rounded to the foreground dtype when that is a float type
 * (numpy compares `foreground > threshold` in the array's own precision). */
typedef struct ulk_seg_params {
    double threshold;
    double min_area;
    double max_area;
    double min_area_factor;
    double min_frontier;
    int32_t ws_hierarchy;      /* 0 = area, 1 = dynamics, 2 = volume (hg.watershed_hierarchy_by_*) */
    int32_t reserved;
} ulk_seg_params;

typedef struct ulk_seg_stats {
    int64_t n_basins, n_rag_edges, plateau_voxels;
    int32_t flow_rounds, k3_rounds;
    int64_t n_foreground;        /* voxels above the threshold                                    */
    int64_t pass_a_voxels;       /* voxels of components above max_area (oversegmentation pass)   */
    int64_t k3_voxels;           /* voxels of basins whose raster-key bottleneck needed a check   */
} ulk_seg_stats;

/* Candidate nodes of one frame, device-resident struct of arrays.  Node order is the
 * reference's emission order (t_node_id = index + 1). */
typedef struct ulk_frame_nodes {
    int32_t n_nodes, n_overlaps, n_hier, fix_shift;
    int32_t shape[3];
    int32_t pad0;
    int64_t mask_words;
    int32_t *bbox;        /* [n,6]  z0,y0,x0,z1,y1,x1 (max exclusive), frame coordinates        */
    int32_t *centroid;    /* [n,3]  z,y,x (node.py:229-233)                                     */
    int32_t *area;        /* [n]                                                                */
    double *frontier;     /* [n]    hg.attribute_contour_strength, -1 for the root fallback     */
    double *height;       /* [n]    hg.attribute_height,           -1 for the root fallback     */
    int32_t *hier_id;     /* [n]    t_hier_id (1-based region index)                            */
    int32_t *parent;      /* [n]    node index of the hierarchy parent, -1 if none emitted      */
    int32_t *hidx;        /* [n]    Node._h_node_index                                          */
    int64_t *mask_off;    /* [n+1]  offset of each node's mask in 32-bit words                  */
    uint32_t *mask;       /* bit-packed masks: rows of ceil(bw/32) words over the bbox          */
    int32_t *overlaps;    /* [m,2]  (node index, ancestor index) rows of the overlaps table     */
    int32_t *labels;      /* [D,H,W] region labels after oversegmentation, or NULL              */
    ulk_seg_stats stats;
} ulk_frame_nodes;

/* LinkingConfig fields compute_spatial_neighbors reads (config/config.py:23-49) */
typedef struct ulk_link_params {
    double max_distance;
    double distance_weight;
    int32_t max_neighbors;
    int32_t has_scale;
    double scale[3];          /* z,y,x, applied to both centroids when has_scale != 0            */
    /* secondary callers of compute_spatial_neighbors (processing.py:183-184,261-264):            */
    int32_t weight_kind;      /* 0 = Node.IoU, 1 = Node.intersection (GT matching, match_gt.py:148-168) */
    int32_t edge_must_be_positive; /* drop candidates whose weight is <= 0                        */
    /* colour z-score gate (color_filtering_mask, processing.py:52-104); n_channels = 0: off.
     * Device arrays of per-node statistics from ulk_node_stats, row-major [n, n_channels]:
     * a candidate survives iff max_c |dst_mean - src_mean| / dst_std' <= z_score_threshold,
     * dst_std' = 1 where dst_std <= 1e-6.                                                        */
    int32_t n_channels;
    int32_t pad0;
    double z_score_threshold;
    const double *src_mean;
    const double *dst_mean;
    const double *dst_std;
} ulk_link_params;

typedef struct ulk_links {
    int64_t n_links;
    int32_t *source;          /* [n] node index in the source frame                             */
    int32_t *target;          /* [n] node index in the target frame                             */
    double *weight;           /* [n] IoU (or intersection) - distance_weight * distance         */
    int64_t n_pairs;          /* candidate pairs evaluated (for the byte model)                 */
} ulk_links;

ULK_API const char *ulk_last_error(void);
ULK_API int ulk_device_count(void);
/* binds the calling thread to a device (one process per GPU: LOCAL_RANK) */
ULK_API int ulk_set_device(int device);
/* kernels launched by this library on the calling process so far */
ULK_API int64_t ulk_launch_count(void);
/* scratch workspace of the calling process, bytes: [0] retained slab capacity, [1] high-water
 * mark of scratch in use during a call, [2] most bytes ever served outside the slab */
ULK_API void ulk_workspace_stats(int64_t out[3]);

/* per-kernel timing for bench.py's roofline line: when enabled every launch is bracketed by a
 * CUDA event pair on its own stream.  report: "kernel name\tlaunches\ttotal ms\n" per kernel;
 * returns the buffer size the full report needs. */
ULK_API void ulk_profile_enable(int on);
ULK_API void ulk_profile_reset(void);
ULK_API int ulk_profile_report(char *buf, int64_t cap);

/* stream: a cudaStream_t (0 = default stream); all work of a call is ordered on it and the
 * call returns after the results are complete. */
ULK_API int ulk_segment_frame(const void *foreground, int fg_dtype, const void *contours, int ct_dtype,
                      const int64_t shape[3], const ulk_seg_params *params, int keep_labels,
                      void *stream, ulk_frame_nodes **out);
/* same, inputs in (ideally pinned) host memory: the copies are part of the call */
ULK_API int ulk_segment_frame_host(const void *foreground, int fg_dtype, const void *contours, int ct_dtype,
                           const int64_t shape[3], const ulk_seg_params *params, int keep_labels,
                           void *stream, ulk_frame_nodes **out);
/* pipelined ingest (processing.py:204,212 read `foreground[t]` / `contours[t]` per frame):
 * ulk_upload_begin starts the host->device copy of a frame on the library's own copy stream and
 * returns at once (asynchronous only for pinned host memory); ulk_segment_frame_uploaded waits
 * for it on `stream`, segments the frame and releases the upload.  Begin frame t+1 before
 * segmenting frame t and the copy is hidden behind the kernels.  At most two uploads in flight. */
typedef struct ulk_frame_upload ulk_frame_upload;
ULK_API int ulk_upload_begin(const void *foreground_host, int fg_dtype, const void *contours_host, int ct_dtype,
                     const int64_t shape[3], ulk_frame_upload **out);
ULK_API int ulk_segment_frame_uploaded(ulk_frame_upload *upload, const ulk_seg_params *params, int keep_labels,
                               void *stream, ulk_frame_nodes **out);
ULK_API void ulk_upload_free(ulk_frame_upload *upload);
ULK_API void ulk_frame_nodes_free(ulk_frame_nodes *nodes);
/* copies the device arrays into caller-provided host buffers (any may be NULL) */
ULK_API int ulk_frame_nodes_to_host(const ulk_frame_nodes *nodes, int32_t *bbox, int32_t *centroid, int32_t *area,
                            double *frontier, double *height, int32_t *hier_id, int32_t *parent,
                            int32_t *hidx, int64_t *mask_off, uint32_t *mask, int32_t *overlaps,
                            int32_t *labels, void *stream);
/* builds a device node table from host arrays (a boundary frame received from another rank,
 * or nodes loaded back from the database) */
ULK_API int ulk_frame_nodes_from_host(int32_t n_nodes, const int32_t shape[3], const int32_t *bbox,
                              const int32_t *centroid, const int32_t *area, const int64_t *mask_off,
                              const uint32_t *mask, void *stream, ulk_frame_nodes **out);

/* same two, for device-resident buffers owned by the caller (a torch tensor received over
 * NCCL from the rank that owns the next timepoint shard): d2d copies, no host round trip */
ULK_API int ulk_frame_nodes_to_device(const ulk_frame_nodes *nodes, int32_t *bbox, int32_t *centroid, int32_t *area,
                              int64_t *mask_off, uint32_t *mask, void *stream);
ULK_API int ulk_frame_nodes_from_device(int32_t n_nodes, const int32_t shape[3], const int32_t *bbox,
                                const int32_t *centroid, const int32_t *area, const int64_t *mask_off,
                                int64_t mask_words, const uint32_t *mask, void *stream, ulk_frame_nodes **out);

/* dst_shift: host [n_dst,3] float64 (z,y,x shifts of the target nodes) or NULL; for 2-D
 * frames the z column must be 0 (the reference slices `shift[:, -ndim:]`) */
ULK_API int ulk_link_frames(const ulk_frame_nodes *src, const ulk_frame_nodes *dst, const double *dst_shift_host,
                    const ulk_link_params *params, void *stream, ulk_links **out);
/* Per-node reductions over the node masks.  With `image` (a device [D,H,W] volume of
 * `image_dtype`): Node.intensity_mean / Node.intensity_std (core/segmentation/node.py:280-296)
 * and the intensity min / max of the regionprops features (processing.py:107-157) into column
 * `channel` of the device double arrays mean / std / vmin / vmax [n_nodes, n_channels];
 * `centroid` [n_nodes, 3] receives the float64 mean (z, y, x) of the mask coordinates.  Any
 * output may be NULL; `image` may be NULL when only the centroid is wanted. */
ULK_API int ulk_node_stats(const ulk_frame_nodes *nodes, const void *image, int image_dtype,
                   int32_t channel, int32_t n_channels, double *mean, double *std, double *vmin, double *vmax,
                   double *centroid, void *stream);
ULK_API int ulk_links_to_host(const ulk_links *links, int32_t *source, int32_t *target, double *weight, void *stream);
ULK_API void ulk_links_free(ulk_links *links);

/* Bulk writer of the `nodes.pickle` column: pickle.dumps(Node) of every node of a
 * frame (core/segmentation/processing.py:254-267; Node.__getstate__, node.py:270-278) in one
 * kernel.  Blob i = head template + the node's bool mask as raw bytes + tail template, with the
 * per-node fields patched into fixed-width holes at the offsets of `spec` (built by
 * ultrack_b200/node.py:BlobTemplate, which documents the pickle stream).  blob_off_host[n+1] =
 * byte offset of every blob in `out_host` (blob size = head_len + bbox volume + tail_len);
 * ids are id_base + node index + 1 (processing.py:42-59). */
typedef struct ulk_blob_spec {
    int32_t head_len, tail_len, ndim, payload_start, payload_tail;
    int32_t off_area, off_frontier, off_height, off_hidx, off_bbox, off_chunk_len;
    int32_t off_blosc_nbytes, off_blosc_blocksize, off_blosc_cbytes, off_shape, off_raw_len;
    int32_t toff_id, toff_time, toff_centroid;
} ulk_blob_spec;
ULK_API int ulk_frame_blobs_to_host(const ulk_frame_nodes *nodes, const uint8_t *head, const uint8_t *tail,
                            const ulk_blob_spec *spec, int64_t id_base, int32_t time,
                            const int64_t *blob_off_host, uint8_t *out_host, void *stream);

/* utils/edge.py:18-96 for one frame: n_sets label images [n_sets,D,H,W] (int32, device) ->
 * foreground uint8 [D,H,W] and contours float32 [D,H,W] (sigma = None path) */
ULK_API int ulk_labels_to_contours(const int32_t *labels, int32_t n_sets, const int64_t shape[3],
                           uint8_t *foreground, float *contours, void *stream);

/* the `sigma` branch of labels_to_contours (utils/edge.py:89-91), in place on a device float32
 * [D,H,W] volume: scipy.ndimage.gaussian_filter(contours, sigma) (reflect boundary, window radius
 * int(4 sigma + 0.5), float64 taps, float32 result per axis) then division by the global maximum.
 * sigma: z, y, x (z ignored when D == 1). */
ULK_API int ulk_smooth_contours(float *contours, const int64_t shape[3], const double sigma[3], void *stream);

/* No counterpart in the reference: added for callers that
 * cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
 * ulk_smooth_contours with NO allocation of any kind: the ping-pong volume and the maximum live in
 * a file-scope device array of ULK_SMOOTH_STATIC_MAX_VOXELS float32 (+ 256 B), held by every GPU
 * that loads the library until exit; the Gaussian taps travel in the kernel parameters.  Results
 * are bit-identical to ulk_smooth_contours.  Limits: D*H*W <= ULK_SMOOTH_STATIC_MAX_VOXELS and
 * int(4 sigma + 0.5) <= ULK_SMOOTH_STATIC_MAX_RADIUS per filtered axis; beyond them the call fails
 * with code -12 and a message (use ulk_smooth_contours).  The call only ENQUEUES work on `stream`
 * and does not synchronise it (no host read-back: the maximum stays on the device).  All calls on
 * a device share the one array: calls that overlap on the GPU (different streams or host threads)
 * race -- serialise them. */
#define ULK_SMOOTH_STATIC_MAX_VOXELS 4194304        /* 16 MiB of float32: 2048 x 2048 or 64 x 256 x 256 */
#define ULK_SMOOTH_STATIC_MAX_RADIUS 127            /* sigma < 31.875 */
ULK_API int ulk_smooth_contours_static(float *contours, const int64_t shape[3], const double sigma[3], void *stream);
ULK_API int64_t ulk_smooth_contours_static_max_voxels(void);

/* add_flow (imgproc/flow.py:639-751), the per-timepoint kernel: for every node of `nodes` the
 * vector field sampled at its centroid mapped into field coordinates (round half to even of
 * centroid * (field_dim - 1) / (image_dim - 1), clipped: flow.py:716-720), scaled to voxels
 * (flow.py:737).  `field` is a device array [n_vec, FD, FH, FW] of `field_dtype` for ONE
 * timepoint; with fewer vectors than dimensions the last axes have priority (the last vector is
 * x).  shift_host receives [n,3] doubles (z, y, x), the layout ulk_link_frames takes. */
ULK_API int ulk_node_shifts(const ulk_frame_nodes *nodes, const void *field, int field_dtype, int32_t n_vec,
                    const int64_t field_shape[3], double *shift_host, void *stream);

/* seeded synthetic frame on the device (ultrack_b200/synth.py recipe, variant "S"):
 * foreground float32 and contours float32 [D,H,W] */
ULK_API int ulk_synth_frame(const int64_t shape[3], int32_t t, uint64_t seed, double radius, double fill,
                    double motion, double noise, float *foreground, float *contours, void *stream);

#ifdef __cplusplus
}
#endif
#endif
