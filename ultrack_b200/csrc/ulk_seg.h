// ulk_seg.h -- internal declarations shared by the pipeline files.
#pragma once
#include "../../include/ultrack_b200.h"
#include "ulk_rt.h"

int ulk_segment_frame_impl(ulk_stream_t st, const void *fg, int fg_dtype, const void *ct, int ct_dtype,
                           const int64_t shape[3], const ulk_seg_params *prm, int keep_labels,
                           ulk_frame_nodes **result);
void ulk_frame_nodes_free_impl(ulk_frame_nodes *f);
int ulk_link_frames_impl(ulk_stream_t st, const ulk_frame_nodes *src, const ulk_frame_nodes *dst,
                         const double *dst_shift_host, const ulk_link_params *prm, ulk_links **out);
int ulk_node_stats_impl(ulk_stream_t st, const ulk_frame_nodes *f, const void *image, int dtype,
                        int32_t channel, int32_t n_channels, double *mean, double *sd, double *vmin, double *vmax,
                        double *centroid);
int ulk_frame_blobs_impl(ulk_stream_t st, const ulk_frame_nodes *f, const uint8_t *head, const uint8_t *tail,
                         const ulk_blob_spec *sp, int64_t id_base, int32_t time, const int64_t *blob_off_host,
                         uint8_t *out_host);
int ulk_node_shifts_impl(ulk_stream_t st, const ulk_frame_nodes *nodes, const void *field, int field_dtype, int32_t n_vec,
                         const int64_t field_shape[3], double *shift_host);
int ulk_labels_to_contours_impl(ulk_stream_t st, const int32_t *labels, int32_t n_sets, const int64_t shape[3],
                                uint8_t *foreground, float *contours);
int ulk_smooth_contours_impl(ulk_stream_t st, float *contours, const int64_t shape[3], const double sigma[3]);
int ulk_synth_frame_impl(ulk_stream_t st, const int64_t shape[3], int32_t t, uint64_t seed, double radius,
                         double fill, double motion, double noise, float *fg, float *ct);
const char *ulk_error_cstr();
void ulk_set_alloc_stream(ulk_stream_t s);
