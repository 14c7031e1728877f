// ulk_link.cu -- adjacent-frame candidate linking on device-resident node tables.
//
// Replaces ultrack/core/linking/processing.py:172-291 (compute_spatial_neighbors) and the
// numba IoU kernels of ultrack/core/segmentation/node.py:23-118:
//   * scipy KDTree.query(k = 2*max_neighbors, distance_upper_bound = max_distance) becomes a
//     uniform-grid search (cell edge >= max_distance) keeping the k nearest by (d^2, index);
//   * IoU of bbox-aligned bool masks becomes AND + popcount over bit-packed rows;
//   * per target: weight = IoU - distance_weight * distance, keep the max_neighbors best by
//     (weight, -distance, source id) descending (processing.py:269).
#include "ulk_seg.h"
#include "ulk_common.h"
#include <cmath>

#define CHK(x) do { int rc__ = (x); if (rc__) return rc__; } while (0)
#define CHKA(a) do { if ((a).failed) ULK_FAIL(-3, "device allocation failed"); } while (0)

struct LinkGeom {
    double mn[3];      // grid origin
    double cs;         // cell edge
    i32 nc[3];         // cells per axis
    double scale[3];
    int has_scale;
    double ub, ub2;    // max_distance and its square
};

// processing.py:187-219: sources are integer centroids, targets float32 centroids + shift,
// both scaled afterwards
ULK_HD ULK_INLINE void src_pos(const i32 *c, const LinkGeom &g, double p[3])
{
    for (int d = 0; d < 3; ++d) p[d] = g.has_scale ? (double)c[d] * g.scale[d] : (double)c[d];
}
ULK_HD ULK_INLINE void dst_pos(const i32 *c, const double *shift, const LinkGeom &g, double p[3])
{
    for (int d = 0; d < 3; ++d) {
        float f = (float)c[d];
        if (shift) f = (float)((double)f + shift[d]);
        p[d] = g.has_scale ? (double)f * g.scale[d] : (double)f;
    }
}
ULK_HD ULK_INLINE i32 cell_of(double p, double mn, double cs, i32 nc)
{
    double q = floor((p - mn) / cs);
    if (q < 0) q = 0;
    if (q > nc - 1) q = nc - 1;
    return (i32)q;
}

struct K_SrcBounds {
    const i32 *cen; LinkGeom g; u64 *mn, *mx;
    ULK_D void operator()(i64 i) const
    {
        double p[3];
        src_pos(cen + 3 * i, g, p);
        for (int d = 0; d < 3; ++d) {
            ulk_atomic_min(&mn[d], ulk_d2ord(p[d]));
            ulk_atomic_max(&mx[d], ulk_d2ord(p[d]));
        }
    }
#ifndef ULK_EMU
    ULK_D void warp(i64 i, bool valid) const     // one pair of atomics per axis and warp
    {
        u64 lo[3] = {ULK_NONE64, ULK_NONE64, ULK_NONE64}, hi[3] = {0, 0, 0};
        if (valid) {
            double p[3];
            src_pos(cen + 3 * i, g, p);
            for (int d = 0; d < 3; ++d) lo[d] = hi[d] = ulk_d2ord(p[d]);
        }
        for (int d = 0; d < 3; ++d) {
            for (int o = 16; o > 0; o >>= 1) {
                u64 a = __shfl_xor_sync(0xffffffffu, lo[d], o), b = __shfl_xor_sync(0xffffffffu, hi[d], o);
                lo[d] = a < lo[d] ? a : lo[d];
                hi[d] = b > hi[d] ? b : hi[d];
            }
            if ((threadIdx.x & 31) == 0 && lo[d] != ULK_NONE64) {
                if (lo[d] < mn[d]) ulk_atomic_min(&mn[d], lo[d]);
                if (hi[d] > mx[d]) ulk_atomic_max(&mx[d], hi[d]);
            }
        }
    }
#endif
};
struct K_SrcCell {
    const i32 *cen; LinkGeom g; u32 *cell; u32 *idx; i32 *hist;
    ULK_D void operator()(i64 i) const
    {
        double p[3];
        src_pos(cen + 3 * i, g, p);
        i32 c = (cell_of(p[0], g.mn[0], g.cs, g.nc[0]) * g.nc[1] + cell_of(p[1], g.mn[1], g.cs, g.nc[1])) * g.nc[2]
                + cell_of(p[2], g.mn[2], g.cs, g.nc[2]);
        cell[i] = (u32)c;
        idx[i] = (u32)i;
        ulk_atomic_add(&hist[c], 1);
    }
};
// k nearest sources of every target, strict d < max_distance (scipy's upper bound is
// exclusive), ordered by (d^2, source index)
struct K_KnnSelect {
    const i32 *scen; const i32 *dcen; const double *shift; LinkGeom g;
    const u32 *sorted_idx; const i32 *cstart; i32 k;
    i32 *cidx; double *cd2; i32 *ccount;
    ULK_D void operator()(i64 t) const
    {
        double q[3];
        dst_pos(dcen + 3 * t, shift ? shift + 3 * t : nullptr, g, q);
        i32 lo[3], hi[3];
        for (int d = 0; d < 3; ++d) {
            lo[d] = cell_of(q[d] - g.ub, g.mn[d], g.cs, g.nc[d]);
            hi[d] = cell_of(q[d] + g.ub, g.mn[d], g.cs, g.nc[d]);
        }
        i32 *ci = cidx + t * k;
        double *cd = cd2 + t * k;
        i32 n = 0;
        for (i32 cz = lo[0]; cz <= hi[0]; ++cz)
            for (i32 cy = lo[1]; cy <= hi[1]; ++cy)
                for (i32 cx = lo[2]; cx <= hi[2]; ++cx) {
                    i32 c = (cz * g.nc[1] + cy) * g.nc[2] + cx;
                    for (i32 j = cstart[c]; j < cstart[c + 1]; ++j) {
                        i32 s = (i32)sorted_idx[j];
                        double p[3];
                        src_pos(scen + 3 * s, g, p);
                        double d2 = 0.0;
                        for (int d = 0; d < 3; ++d) { double df = p[d] - q[d]; d2 = d2 + df * df; }
                        if (!(d2 < g.ub2)) continue;
                        // insertion into the sorted top-k
                        i32 pos = n < k ? n : k;
                        while (pos > 0 && (cd[pos - 1] > d2 || (cd[pos - 1] == d2 && ci[pos - 1] > s))) --pos;
                        if (pos >= k) continue;
                        i32 last = n < k ? n : k - 1;
                        for (i32 m = last; m > pos; --m) { cd[m] = cd[m - 1]; ci[m] = ci[m - 1]; }
                        cd[pos] = d2; ci[pos] = s;
                        if (n < k) ++n;
                    }
                }
        ccount[t] = n;
    }
};

// node.py:23-118 on bit-packed masks.  bbox rows are packed along x, 32 voxels per word.
ULK_D ULK_INLINE u32 mask_bits32(const u32 *row, i64 wpr, i64 bit)
{
    // 32 bits of the row starting at `bit` (bits past the row read as 0)
    i64 w = bit >> 5;
    u32 s = (u32)(bit & 31);
    u32 lo = w < wpr ? row[w] : 0u;
    u32 hi = (w + 1) < wpr ? row[w + 1] : 0u;
    return ulk_funnel_r(lo, hi, s);
}
struct K_PairIoU {
    const i32 *sbb, *sarea; const i64 *soff; const u32 *smask;
    const i32 *dbb, *darea; const i64 *doff; const u32 *dmask;
    const double *shift; i32 k; const i32 *cidx; const double *cd2; const i32 *ccount;
    double dw; double *cw; double *cdist; i64 slots;
    int weight_kind, positive_only;        // processing.py:183-184,261-264; match_gt.py:148-168
    int nch; double zthr; const double *smean, *dmean, *dstd;   // colour z-score gate, processing.py:52-104
    u8 *cvalid;
    // the candidate's weight, and whether it survives the colour gate and the positive-only rule
    ULK_D ULK_INLINE void finish(i64 i, i64 t, i32 s, i64 inter) const
    {
        const double dist = sqrt(cd2[i]);
        cdist[i] = dist;
        double wgt = 0.0;
        if (weight_kind == 1) wgt = (double)inter;
        else if (inter > 0) wgt = (double)inter / (double)((i64)darea[t] + (i64)sarea[s] - inter);
        const double pen = dw * dist;
        wgt = wgt - pen;
        cw[i] = wgt;
        bool ok = !(positive_only && wgt <= 0.0);
        if (ok && nch > 0) {
            double worst = 0.0;
            for (int c = 0; c < nch; ++c) {
                double sd = dstd[t * nch + c];
                if (sd <= 1e-6) sd = 1.0;
                double df = (dmean[t * nch + c] - smean[(i64)s * nch + c]) / sd;
                df = df < 0 ? -df : df;
                if (df > worst || df != df) worst = df;
            }
            ok = worst <= zthr;
        }
        cvalid[i] = ok ? 1 : 0;
    }
#ifndef ULK_EMU
    // one warp per candidate pair: lanes take the (z, y) rows of the bbox overlap window
    ULK_D void warp(i64 ii, bool) const
    {
        const int lane = threadIdx.x & 31;
        const i64 i = ii >> 5;
        if (i >= slots) return;
        i64 t = i / k;
        i32 slot = (i32)(i % k);
        if (slot >= ccount[t]) return;
        i32 s = cidx[i];
        i32 b1[6], b2[6];
        for (int d = 0; d < 6; ++d) { b1[d] = dbb[6 * t + d]; b2[d] = sbb[6 * s + d]; }
        if (shift)
            for (int d = 0; d < 3; ++d) { i32 r = (i32)rint(shift[3 * t + d]); b1[d] += r; b1[d + 3] += r; }
        i32 lo[3], hi[3];
        bool hit = true;
        for (int d = 0; d < 3; ++d) {
            lo[d] = b1[d] > b2[d] ? b1[d] : b2[d];
            hi[d] = b1[d + 3] < b2[d + 3] ? b1[d + 3] : b2[d + 3];
            if (lo[d] >= hi[d]) hit = false;
        }
        u32 cnt = 0;
        if (hit) {
            i64 h1 = b1[4] - b1[1], w1 = (b1[5] - b1[2] + 31) / 32;
            i64 h2 = b2[4] - b2[1], w2 = (b2[5] - b2[2] + 31) / 32;
            const u32 *m1 = dmask + doff[t], *m2 = smask + soff[s];
            const i32 nx = hi[2] - lo[2], ny = hi[1] - lo[1];
            const i32 rows = (hi[0] - lo[0]) * ny;      // a window of one node: far below 2^31
            for (i32 r = lane; r < rows; r += 32) {
                i32 z = lo[0] + r / ny, y = lo[1] + r % ny;
                const u32 *r1 = m1 + ((i64)(z - b1[0]) * h1 + (y - b1[1])) * w1;
                const u32 *r2 = m2 + ((i64)(z - b2[0]) * h2 + (y - b2[1])) * w2;
                for (i32 x = 0; x < nx; x += 32) {
                    u32 a = mask_bits32(r1, w1, (i64)(lo[2] - b1[2]) + x);
                    u32 b = mask_bits32(r2, w2, (i64)(lo[2] - b2[2]) + x);
                    u32 m = a & b;
                    i32 rem = nx - x;
                    if (rem < 32) m &= (1u << rem) - 1u;
                    cnt += ulk_popc(m);
                }
            }
        }
        // per-lane counts stay below 2^32 for any window that fits a frame of < 2^31 voxels
        i64 inter = (i64)__reduce_add_sync(0xffffffffu, cnt);
        if (lane != 0) return;
        finish(i, t, s, inter);
    }
#endif
    ULK_D void operator()(i64 i) const
    {
        i64 t = i / k;
        i32 slot = (i32)(i % k);
        if (slot >= ccount[t]) return;
        i32 s = cidx[i];
        i32 b1[6], b2[6];
        for (int d = 0; d < 6; ++d) { b1[d] = dbb[6 * t + d]; b2[d] = sbb[6 * s + d]; }
        if (shift)
            for (int d = 0; d < 3; ++d) { i32 r = (i32)rint(shift[3 * t + d]); b1[d] += r; b1[d + 3] += r; }
        i32 lo[3], hi[3];
        bool hit = true;
        for (int d = 0; d < 3; ++d) {
            lo[d] = b1[d] > b2[d] ? b1[d] : b2[d];
            hi[d] = b1[d + 3] < b2[d + 3] ? b1[d + 3] : b2[d + 3];
            if (lo[d] >= hi[d]) hit = false;
        }
        i64 inter = 0;
        if (hit) {
            i64 h1 = b1[4] - b1[1], w1 = (b1[5] - b1[2] + 31) / 32;
            i64 h2 = b2[4] - b2[1], w2 = (b2[5] - b2[2] + 31) / 32;
            const u32 *m1 = dmask + doff[t], *m2 = smask + soff[s];
            i32 nx = hi[2] - lo[2];
            for (i32 z = lo[0]; z < hi[0]; ++z)
                for (i32 y = lo[1]; y < hi[1]; ++y) {
                    const u32 *r1 = m1 + ((i64)(z - b1[0]) * h1 + (y - b1[1])) * w1;
                    const u32 *r2 = m2 + ((i64)(z - b2[0]) * h2 + (y - b2[1])) * w2;
                    for (i32 x = 0; x < nx; x += 32) {
                        u32 a = mask_bits32(r1, w1, (i64)(lo[2] - b1[2]) + x);
                        u32 b = mask_bits32(r2, w2, (i64)(lo[2] - b2[2]) + x);
                        u32 m = a & b;
                        i32 rem = nx - x;
                        if (rem < 32) m &= (1u << rem) - 1u;
                        inter += ulk_popc(m);
                    }
                }
        }
        finish(i, t, s, inter);
    }
};
// processing.py:269: sorted(reverse=True)[:max_neighbors] on (weight, -dist, source_id)
struct K_LinkTopK {
    i32 k, maxn; const i32 *ccount; i32 *cidx; double *cw; double *cdist; i32 *nout; const u8 *cvalid;
    ULK_D void operator()(i64 t) const
    {
        i32 n = ccount[t];
        i32 *ci = cidx + t * k;
        double *w = cw + t * k, *ds = cdist + t * k;
        const u8 *ok = cvalid + t * k;
        i32 m = 0;
        for (i32 a = 0; a < n; ++a)
            if (ok[a]) { ci[m] = ci[a]; w[m] = w[a]; ds[m] = ds[a]; ++m; }
        n = m;
        for (i32 a = 1; a < n; ++a) {     // insertion sort, n <= k is small
            i32 si = ci[a]; double sw = w[a], sd = ds[a];
            i32 p = a;
            while (p > 0) {
                bool before = sw > w[p - 1] || (sw == w[p - 1] && (sd < ds[p - 1] || (sd == ds[p - 1] && si > ci[p - 1])));
                if (!before) break;
                ci[p] = ci[p - 1]; w[p] = w[p - 1]; ds[p] = ds[p - 1];
                --p;
            }
            ci[p] = si; w[p] = sw; ds[p] = sd;
        }
        nout[t] = n < maxn ? n : maxn;
    }
};
struct K_LinkGather {
    i32 k; const i32 *nout; const i64 *ooff; const i32 *cidx; const double *cw;
    i32 *src, *dst; double *wt;
    ULK_D void operator()(i64 t) const
    {
        i64 o = ooff[t];
        for (i32 j = 0; j < nout[t]; ++j) { src[o + j] = cidx[t * k + j]; dst[o + j] = (i32)t; wt[o + j] = cw[t * k + j]; }
    }
};
struct K_I32ToI64 { const i32 *a; i64 *b; ULK_D void operator()(i64 i) const { b[i] = a[i]; } };

// Per-node reductions over the bit-packed masks: Node.intensity_mean / Node.intensity_std
// (core/segmentation/node.py:280-296) of one image channel for the colour gate, plus what the
// regionprops features of processing.py:107-157 need (intensity min / max, float centroid).
// One warp per node, lanes over the (z, y) rows of its bbox; two passes over the mask (numpy's
// std is the root of the mean squared deviation from the mean).
struct UlkNodeAcc { double s, mn, mx; i64 cnt, cz, cy, cx; };
struct K_NodeStats {
    const i32 *bb; const i64 *moff; const u32 *mask; const void *img; int dt; i32 H, W;
    i32 nch, ch; double *mean, *sd, *vmin, *vmax, *cen; i64 n;
    ULK_D ULK_INLINE void pass1(const i32 *b, const u32 *m, int lane, int step, UlkNodeAcc &a) const
    {
        const i32 h = b[4] - b[1], w = b[5] - b[2], wpr = (w + 31) / 32;
        const i32 nrow = (b[3] - b[0]) * h;
        for (i32 r = lane; r < nrow; r += step) {
            const i32 z = b[0] + r / h, y = b[1] + r % h;
            const i64 base = ((i64)z * H + y) * W + b[2];
            const u32 *row = m + (i64)r * wpr;
            for (i32 q = 0; q < wpr; ++q)
                for (u32 bits = row[q]; bits; bits &= bits - 1) {
                    const int bit = ulk_popc((bits & (0u - bits)) - 1u);
                    const i32 xo = q * 32 + bit;
                    ++a.cnt; a.cz += z; a.cy += y; a.cx += b[2] + xo;
                    if (img) {
                        const double v = ulk_load_as_double(img, dt, base + xo);
                        a.s += v;
                        if (v < a.mn) a.mn = v;
                        if (v > a.mx) a.mx = v;
                    }
                }
        }
    }
    ULK_D ULK_INLINE double pass2(const i32 *b, const u32 *m, int lane, int step, double mu) const
    {
        const i32 h = b[4] - b[1], w = b[5] - b[2], wpr = (w + 31) / 32;
        const i32 nrow = (b[3] - b[0]) * h;
        double acc = 0.0;
        for (i32 r = lane; r < nrow; r += step) {
            const i32 z = b[0] + r / h, y = b[1] + r % h;
            const i64 base = ((i64)z * H + y) * W + b[2];
            const u32 *row = m + (i64)r * wpr;
            for (i32 q = 0; q < wpr; ++q)
                for (u32 bits = row[q]; bits; bits &= bits - 1) {
                    const int bit = ulk_popc((bits & (0u - bits)) - 1u);
                    const double d = ulk_load_as_double(img, dt, base + q * 32 + bit) - mu;
                    acc += d * d;
                }
        }
        return acc;
    }
    ULK_D ULK_INLINE void store(i64 i, const UlkNodeAcc &a, double mu, double q) const
    {
        const double c = a.cnt > 0 ? (double)a.cnt : 1.0;
        if (cen) { cen[3 * i] = (double)a.cz / c; cen[3 * i + 1] = (double)a.cy / c; cen[3 * i + 2] = (double)a.cx / c; }
        if (!img) return;
        if (mean) mean[i * nch + ch] = mu;
        if (sd) sd[i * nch + ch] = a.cnt > 0 ? sqrt(q / c) : 0.0;
        if (vmin) vmin[i * nch + ch] = a.mn;
        if (vmax) vmax[i * nch + ch] = a.mx;
    }
#ifndef ULK_EMU
    ULK_D void warp(i64 ii, bool) const
    {
        const int lane = threadIdx.x & 31;
        const i64 i = ii >> 5;
        if (i >= n) return;
        i32 b[6];
        for (int d = 0; d < 6; ++d) b[d] = bb[6 * i + d];
        const u32 *m = mask + moff[i];
        UlkNodeAcc a{0.0, 1.0 / 0.0, -1.0 / 0.0, 0, 0, 0, 0};
        pass1(b, m, lane, 32, a);
        for (int o = 16; o > 0; o >>= 1) {
            a.s += __shfl_xor_sync(0xffffffffu, a.s, o);
            const double mn2 = __shfl_xor_sync(0xffffffffu, a.mn, o), mx2 = __shfl_xor_sync(0xffffffffu, a.mx, o);
            a.mn = mn2 < a.mn ? mn2 : a.mn; a.mx = mx2 > a.mx ? mx2 : a.mx;
            a.cnt += __shfl_xor_sync(0xffffffffu, a.cnt, o);
            a.cz += __shfl_xor_sync(0xffffffffu, a.cz, o);
            a.cy += __shfl_xor_sync(0xffffffffu, a.cy, o);
            a.cx += __shfl_xor_sync(0xffffffffu, a.cx, o);
        }
        const double mu = a.cnt > 0 ? a.s / (double)a.cnt : 0.0;
        double q = 0.0;
        if (img && sd) {
            q = pass2(b, m, lane, 32, mu);
            for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
        }
        if (lane == 0) store(i, a, mu, q);
    }
#endif
    ULK_D void operator()(i64 i) const
    {
        i32 b[6];
        for (int d = 0; d < 6; ++d) b[d] = bb[6 * i + d];
        const u32 *m = mask + moff[i];
        UlkNodeAcc a{0.0, 1.0 / 0.0, -1.0 / 0.0, 0, 0, 0, 0};
        pass1(b, m, 0, 1, a);
        const double mu = a.cnt > 0 ? a.s / (double)a.cnt : 0.0;
        const double q = (img && sd) ? pass2(b, m, 0, 1, mu) : 0.0;
        store(i, a, mu, q);
    }
};

int ulk_node_stats_impl(ulk_stream_t st, const ulk_frame_nodes *f, const void *image, int dtype,
                        int32_t channel, int32_t n_channels, double *mean, double *sd, double *vmin, double *vmax,
                        double *centroid)
{
    const i64 n = f->n_nodes;
    if (n <= 0) return 0;
    K_NodeStats k{f->bbox, f->mask_off, f->mask, image, dtype, f->shape[1], f->shape[2], n_channels, channel,
                  mean, sd, vmin, vmax, centroid, n};
#ifdef ULK_EMU
    ulk_launch(n, st, k);
#else
    ulk_launch_warp(n * 32, st, k);
#endif
    return ulk_sync(st);
}

int ulk_link_frames_impl(ulk_stream_t st, const ulk_frame_nodes *src, const ulk_frame_nodes *dst,
                         const double *dst_shift_host, const ulk_link_params *prm, ulk_links **out)
{
    ulk_links *L = new ulk_links();
    memset(L, 0, sizeof(*L));
    *out = L;
    const i32 Ns = src->n_nodes, Nd = dst->n_nodes;
    const i32 k = 2 * prm->max_neighbors;
    if (Ns <= 0 || Nd <= 0 || k <= 0 || !(prm->max_distance > 0)) return 0;
    if (prm->weight_kind < 0 || prm->weight_kind > 1) ULK_FAIL(-1, "weight_kind=%d (0 IoU, 1 intersection)", prm->weight_kind);
    const int nch = prm->n_channels;
    if (nch < 0 || (nch > 0 && (!prm->src_mean || !prm->dst_mean || !prm->dst_std)))
        ULK_FAIL(-1, "colour gate: n_channels=%d needs src_mean, dst_mean and dst_std", nch);
    UlkArena A;
    LinkGeom g;
    g.has_scale = prm->has_scale;
    for (int d = 0; d < 3; ++d) g.scale[d] = prm->scale[d];
    g.ub = prm->max_distance;
    g.ub2 = prm->max_distance * prm->max_distance;
    double *shift = nullptr;
    if (dst_shift_host) {
        shift = A.get<double>(3 * (i64)Nd);
        CHKA(A);
        CHK(ulk_h2d(shift, dst_shift_host, (size_t)Nd * 3 * sizeof(double), st));
    }
    // ---- uniform grid over the source centroids
    u64 *bounds = A.get<u64>(6);
    CHKA(A);
    CHK(ulk_memset(bounds, 0xff, 3 * 8, st));
    CHK(ulk_memset(bounds + 3, 0, 3 * 8, st));
    for (int d = 0; d < 3; ++d) { g.mn[d] = 0; g.nc[d] = 1; }
    g.cs = 1;
    ulk_launch_warp(Ns, st, K_SrcBounds{src->centroid, g, bounds, bounds + 3});
    u64 hb[6];
    CHK(ulk_read(hb, bounds, 6, st));
    double mx[3];
    for (int d = 0; d < 3; ++d) {
        auto ord2d = [](u64 o) { u64 b = (o & 0x8000000000000000ull) ? (o & 0x7fffffffffffffffull) : ~o; double v; memcpy(&v, &b, 8); return v; };
        g.mn[d] = ord2d(hb[d]);
        mx[d] = ord2d(hb[3 + d]);
    }
    g.cs = prm->max_distance;
    for (;;) {
        double total = 1;
        for (int d = 0; d < 3; ++d) {
            double n = floor((mx[d] - g.mn[d]) / g.cs) + 1.0;
            if (n < 1) n = 1;
            if (n > 2000000000.0) n = 2000000000.0;
            g.nc[d] = (i32)n;
            total *= n;
        }
        if (total <= 16777216.0) break;
        g.cs *= 2.0;
    }
    i64 ncell = (i64)g.nc[0] * g.nc[1] * g.nc[2];
    u32 *cell = A.get<u32>(Ns), *idx = A.get<u32>(Ns), *cell2 = A.get<u32>(Ns), *sidx = A.get<u32>(Ns);
    i32 *hist = A.get<i32>(ncell + 1), *cstart = A.get<i32>(ncell + 1);
    CHKA(A);
    CHK(ulk_memset(hist, 0, (size_t)(ncell + 1) * sizeof(i32), st));
    ulk_launch(Ns, st, K_SrcCell{src->centroid, g, cell, idx, hist});
    CHK(ulk_sort_pairs_u32(cell, cell2, idx, sidx, Ns, st));
    CHK(ulk_exclusive_scan_i32(hist, cstart, ncell + 1, st));
    // ---- k-NN, IoU, top-k
    i64 slots = (i64)Nd * k;
    i32 *cidx = A.get<i32>(slots), *ccount = A.get<i32>(Nd), *nout = A.get<i32>(Nd + 1);
    double *cd2 = A.get<double>(slots), *cw = A.get<double>(slots), *cdist = A.get<double>(slots);
    u8 *cvalid = A.get<u8>(slots);
    i64 *nout64 = A.get<i64>(Nd + 1), *ooff = A.get<i64>(Nd + 1);
    CHKA(A);
    ulk_launch(Nd, st, K_KnnSelect{src->centroid, dst->centroid, shift, g, sidx, cstart, k, cidx, cd2, ccount});
    K_PairIoU piou{src->bbox, src->area, src->mask_off, src->mask, dst->bbox, dst->area,
                   dst->mask_off, dst->mask, shift, k, cidx, cd2, ccount, prm->distance_weight, cw, cdist, slots,
                   prm->weight_kind, prm->edge_must_be_positive, nch, prm->z_score_threshold,
                   prm->src_mean, prm->dst_mean, prm->dst_std, cvalid};
#ifdef ULK_EMU
    ulk_launch(slots, st, piou);
#else
    ulk_launch_warp(slots * 32, st, piou);
#endif
    ulk_launch(Nd, st, K_LinkTopK{k, prm->max_neighbors, ccount, cidx, cw, cdist, nout, cvalid});
    CHK(ulk_memset(nout + Nd, 0, sizeof(i32), st));
    ulk_launch(Nd + 1, st, K_I32ToI64{nout, nout64});
    CHK(ulk_exclusive_scan_i64(nout64, ooff, (i64)Nd + 1, st));
    i64 n_links = 0;
    CHK(ulk_read(&n_links, ooff + Nd, 1, st));
    L->n_links = n_links;
    void *p;
    if (ulk_dev_alloc(&p, (size_t)(n_links > 0 ? n_links : 1) * 4)) return -3;
    L->source = (i32 *)p;
    if (ulk_dev_alloc(&p, (size_t)(n_links > 0 ? n_links : 1) * 4)) return -3;
    L->target = (i32 *)p;
    if (ulk_dev_alloc(&p, (size_t)(n_links > 0 ? n_links : 1) * 8)) return -3;
    L->weight = (double *)p;
    ulk_launch(Nd, st, K_LinkGather{k, nout, ooff, cidx, cw, L->source, L->target, L->weight});
    // pairs evaluated, for the byte model
    {
        i64 *c64 = A.get<i64>(Nd + 1), *cpre = A.get<i64>(Nd + 1);
        CHKA(A);
        ulk_launch(Nd, st, K_I32ToI64{ccount, c64});
        CHK(ulk_memset(c64 + Nd, 0, sizeof(i64), st));
        CHK(ulk_exclusive_scan_i64(c64, cpre, (i64)Nd + 1, st));
        CHK(ulk_read(&L->n_pairs, cpre + Nd, 1, st));
    }
    CHK(ulk_sync(st));
    return 0;
}
