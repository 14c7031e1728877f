// ulk_aux.cu -- the step in front of the path (labels_to_contours) and the device-side
// synthetic frame generator used by bench.py for frames that never exist on the host.
#include "ulk_seg.h"
#include "ulk_common.h"
#include <cmath>

// ------------------------------------------------------------------ labels_to_contours
// ultrack/utils/edge.py:18-96 with sigma=None:
//   foreground |= labels > 0 ; contours += find_boundaries(labels, mode="outer") ; /= n_sets
// skimage.segmentation.find_boundaries(mode="outer") [UPSTREAM-RECALL]:
//   thick  = max(cross nbhd) != min(cross nbhd)
//   adj    = foreground voxel whose full 3^d nbhd holds two different non-zero labels
//   outer  = thick & (background | adj)
// (out-of-frame neighbours are ignored: skimage pads by reflection).
struct K_LabelsToContours {
    const i32 *lab; i32 nsets; UlkDims dm; u8 *fg; float *ct;
    ULK_D void operator()(i64 vv) const
    {
        i32 v = (i32)vv;
        i32 z, y, x;
        dm.coords(v, z, y, x);
        int anyfg = 0;
        float acc = 0.0f;
        for (i32 s = 0; s < nsets; ++s) {
            const i32 *L = lab + (i64)s * dm.V;
            i32 c = L[v];
            if (c > 0) anyfg = 1;
            i32 cmin = c, cmax = c;
            for (int d = 0; d < 3; ++d) {
                i32 st = dm.stride(d);
                if (dm.bwd(d, z, y, x)) { i32 u = L[v - st]; cmin = u < cmin ? u : cmin; cmax = u > cmax ? u : cmax; }
                if (dm.fwd(d, z, y, x)) { i32 u = L[v + st]; cmin = u < cmin ? u : cmin; cmax = u > cmax ? u : cmax; }
            }
            bool thick = cmin != cmax;
            if (!thick) continue;
            bool outer = c == 0;
            if (!outer) {
                i32 fmin = 0x7fffffff, fmax = 0;
                for (i32 dz = -1; dz <= 1; ++dz) {
                    i32 zz = z + dz;
                    if (zz < 0 || zz >= dm.D) continue;
                    for (i32 dy = -1; dy <= 1; ++dy) {
                        i32 yy = y + dy;
                        if (yy < 0 || yy >= dm.H) continue;
                        for (i32 dx = -1; dx <= 1; ++dx) {
                            i32 xx = x + dx;
                            if (xx < 0 || xx >= dm.W) continue;
                            i32 u = L[((i64)zz * dm.H + yy) * dm.W + xx];
                            fmax = u > fmax ? u : fmax;
                            if (u != 0 && u < fmin) fmin = u;
                        }
                    }
                }
                outer = fmax != fmin;
            }
            if (outer) acc += 1.0f;
        }
        fg[v] = (u8)anyfg;
        ct[v] = acc / (float)nsets;
    }
};

int ulk_labels_to_contours_impl(ulk_stream_t st, const int32_t *labels, int32_t n_sets, const int64_t shape[3],
                                uint8_t *foreground, float *contours)
{
    UlkDims dm;
    dm.init(shape[0], shape[1], shape[2]);
    ulk_launch(dm.V, st, K_LabelsToContours{labels, n_sets, dm, foreground, contours});
    return ulk_sync(st);
}

// ------------------------------------------------------------------ contour smoothing
// utils/edge.py:89-91: contours = ndi.gaussian_filter(contours, sigma); contours /= contours.max().
// scipy semantics: per axis (0 first), 1-D correlation with the normalised Gaussian window of
// radius int(4 * sigma + 0.5), boundary mode "reflect" (d c b a | a b c d | d c b a), taps
// accumulated in float64 and every axis pass rounded to the array's float32.
struct K_Gauss1D {
    const float *in; float *out; const double *w; i32 r; UlkDims dm; int axis;   // axis: 0 = x, 1 = y, 2 = z
    ULK_D void operator()(i64 vv) const
    {
        i32 z, y, x;
        dm.coords((i32)vv, z, y, x);
        const i32 n = axis == 0 ? dm.W : (axis == 1 ? dm.H : dm.D);
        const i32 c = axis == 0 ? x : (axis == 1 ? y : z);
        const i64 st = dm.stride(axis);
        const float *base = in + (vv - (i64)c * st);
        auto at = [&](i32 k) {
            // "reflect": period 2n, mirrored about the half-sample beyond each edge
            i32 p = k % (2 * n);
            if (p < 0) p += 2 * n;
            if (p >= n) p = 2 * n - 1 - p;
            return (double)base[(i64)p * st];
        };
        double acc = at(c) * w[r];
        for (i32 j = r; j >= 1; --j) acc += (at(c - j) + at(c + j)) * w[r - j];     // scipy's order: outermost pair first
        out[vv] = (float)acc;
    }
};
struct K_MaxF32 {
    const float *p; u32 *mx;
    ULK_D void operator()(i64 i) const
    {
        const float v = p[i];
        if (v > 0.0f) { u32 b; memcpy(&b, &v, 4); if (b > *mx) ulk_atomic_max(mx, b); }
    }
};
struct K_DivF32 { float *p; float d; ULK_D void operator()(i64 i) const { p[i] = p[i] / d; } };

int ulk_smooth_contours_impl(ulk_stream_t st, float *contours, const int64_t shape[3], const double sigma[3])
{
    UlkDims dm;
    dm.init(shape[0], shape[1], shape[2]);
    UlkArena A;
    float *tmp = A.get<float>(dm.V);
    u32 *mx = A.get<u32>(1);
    if (A.failed) ULK_FAIL(-3, "device allocation failed");
    float *src = contours, *dst = tmp;
    for (int ax = 2; ax >= 0; --ax) {          // scipy filters axis 0 (z) first
        const double sg = sigma[2 - ax];
        const i64 n = ax == 0 ? dm.W : (ax == 1 ? dm.H : dm.D);
        if (!(sg > 0.0) || n <= 1) continue;    // 2-D frames carry a unit z axis
        const i32 r = (i32)(4.0 * sg + 0.5);
        std::vector<double> w((size_t)(2 * r + 1));
        double sum = 0.0;
        for (i32 k = -r; k <= r; ++k) { w[(size_t)(k + r)] = exp(-0.5 / (sg * sg) * (double)k * (double)k); sum += w[(size_t)(k + r)]; }
        for (double &v : w) v /= sum;
        double *dw = A.get<double>(2 * r + 1);
        if (A.failed) ULK_FAIL(-3, "device allocation failed");
        int rc = ulk_h2d(dw, w.data(), w.size() * sizeof(double), st);
        if (rc) return rc;
        ulk_launch(dm.V, st, K_Gauss1D{src, dst, dw, r, dm, ax});
        float *t = src; src = dst; dst = t;
    }
    if (src != contours) { int rc = ulk_d2d(contours, src, (size_t)dm.V * 4, st); if (rc) return rc; }
    int rc = ulk_memset(mx, 0, 4, st);
    if (rc) return rc;
    ulk_launch(dm.V, st, K_MaxF32{contours, mx});
    u32 hb = 0;
    rc = ulk_read(&hb, mx, 1, st);
    if (rc) return rc;
    float m;
    memcpy(&m, &hb, 4);
    // numpy divides by max() whatever it is (0 -> nan); keep that behaviour
    ulk_launch(dm.V, st, K_DivF32{contours, m});
    return ulk_sync(st);
}

// ------------------------------------------------------------------ synthetic frames
ULK_HD ULK_INLINE u64 synth_mix(u64 x, u64 k)
{
    x = (x ^ k) * 0xBF58476D1CE4E5B9ull;
    x ^= x >> 31;
    x *= 0x94D049BB133111EBull;
    x ^= x >> 29;
    return x;
}
ULK_HD ULK_INLINE double synth_u01(u64 seed, u64 a, u64 b, u64 c, int nk)
{
    u64 x = seed * 0x9E3779B97F4A7C15ull;
    x = synth_mix(x, a);
    if (nk > 1) x = synth_mix(x, b);
    if (nk > 2) x = synth_mix(x, c);
    return (double)(x >> 11) * (1.0 / 9007199254740992.0);
}
// ultrack_b200/synth.py make_frame(variant="S"): balls on a jittered lattice; contours low
// in the centre, 1 on the rim and between touching cells, plus a small hashed noise
struct K_SynthFrame {
    UlkDims dm; int ndim; i32 t; u64 seed; double radius, pitch, motion, noise; i32 ncell[3];
    float *fg, *ct;
    ULK_D void operator()(i64 vv) const
    {
        i32 v = (i32)vv;
        i32 c[3];
        dm.coords(v, c[0], c[1], c[2]);
        int a0 = 3 - ndim;
        i32 cell[3] = {0, 0, 0};
        for (int a = a0; a < 3; ++a) cell[a] = (i32)floor((double)c[a] / pitch);
        double s1 = 1e30, s2 = 1e30;
        i32 lo0 = ndim == 3 ? -1 : 0, hi0 = ndim == 3 ? 1 : 0;
        for (i32 o0 = lo0; o0 <= hi0; ++o0)
            for (i32 o1 = -1; o1 <= 1; ++o1)
                for (i32 o2 = -1; o2 <= 1; ++o2) {
                    i32 cc[3] = {cell[0] + o0, cell[1] + o1, cell[2] + o2};
                    i64 cid = 0;
                    for (int a = a0; a < 3; ++a) cid = cid * ncell[a] + (cc[a] + 1);
                    double d2 = 0;
                    for (int a = a0; a < 3; ++a) {
                        int ax = a - a0;
                        double jit = (synth_u01(seed, (u64)cid, (u64)(17 + ax), 0, 2) - 0.5) * 0.5 * pitch;
                        double mov = motion > 0 ? (synth_u01(seed, (u64)cid, (u64)(101 + ax), (u64)t, 3) - 0.5) * 2.0 * motion : 0.0;
                        double ctrp = ((double)cc[a] + 0.5) * pitch + jit + mov;
                        double df = (double)c[a] - ctrp;
                        d2 += df * df;
                    }
                    double u1 = synth_u01(seed, (u64)cid, 7, 0, 2), u2 = synth_u01(seed, (u64)cid, 8, 0, 2);
                    if (u1 < 1e-12) u1 = 1e-12;
                    double g = sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
                    double r = radius * (1.0 + 0.15 * g);
                    if (r < 0.6 * radius) r = 0.6 * radius;
                    if (r > 1.4 * radius) r = 1.4 * radius;
                    double s = sqrt(d2) / r;
                    if (s < s1) { s2 = s1; s1 = s; }
                    else if (s < s2) s2 = s;
                }
        double nz = synth_u01(seed, (u64)v, (u64)(1000 + t), 0, 2) * noise;
        double cv = 1.0 - 3.0 * (s2 - s1);
        if (s1 > cv) cv = s1;
        if (cv < 0) cv = 0;
        if (cv > 1) cv = 1;
        double f = 4.0 * (1.0 - s1);
        if (f < 0) f = 0;
        if (f > 1) f = 1;
        fg[v] = (float)f;
        ct[v] = (float)(cv + nz);
    }
};

int ulk_synth_frame_impl(ulk_stream_t st, const int64_t shape[3], int32_t t, uint64_t seed, double radius,
                         double fill, double motion, double noise, float *fg, float *ct)
{
    K_SynthFrame k;
    k.dm.init(shape[0], shape[1], shape[2]);
    k.ndim = k.dm.D == 1 ? 2 : 3;
    double unit = k.ndim == 2 ? 3.141592653589793 : 4.0 / 3.0 * 3.141592653589793;
    k.pitch = radius * pow(unit / fill, 1.0 / k.ndim);
    k.t = t; k.seed = seed; k.radius = radius; k.motion = motion; k.noise = noise;
    i32 dims[3] = {k.dm.D, k.dm.H, k.dm.W};
    for (int a = 0; a < 3; ++a) k.ncell[a] = (i32)ceil(dims[a] / k.pitch) + 2;
    k.fg = fg; k.ct = ct;
    ulk_launch(k.dm.V, st, k);
    return ulk_sync(st);
}

// ------------------------------------------------------------------ add_flow (imgproc/flow.py:639-751)
// coords = clip(round(centroid * (field_dim - 1) / (image_dim - 1)), 0, field_dim - 1)
// (flow.py:716-720); column j of (x, y, z) takes vector n_vec-1-j (flow.py:711,726-731) times the
// image extent of that axis (flow.py:713,737).  Output rows are (z, y, x).
struct K_NodeShifts {
    const i32 *cen; const void *field; int dt; i32 nvec;
    i64 fd, fh, fw; double cs[3]; double vs[3];      // cs: coordinate scaling z,y,x; vs: extent of x, y, z
    double *out;
    ULK_D void operator()(i64 n) const
    {
        i64 c[3];
        const i64 lim[3] = {fd, fh, fw};
        for (int d = 0; d < 3; ++d) {
            double q = rint((double)cen[3 * n + d] * cs[d]);
            i64 k = (i64)q;
            if (k < 0) k = 0;
            if (k > lim[d] - 1) k = lim[d] - 1;
            c[d] = k;
        }
        const i64 idx = (c[0] * fh + c[1]) * fw + c[2], fv = fd * fh * fw;
        out[3 * n] = out[3 * n + 1] = out[3 * n + 2] = 0.0;
        for (i32 j = 0; j < nvec; ++j)
            out[3 * n + (2 - j)] = ulk_load_as_double(field, dt, (i64)(nvec - 1 - j) * fv + idx) * vs[j];
    }
};
int ulk_node_shifts_impl(ulk_stream_t st, const ulk_frame_nodes *nodes, const void *field, int field_dtype, int32_t n_vec,
                         const int64_t field_shape[3], double *shift_host)
{
    const i32 N = nodes->n_nodes;
    if (N <= 0) return 0;
    UlkArena A;
    double *out = A.get<double>(3 * (i64)N);
    if (A.failed) ULK_FAIL(-3, "device allocation failed");
    K_NodeShifts k;
    k.cen = nodes->centroid; k.field = field; k.dt = field_dtype; k.nvec = n_vec;
    k.fd = field_shape[0]; k.fh = field_shape[1]; k.fw = field_shape[2];
    for (int d = 0; d < 3; ++d) {
        const i64 img = nodes->shape[d];
        k.cs[d] = img > 1 ? (double)(field_shape[d] - 1) / (double)(img - 1) : 0.0;
    }
    k.vs[0] = (double)nodes->shape[2]; k.vs[1] = (double)nodes->shape[1]; k.vs[2] = (double)nodes->shape[0];
    k.out = out;
    ulk_launch(N, st, k);
    return ulk_d2h(shift_host, out, (size_t)N * 3 * sizeof(double), st);
}

// ------------------------------------------------------------------ bulk node blobs (f-2)
// pickle.dumps(Node) of every node of a frame (core/segmentation/processing.py:254-267,
// core/segmentation/node.py:270-278), written by the device: HEAD template + raw bool mask bytes
// expanded from the bit-packed mask + TAIL template, fixed-width holes patched per node
// (ultrack_b200/node.py BlobTemplate builds the templates and documents the stream).
// One warp per node; byte j of a blob is written by lane j % 32.
#ifdef ULK_EMU
#define ULK_SYNCWARP()
#else
#define ULK_SYNCWARP() __syncwarp()
#endif
struct K_NodeBlobs {
    const i32 *bbox, *centroid, *area, *hidx; const double *frontier, *height; const i64 *moff; const u32 *mask;
    const u8 *head, *tail; ulk_blob_spec sp; i64 id_base; i32 time; const i64 *boff; u8 *out; i64 n;
    static ULK_D ULK_INLINE void put_le(u8 *p, u64 v, int nb) { for (int k = 0; k < nb; ++k) p[k] = (u8)(v >> (8 * k)); }
    static ULK_D ULK_INLINE void put_be_double(u8 *p, double d)
    {
        u64 v;
        memcpy(&v, &d, 8);
        for (int k = 0; k < 8; ++k) p[k] = (u8)(v >> (8 * (7 - k)));
    }
    ULK_D ULK_INLINE void write(i64 i, int lane, int step) const
    {
        const int nd = sp.ndim, lo = 3 - nd;
        const i32 *bb = bbox + 6 * i;
        const i32 d = bb[3] - bb[0], h = bb[4] - bb[1], w = bb[5] - bb[2];
        const i64 vol = (i64)d * h * w;
        u8 *o = out + boff[i];
        for (i32 j = lane; j < sp.head_len; j += step) o[j] = head[j];
        u8 *om = o + sp.head_len;
        const u32 *m = mask + moff[i];
        const i32 wpr = (w + 31) / 32;
        const i64 rows = (i64)d * h;
        // lanes walk the dense [d*h, w] byte image; consecutive lanes write consecutive bytes
        for (i64 j = lane; j < vol; j += step) {
            const i64 r = j / w;
            const i32 x = (i32)(j - r * w);
            om[j] = (u8)((m[r * wpr + (x >> 5)] >> (x & 31)) & 1u);
        }
        (void)rows;
        u8 *ot = om + vol;
        for (i32 j = lane; j < sp.tail_len; j += step) ot[j] = tail[j];
        ULK_SYNCWARP();
        if (lane != 0) return;
        put_le(o + sp.off_area, (u64)(u32)area[i], 4);
        put_be_double(o + sp.off_frontier, frontier ? frontier[i] : -1.0);
        put_be_double(o + sp.off_height, height ? height[i] : -1.0);
        put_le(o + sp.off_hidx, (u64)(u32)(hidx ? hidx[i] : -1), 4);
        for (int k = 0; k < nd; ++k) {
            put_le(o + sp.off_bbox + 8 * k, (u64)(i64)bb[lo + k], 8);
            put_le(o + sp.off_bbox + 8 * (nd + k), (u64)(i64)bb[3 + lo + k], 8);
            put_le(o + sp.off_shape + 5 * k, (u64)(u32)(bb[3 + lo + k] - bb[lo + k]), 4);
            put_le(ot + sp.toff_centroid + 8 * k, (u64)(i64)centroid[3 * i + lo + k], 8);
        }
        const u64 payload = (u64)(sp.head_len - sp.payload_start) + (u64)vol + (u64)sp.payload_tail;
        put_le(o + sp.off_chunk_len, payload + 16, 4);
        put_le(o + sp.off_blosc_nbytes, payload, 4);
        put_le(o + sp.off_blosc_blocksize, payload, 4);
        put_le(o + sp.off_blosc_cbytes, payload + 16, 4);
        put_le(o + sp.off_raw_len, (u64)vol, 4);
        put_le(ot + sp.toff_id, (u64)(id_base + i + 1), 8);
        put_le(ot + sp.toff_time, (u64)(u32)time, 4);
    }
#ifndef ULK_EMU
    ULK_D void warp(i64 ii, bool) const
    {
        const i64 i = ii >> 5;
        if (i < n) write(i, threadIdx.x & 31, 32);
    }
#endif
    ULK_D void operator()(i64 i) const { write(i, 0, 1); }
};

int ulk_frame_blobs_impl(ulk_stream_t st, const ulk_frame_nodes *f, const uint8_t *head, const uint8_t *tail,
                         const ulk_blob_spec *sp, int64_t id_base, int32_t time, const int64_t *blob_off_host,
                         uint8_t *out_host)
{
    const i64 n = f->n_nodes;
    if (n <= 0) return 0;
    const i64 total = blob_off_host[n];
    for (i64 i = 0; i < n; ++i)
        if (blob_off_host[i + 1] - blob_off_host[i] >= 2147483000ll) ULK_FAIL(-1, "node %lld: mask too large for one Blosc chunk", (long long)i);
    UlkArena A;
    u8 *dh = A.get<u8>(sp->head_len), *dt = A.get<u8>(sp->tail_len), *dout = A.get<u8>(total);
    i64 *doff = A.get<i64>(n + 1);
    if (A.failed) ULK_FAIL(-3, "device allocation failed");
    int rc = 0;
    rc |= ulk_h2d(dh, head, (size_t)sp->head_len, st);
    rc |= ulk_h2d(dt, tail, (size_t)sp->tail_len, st);
    rc |= ulk_h2d(doff, blob_off_host, (size_t)(n + 1) * 8, st);
    if (rc) return rc;
    K_NodeBlobs k{f->bbox, f->centroid, f->area, f->hidx, f->frontier, f->height, f->mask_off, f->mask,
                  dh, dt, *sp, id_base, time, doff, dout, n};
#ifdef ULK_EMU
    ulk_launch(n, st, k);
#else
    ulk_launch_warp(n * 32, st, k);
#endif
    return ulk_d2h(out_host, dout, (size_t)total, st);
}
