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
