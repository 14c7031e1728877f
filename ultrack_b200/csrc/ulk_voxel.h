// ulk_voxel.h -- voxel-scale kernels of the segmentation path (functor bodies).
//
// Every kernel is a functor launched one thread per voxel (or per list entry)
// through ulk_launch<>.  Vocabulary follows ultrack / higra: a *region* is one
// hierarchy's support (core/segmentation/hierarchy.py:89-90), a *basin* is a
// catchment basin of the watershed cut of a region (altitude-0 node of
// hg.watershed_hierarchy_by_*), the *RAG* is the basin adjacency graph.
#pragma once
#include "ulk_common.h"

#define ULK_JK_CORE 0xFFFFFFFEu   // join key of voxels inside the regional minimum

// ------------------------------------------------------------------ union-find on voxels
ULK_D ULK_INLINE i32 ulk_vfind(const i32 *P, i32 x)
{
    i32 p = P[x];
    while (p != x) { x = p; p = P[x]; }
    return x;
}
// find with path halving: every store replaces a parent by one of its ancestors (parents only
// ever decrease), so plain stores are safe next to the atomicMin hooks of ulk_vunion
ULK_D ULK_INLINE i32 ulk_vfind_halve(i32 *P, i32 x)
{
    for (;;) {
        i32 p = P[x];
        if (p == x) return x;
        i32 g = P[p];
        if (g == p) return p;
        P[x] = g;
        x = g;
    }
}
ULK_D ULK_INLINE void ulk_vunion(i32 *P, i32 a, i32 b)
{
    for (;;) {
        a = ulk_vfind_halve(P, a);
        b = ulk_vfind_halve(P, b);
        if (a == b) return;
        if (a < b) { i32 t = a; a = b; b = t; }
        i32 old = ulk_atomic_min(&P[a], b);   // hook the larger root under the smaller
        if (old == a) return;
        a = old;
    }
}

// ------------------------------------------------------------------ K1: threshold + contour staging
// processing.py:212 `foreground[time] > threshold`; vendored/hierarchy.py:158-159,350 up-casts.
// The component forest starts with every x-run of foreground voxels inside a 32-voxel warp
// segment already hooked to its first voxel, so K_CclUnion only has to join runs.
#ifndef ULK_EMU
// first lane of the chain of `linked` lanes that ends at `lane` (bit l of linked: lane l is
// joined to lane l-1)
ULK_D ULK_INLINE int ulk_run_start(u32 linked, int lane)
{
    u32 stop = ~linked & (0xffffffffu >> (31 - lane));   // lanes <= lane that start a run
    return 31 - __clz(stop);                               // lane 0 never has its link bit set
}
#endif
struct K_Threshold {
    const void *fg; int fgt; double thr;
    const void *ct; int ctt;
    float *ctf; double *ctd;
    i32 *P; u32 *cmax;
    ULK_D ULK_INLINE bool stage(i64 v, u32 &bits) const
    {
        double f = ulk_load_as_double(fg, fgt, v);
        bits = 0;
        if (ctt == ULK_F32) {       // read in place, never staged
            float a = fabsf(((const float *)ct)[v]);
            if (a <= 3.0e38f) memcpy(&bits, &a, 4);
            return f > thr;
        }
        double c = ulk_load_as_double(ct, ctt, v);
        if (ctf) ctf[v] = (float)c; else if (ctd) ctd[v] = c;
        float a = (float)(c < 0 ? -c : c);
        if (a == a && a <= 3.0e38f) memcpy(&bits, &a, 4);
        return f > thr;
    }
#ifdef ULK_EMU
    void operator()(i64 v, i32, i32, i32 x) const
    {
        u32 bits;
        bool on = stage(v, bits);
        if (bits > *cmax) *cmax = bits;
        bool link = on && (x & 31) != 0 && P[v - 1] >= 0;
        P[v] = on ? (link ? P[v - 1] : (i32)v) : -1;
    }
#else
    ULK_D void warp(i64 v, i32, i32, i32, bool valid) const
    {
        const int lane = threadIdx.x & 31;     // == x & 31
        u32 bits = 0;
        bool on = valid && stage(v, bits);
        u32 m = __reduce_max_sync(0xffffffffu, bits);
        if (lane == 0 && m > *cmax) atomicMax(cmax, m);
        bool prev = __shfl_up_sync(0xffffffffu, on ? 1 : 0, 1) != 0;
        u32 linked = __ballot_sync(0xffffffffu, on && prev && lane > 0);
        if (valid) P[v] = on ? (i32)v - (lane - ulk_run_start(linked, lane)) : -1;
    }
#endif
};

// ------------------------------------------------------------------ connected components (ndi.label)
// joins the runs: an x edge only across a warp-segment boundary, a y / z edge only where the
// edge one voxel to the left does not already join the same two runs
struct K_CclUnion {
    i32 *P; UlkDims dm; i32 *cnt;
    ULK_D void operator()(i64 vv, i32 z, i32 y, i32 x) const
    {
        const i32 v = (i32)vv;
        // every neighbour is loaded up front (clamped to v at the frame border) so the loads
        // overlap; only the sign is used, and the sign of an entry never changes
        const bool hx = x > 0, hy = y > 0, hz = z > 0;
        const i32 uy = hy ? v - dm.W : v, uz = hz ? v - dm.W * dm.H : v;
        const i32 p0 = P[v];
        const i32 pl = P[hx ? v - 1 : v], py = P[uy], pyl = P[hx ? uy - 1 : uy], pz = P[uz], pzl = P[hx ? uz - 1 : uz];
        if (p0 < 0) return;
        if (p0 == v) cnt[v] = 0;      // a final root is never hooked: its size counter starts at 0
        const bool left = hx && pl >= 0;
        if (left && (x & 31) == 0) ulk_vunion(P, v, v - 1);
        if (hy && py >= 0 && !(left && pyl >= 0)) ulk_vunion(P, v, uy);
        if (hz && pz >= 0 && !(left && pzl >= 0)) ulk_vunion(P, v, uz);
    }
};
struct K_Flatten {
    i32 *P;
    ULK_D void operator()(i64 v) const { if (P[v] >= 0) P[v] = ulk_vfind(P, (i32)v); }
};
// flatten, component sizes (one atomic per group of lanes with the same root) and root list
struct K_FlattenCount {
    i32 *P; i32 *cnt; i32 *roots; i32 *nroots;
#ifdef ULK_EMU
    void operator()(i64 v, i32, i32, i32) const
    {
        if (P[v] < 0) return;
        i32 r = ulk_vfind(P, (i32)v);
        P[v] = r;
        cnt[r] += 1;
        if (r == (i32)v) roots[(*nroots)++] = r;
    }
#else
    ULK_D void warp(i64 v, i32, i32, i32, bool valid) const
    {
        i32 r = -1;
        if (valid && P[v] >= 0) { r = ulk_vfind(P, (i32)v); P[v] = r; }
        u32 grp = __match_any_sync(0xffffffffu, r);
        if (r < 0) return;
        if ((threadIdx.x & 31) == __ffs(grp) - 1) atomicAdd(&cnt[r], __popc(grp));
        if (r == (i32)v) roots[atomicAdd(nroots, 1)] = r;
    }
#endif
};
// per component, over the root list in raster order: morphology.remove_small_objects drops
// components with count <= max_size (only when there is more than one component,
// hierarchy.py:64-73); `big` marks kept components above max_area
struct K_RootFlags {
    const u32 *roots; const i32 *cnt; i32 max_size; int do_remove; i32 max_area;
    i32 *keep; i32 *big; i32 *size;
    ULK_D void operator()(i64 i) const
    {
        i32 s = cnt[roots[i]];
        size[i] = s;
        i32 k = !(do_remove && s <= max_size);
        keep[i] = k;
        big[i] = k && s > max_area;
    }
};
// code[root] = 1-based index among the kept components (0: dropped); cc_big[index] = index
// among the components above max_area or -1; big_cc = the inverse of cc_big
struct K_RootCode {
    const u32 *roots; const i32 *keep, *keeppre, *big, *bigpre; i32 *code; i32 *cc_big; i32 *big_cc;
    ULK_D void operator()(i64 i) const
    {
        u32 r = roots[i];
        if (!keep[i]) { code[r] = 0; return; }
        i32 c = keeppre[i];
        code[r] = c + 1;
        cc_big[c] = big[i] ? bigpre[i] : -1;
        if (big[i]) big_cc[bigpre[i]] = c;
    }
};
// Component labels, written in place over the flattened component forest (the background keeps
// its -1).  mode 0 (no component above max_area): the 1-based component index, 0 if dropped.
// mode 1: components above max_area get their 1-based index among themselves (the region labels
// of the oversegmentation pass), the other kept components -(index + 2), dropped ones -1.
struct K_LabelCc {
    const i32 *code; const i32 *cc_big; int mode; i32 *lab;
    ULK_D void operator()(i64 v, i32, i32, i32) const
    {
        i32 c = code[lab[v]];
        if (mode == 0) { lab[v] = c; return; }
        if (c == 0) { lab[v] = -1; return; }
        i32 g = cc_big[c - 1];
        lab[v] = g >= 0 ? g + 1 : -(c - 1) - 2;
    }
};
struct K_FlagRoots {
    const i32 *P; i32 *flag;
    ULK_D void operator()(i64 v) const { flag[v] = P[v] == (i32)v ? 1 : 0; }
};

// ------------------------------------------------------------------ voxel stage: steepest edges
// For every active voxel the minimum incident edge by (weight, edge key); edge key =
// 3*source + direction reproduces the enumeration order of edges_from_mask
// (vendored/graph.py:41-57), i.e. higra's stable tie-break.
// Candidates are visited in increasing edge-key order (-z, -y, -x, +x, +y, +z), so a strict
// comparison of the weights alone yields the minimum of (weight, key).
struct K_MinEdge {
    const i32 *lab; UlkCt ct; UlkDims dm;
    u32 *mw, *pk;
    ULK_D void operator()(i64 vv, i32 z, i32 y, i32 x) const
    {
        const i32 v = (i32)vv;
        const i32 sy = dm.W, sz = dm.W * dm.H;
        // neighbour indices clamped to v at the frame border: all loads are issued together
        const bool h0 = z > 0, h1 = y > 0, h2 = x > 0, h3 = x + 1 < dm.W, h4 = y + 1 < dm.H, h5 = z + 1 < dm.D;
        const i32 u0 = h0 ? v - sz : v, u1 = h1 ? v - sy : v, u2 = h2 ? v - 1 : v;
        const i32 u3 = h3 ? v + 1 : v, u4 = h4 ? v + sy : v, u5 = h5 ? v + sz : v;
        const i32 l = lab[v];
        const i32 l0 = lab[u0], l1 = lab[u1], l2 = lab[u2], l3 = lab[u3], l4 = lab[u4], l5 = lab[u5];
        if (l <= 0) return;          // mw / pk are only ever read where lab > 0
        const bool s0 = h0 && l0 == l, s1 = h1 && l1 == l, s2 = h2 && l2 == l;
        const bool s3 = h3 && l3 == l, s4 = h4 && l4 == l, s5 = h5 && l5 == l;
        u32 bw = ULK_NONE32, bk = ULK_NONE32, w;
        if (ct.f) {
            const float c = ct.f[v];
            const float c0 = ct.f[u0], c1 = ct.f[u1], c2 = ct.f[u2], c3 = ct.f[u3], c4 = ct.f[u4], c5 = ct.f[u5];
            if (s0) { w = ulk_f2ord(ulk_wsum(c0 + c)); if (w < bw) { bw = w; bk = 3u * (u32)u0 + 2u; } }
            if (s1) { w = ulk_f2ord(ulk_wsum(c1 + c)); if (w < bw) { bw = w; bk = 3u * (u32)u1 + 1u; } }
            if (s2) { w = ulk_f2ord(ulk_wsum(c2 + c)); if (w < bw) { bw = w; bk = 3u * (u32)u2; } }
            if (s3) { w = ulk_f2ord(ulk_wsum(c + c3)); if (w < bw) { bw = w; bk = 3u * (u32)v; } }
            if (s4) { w = ulk_f2ord(ulk_wsum(c + c4)); if (w < bw) { bw = w; bk = 3u * (u32)v + 1u; } }
            if (s5) { w = ulk_f2ord(ulk_wsum(c + c5)); if (w < bw) { bw = w; bk = 3u * (u32)v + 2u; } }
        } else {
            if (s0) { w = ulk_f2ord(ct.w(u0, v)); if (w < bw) { bw = w; bk = 3u * (u32)u0 + 2u; } }
            if (s1) { w = ulk_f2ord(ct.w(u1, v)); if (w < bw) { bw = w; bk = 3u * (u32)u1 + 1u; } }
            if (s2) { w = ulk_f2ord(ct.w(u2, v)); if (w < bw) { bw = w; bk = 3u * (u32)u2; } }
            if (s3) { w = ulk_f2ord(ct.w(v, u3)); if (w < bw) { bw = w; bk = 3u * (u32)v; } }
            if (s4) { w = ulk_f2ord(ct.w(v, u4)); if (w < bw) { bw = w; bk = 3u * (u32)v + 1u; } }
            if (s5) { w = ulk_f2ord(ct.w(v, u5)); if (w < bw) { bw = w; bk = 3u * (u32)v + 2u; } }
        }
        mw[v] = bw;
        pk[v] = bk;
    }
};
ULK_HD ULK_INLINE i32 ulk_edge_other(u32 key, i32 v, const UlkDims &dm)
{
    i32 a = (i32)(key / 3u);
    i32 b = a + dm.stride((int)(key % 3u));
    return a == v ? b : a;
}
// first Boruvka round: every voxel hooks along its steepest edge; mutual pairs keep the
// smaller voxel as root.  state: 0 growing, 1 minimum (basin seed)
struct K_Hook1 {
    const i32 *lab; const u32 *mw, *pk; UlkDims dm;
    i32 *bp; u8 *state; u32 *Lr; u32 *best; i32 *nbasins;
    ULK_D void operator()(i64 vv, i32, i32, i32) const
    {
        i32 v = (i32)vv;
        state[v] = 0;
        if (lab[v] <= 0) { bp[v] = -1; return; }
        u32 k = pk[v];
        if (k == ULK_NONE32) {
            bp[v] = v; state[v] = 1; Lr[v] = ULK_NONE32; best[v] = ULK_NONE32;
            ulk_atomic_add(nbasins, 1);
            return;
        }
        i32 u = ulk_edge_other(k, v, dm);
        if (pk[u] == k && v < u) { bp[v] = v; Lr[v] = mw[v]; best[v] = ULK_NONE32; }
        else bp[v] = u;
    }
};
struct K_Compress {
    i32 *bp;
    ULK_D void operator()(i64 v) const { if (bp[v] >= 0) bp[v] = ulk_vfind(bp, (i32)v); }
};
#ifndef ULK_EMU
// appends `value` of every lane with `flag` to a list: one atomic per warp
ULK_D ULK_INLINE void ulk_warp_append(bool flag, i32 value, i32 *list, i32 *counter)
{
    u32 m = __ballot_sync(0xffffffffu, flag);
    if (!m) return;
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    i32 base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (flag) list[base + __popc(m & ((1u << lane) - 1u))] = value;
}
#endif
// path compression of the hook forest, fused with the lists of the flow rounds: growing
// roots, and candidates = voxels whose steepest weight equals their supervertex level
struct K_CompressCand {
    i32 *bp; const u32 *mw; const u32 *Lr; const u8 *state;
    i32 *cand; i32 *ncand; i32 *groots; i32 *ngroots;
#ifdef ULK_EMU
    void operator()(i64 vv, i32, i32, i32) const
    {
        i32 v = (i32)vv;
        if (bp[v] < 0) return;
        i32 r = ulk_vfind(bp, v);
        bp[v] = r;
        if (r == v && state[v] == 0) groots[(*ngroots)++] = v;
        if (state[r] == 0 && mw[v] == Lr[r]) cand[(*ncand)++] = v;
    }
#else
    ULK_D void warp(i64 vv, i32, i32, i32, bool valid) const
    {
        i32 v = (i32)vv;
        bool is_groot = false, is_cand = false;
        if (valid && bp[v] >= 0) {
            i32 r = ulk_vfind(bp, v);
            bp[v] = r;
            bool growing = state[r] == 0;
            is_groot = r == v && growing;
            is_cand = growing && mw[v] == Lr[r];
        }
        ulk_warp_append(is_groot, v, groots, ngroots);
        ulk_warp_append(is_cand, v, cand, ncand);
    }
#endif
};
// a growing supervertex looks for its minimum-key outgoing edge at its own level
struct K_FindFlow {
    const i32 *cand; const i32 *lab; const i32 *bp; const u32 *mw; const u32 *Lr; const u8 *state;
    UlkCt ct; UlkDims dm;
    u32 *best; i32 *cand2; i32 *ncand2;
    ULK_D void operator()(i64 i) const
    {
        i32 v = cand[i];
        i32 r = ulk_vfind(bp, v);
        if (state[r] != 0 || mw[v] != Lr[r]) return;
        cand2[ulk_atomic_add(ncand2, 1)] = v;
        i32 l = lab[v];
        u32 lv = Lr[r];
        i32 z, y, x;
        dm.coords(v, z, y, x);
        u32 b = ULK_NONE32;
        for (int d = 0; d < 3; ++d) {
            i32 st = dm.stride(d);
            if (dm.bwd(d, z, y, x)) {
                i32 u = v - st;
                if (lab[u] == l && ulk_f2ord(ct.w(u, v)) == lv && ulk_vfind(bp, u) != r) {
                    u32 k = 3u * (u32)u + d;
                    if (k < b) b = k;
                }
            }
            if (dm.fwd(d, z, y, x)) {
                i32 u = v + st;
                if (lab[u] == l && ulk_f2ord(ct.w(v, u)) == lv && ulk_vfind(bp, u) != r) {
                    u32 k = 3u * (u32)v + d;
                    if (k < b) b = k;
                }
            }
        }
        if (b != ULK_NONE32) ulk_atomic_min(&best[r], b);
    }
};
// decide, read-only part: target root of every growing supervertex (-1: becomes a minimum,
// -2: stays root as the winner of a mutual pair)
struct K_Decide1 {
    const i32 *groots; const i32 *bp; const u32 *best; UlkDims dm;
    i32 *tgt;
    ULK_D void operator()(i64 i) const
    {
        i32 r = groots[i];
        u32 k = best[r];
        if (k == ULK_NONE32) { tgt[i] = -1; return; }
        i32 a = (i32)(k / 3u), b = a + dm.stride((int)(k % 3u));
        i32 ra = ulk_vfind(bp, a), rb = ulk_vfind(bp, b);
        i32 t = ra == r ? rb : ra;
        if (best[t] == k && r < t) { tgt[i] = -2; return; }
        tgt[i] = t;
    }
};
struct K_Decide2 {
    const i32 *groots; const i32 *tgt;
    i32 *bp; u8 *state; u32 *best; u8 *fmark;
    i32 *groots2; i32 *ngroots2;
    i32 *flow_root; u32 *flow_key; i32 *nflow; i32 *nbasins;
    ULK_D void operator()(i64 i) const
    {
        i32 r = groots[i];
        i32 t = tgt[i];
        u32 k = best[r];
        best[r] = ULK_NONE32;
        if (t == -1) { state[r] = 1; ulk_atomic_add(nbasins, 1); return; }
        if (t == -2) { groots2[ulk_atomic_add(ngroots2, 1)] = r; return; }
        bp[r] = t;
        i32 j = ulk_atomic_add(nflow, 1);
        flow_root[j] = r; flow_key[j] = k;
        u32 a = k / 3u, d = k % 3u;
        ulk_atomic_or((u32 *)(fmark + (a & ~3u)), (1u << d) << (8u * (a & 3u)));
    }
};

// ------------------------------------------------------------------ basins
// second path compression; every basin root appends (region label, root voxel): sorting
// these keys numbers the basins grouped by region, in raster order of their roots
struct K_CompressRoots {
    i32 *bp; const i32 *lab; u64 *rootkeys; i32 *nroots;
    ULK_D void operator()(i64 vv, i32, i32, i32) const
    {
        i32 v = (i32)vv;
        if (bp[v] < 0) return;
        i32 r = ulk_vfind(bp, v);
        bp[v] = r;
        if (r == v) rootkeys[ulk_atomic_add(nroots, 1)] = ((u64)(u32)lab[v] << 32) | (u32)v;
    }
};
struct K_BasinRegroup {      // sorted position = global basin id
    const u64 *sorted; const u32 *Lr;
    u32 *gid_at_root; i32 *g_root; i32 *g_lab; u32 *g_L; i32 *lab_hist;
    ULK_D void operator()(i64 g) const
    {
        u64 key = sorted[g];
        i32 v = (i32)(u32)key, l = (i32)(key >> 32);
        gid_at_root[v] = (u32)g;
        g_root[g] = v;
        g_lab[g] = l;
        g_L[g] = Lr[v];
        ulk_atomic_add(&lab_hist[l - 1], 1);
    }
};
struct K_AssignBasin {     // background stays at the memset value -1
    const i32 *bp; const u32 *gid_at_root;
    i32 *bas;
    ULK_D void operator()(i64 v, i32, i32, i32) const { i32 r = bp[v]; bas[v] = r < 0 ? -1 : (i32)gid_at_root[r]; }
};
// flow edges are spanning-tree edges of the basin of their supervertex
struct K_FlowTop {
    const i32 *flow_root; const u32 *flow_key; const i32 *bas; UlkCt ct; UlkDims dm; u64 *b_top;
    ULK_D void operator()(i64 j) const
    {
        u32 fk = flow_key[j];
        i32 a = (i32)(fk / 3u), e = a + dm.stride((int)(fk % 3u));
        ulk_atomic_max(&b_top[bas[flow_root[j]]], ((u64)ulk_f2ord(ct.w(a, e)) << 32) | fk);
    }
};
// One pass over the voxels once basin ids exist:
//  * per-basin size, largest spanning-tree key, bbox / coordinate sums, and the number of
//    voxel edges between different basins of a region (sizes the adjacency hash table);
//  * join keys for the area attribute: J(v) = (mw[v], jk[v]) is the Kruskal position at which
//    voxel v joins a component that holds a regional minimum; core voxels (mw == basin level)
//    belong to the minimum, plateau voxels above it go to the relaxation list `hp`;
//  * K0[basin] = max over its voxels of their smallest in-basin edge key, the lower bound of
//    the basin's raster-key bottleneck (see K_K3*).
struct K_BasinAccum {
    const i32 *bas; const i32 *lab; const u32 *mw, *pk; UlkDims dm;
    i32 *b_size; u64 *b_top; u64 *ncross;
    int want_stats; i32 *bbmin, *bbmax; i64 *csum;
    int want_join; const u32 *b_L; u32 *jk; i32 *hp; i32 *nhp;
    int want_k0; u32 *K0;
    struct Own { u64 top; u32 nc; u32 kmin; i32 c[3]; bool plateau; };
    ULK_D ULK_INLINE void own(i32 v, i32 z, i32 y, i32 x, i32 b, Own &o) const
    {
        // neighbour basins first, all loads in flight together (indices clamped at the border)
        const i32 sy = dm.W, sz = dm.W * dm.H;
        const bool f0 = x + 1 < dm.W, f1 = y + 1 < dm.H, f2 = z + 1 < dm.D;
        const i32 a0 = bas[f0 ? v + 1 : v], a1 = bas[f1 ? v + sy : v], a2 = bas[f2 ? v + sz : v];
        i32 c0 = -1, c1 = -1, c2 = -1;
        if (want_k0) { c0 = bas[x > 0 ? v - 1 : v]; c1 = bas[y > 0 ? v - sy : v]; c2 = bas[z > 0 ? v - sz : v]; }
        const u32 m = mw[v], k = pk[v];
        const i32 l = lab[v];
        o.top = k != ULK_NONE32 ? ((u64)m << 32) | k : 0;
        o.c[0] = z; o.c[1] = y; o.c[2] = x;
        o.nc = 0;
        o.kmin = ULK_NONE32;
        if (f0) { if (a0 == b) { u32 kk = 3u * (u32)v; if (kk < o.kmin) o.kmin = kk; } else if (lab[v + 1] == l) ++o.nc; }
        if (want_k0 && x > 0 && c0 == b) { u32 kk = 3u * (u32)(v - 1); if (kk < o.kmin) o.kmin = kk; }
        if (f1) { if (a1 == b) { u32 kk = 3u * (u32)v + 1u; if (kk < o.kmin) o.kmin = kk; } else if (lab[v + sy] == l) ++o.nc; }
        if (want_k0 && y > 0 && c1 == b) { u32 kk = 3u * (u32)(v - sy) + 1u; if (kk < o.kmin) o.kmin = kk; }
        if (f2) { if (a2 == b) { u32 kk = 3u * (u32)v + 2u; if (kk < o.kmin) o.kmin = kk; } else if (lab[v + sz] == l) ++o.nc; }
        if (want_k0 && z > 0 && c2 == b) { u32 kk = 3u * (u32)(v - sz) + 2u; if (kk < o.kmin) o.kmin = kk; }
        o.plateau = false;
        if (want_join) {
            if (m == b_L[b] || k == ULK_NONE32) jk[v] = ULK_JK_CORE;
            else if (mw[ulk_edge_other(k, v, dm)] < m) jk[v] = k;
            else { jk[v] = ULK_NONE32; o.plateau = true; }
        }
    }
#ifdef ULK_EMU
    void operator()(i64 vv, i32 z, i32 y, i32 x) const
    {
        i32 v = (i32)vv;
        i32 b = bas[v];
        if (b < 0) return;
        Own o;
        own(v, z, y, x, b, o);
        b_size[b] += 1;
        if (o.top > b_top[b]) b_top[b] = o.top;
        *ncross += o.nc;
        if (o.plateau) hp[(*nhp)++] = v;
        if (want_k0 && o.kmin != ULK_NONE32 && o.kmin > K0[b]) K0[b] = o.kmin;
        if (want_stats)
            for (int d = 0; d < 3; ++d) {
                ulk_atomic_min(&bbmin[3 * b + d], o.c[d]);
                ulk_atomic_max(&bbmax[3 * b + d], o.c[d]);
                csum[3 * b + d] += o.c[d];
            }
    }
#else
    // lanes of a warp that sit in the same basin AND the same row form a group (usually one or
    // two per warp): z and y are uniform in it and x grows with the lane, so the bounding box and
    // the coordinate sums need no reduction; one lane per group issues the atomics, and only
    // those that can still change the accumulator (a stale cached value only costs an atomic)
    ULK_D void warp(i64 vv, i32 z, i32 y, i32 x, bool valid) const
    {
        const int lane = threadIdx.x & 31;
        i32 v = (i32)vv;
        i32 b = valid ? bas[v] : -1;
        Own o;
        o.top = 0; o.nc = 0; o.kmin = ULK_NONE32; o.c[0] = o.c[1] = o.c[2] = 0; o.plateau = false;
        if (b >= 0) own(v, z, y, x, b, o);
        u32 nct = __reduce_add_sync(0xffffffffu, o.nc);
        if (lane == 0 && nct) atomicAdd((unsigned long long *)ncross, (unsigned long long)nct);
        if (want_join) ulk_warp_append(o.plateau, v, hp, nhp);
        u32 grp = __match_any_sync(0xffffffffu, b < 0 ? ULK_NONE64 : (((u64)(u32)b << 32) | (u32)(v - x)));
        if (b < 0) return;
        const int first = __ffs(grp) - 1;
        const bool lead = lane == first;
        const i32 n = __popc(grp);
        u32 thi = __reduce_max_sync(grp, (u32)(o.top >> 32));
        u32 tlo = __reduce_max_sync(grp, (u32)(o.top >> 32) == thi ? (u32)o.top : 0u);
        if (lead) {
            atomicAdd(&b_size[b], n);
            u64 t = ((u64)thi << 32) | tlo;
            if (t > b_top[b]) atomicMax((unsigned long long *)&b_top[b], (unsigned long long)t);
        }
        if (want_k0) {
            u32 km = __reduce_max_sync(grp, o.kmin == ULK_NONE32 ? 0u : o.kmin);
            if (lead && km > K0[b]) atomicMax(&K0[b], km);
        }
        if (want_stats) {
            const i32 x1 = __shfl_sync(grp, x, 31 - __clz(grp));       // x of the group's last lane
            u32 sx = (x1 - x + 1 == n) ? (u32)(((i64)(x + x1) * n) >> 1) : 0u;   // meaningful in the first lane
            if (__shfl_sync(grp, x1 - x + 1 != n ? 1 : 0, first)) sx = __reduce_add_sync(grp, (u32)x);
            if (lead) {
                i32 *lo = bbmin + 3 * b, *hi = bbmax + 3 * b;
                i64 *cs = csum + 3 * b;
                if (z < lo[0]) atomicMin(&lo[0], z);
                if (z > hi[0]) atomicMax(&hi[0], z);
                if (y < lo[1]) atomicMin(&lo[1], y);
                if (y > hi[1]) atomicMax(&hi[1], y);
                if (x < lo[2]) atomicMin(&lo[2], x);
                if (x1 > hi[2]) atomicMax(&hi[2], x1);
                atomicAdd((unsigned long long *)&cs[0], (unsigned long long)((i64)z * n));
                atomicAdd((unsigned long long *)&cs[1], (unsigned long long)((i64)y * n));
                atomicAdd((unsigned long long *)&cs[2], (unsigned long long)sx);
            }
        }
    }
#endif
};

// ------------------------------------------------------------------ join keys (area hierarchy)
// plateau voxels above the basin level: minimax relaxation along spanning-tree edges
struct K_PlateauRelax {
    const i32 *hp; const i32 *lab; const u32 *mw, *pk; const u8 *fmark; UlkCt ct; UlkDims dm;
    u32 *jk; i32 *changed;
    ULK_D void operator()(i64 i) const
    {
        i32 v = hp[i];
        i32 l = lab[v];
        u32 lv = mw[v];
        u32 cur = jk[v];
        i32 z, y, x;
        dm.coords(v, z, y, x);
        u32 bestc = cur;
        for (int d = 0; d < 3; ++d) {
            i32 st = dm.stride(d);
            for (int s = 0; s < 2; ++s) {
                i32 u, src;
                if (s == 0) { if (!dm.bwd(d, z, y, x)) continue; u = v - st; src = u; }
                else { if (!dm.fwd(d, z, y, x)) continue; u = v + st; src = v; }
                if (lab[u] != l) continue;
                u32 k = 3u * (u32)src + d;
                bool tree = pk[v] == k || pk[u] == k || ((fmark[src] >> d) & 1);
                if (!tree) continue;
                if (ulk_f2ord(ct.w(src, src + st)) != lv) continue;
                u32 c;
                if (mw[u] < lv) c = k;
                else { u32 ju = jk[u]; c = ju > k ? ju : k; }
                if (c < bestc) bestc = c;
            }
        }
        if (bestc < cur) { jk[v] = bestc; *changed = 1; }
    }
};
// late joiners climb the basin BPT: a voxel attaches to the highest ancestor of its basin
// whose merge precedes its join key
struct K_Attach {
    const i32 *bas; const u32 *mw, *jk; const i32 *b_lab; const i32 *bs;
    const i32 *Ipar; const u64 *Iwk; i32 *attach;
    ULK_D void operator()(i64 v, i32, i32, i32) const
    {
        i32 b = bas[v];
        if (b < 0) return;
        u32 j = jk[v];
        if (j == ULK_JK_CORE) return;
        u64 J = ((u64)mw[v] << 32) | j;
        i32 h = b_lab[b] - 1;
        i32 s = bs[h];
        i32 o = 2 * s;
        i32 node = b - s;
        i32 start = node;
        // the parent's merge key and the grandparent index are loaded together: one memory
        // round trip per level of the climb
        i32 p = Ipar[o + node];
        while (p != node) {
            u64 w = Iwk[o + p];
            i32 pp = Ipar[o + p];
            if (!(w < J)) break;
            node = p;
            p = pp;
        }
        if (node != start) {
            ulk_atomic_add(&attach[o + start], -1);
            ulk_atomic_add(&attach[o + node], 1);
        }
    }
};

// ------------------------------------------------------------------ region adjacency graph
struct UlkRagTable {
    u64 *key; u64 *wk; u32 *k3; u32 *cnt; u64 *shi; u64 *slo;
    u64 mask;
};
ULK_HD ULK_INLINE u64 ulk_hash64(u64 x)
{
    x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
    return x;
}
struct K_RagInsert {
    const i32 *bas; const i32 *lab; UlkCt ct; UlkDims dm; UlkRagTable T; double fixscale;
    ULK_D void operator()(i64 vv, i32 z, i32 y, i32 x) const
    {
        const i32 v = (i32)vv;
        const bool f[3] = {x + 1 < dm.W, y + 1 < dm.H, z + 1 < dm.D};
        const i32 u[3] = {f[0] ? v + 1 : v, f[1] ? v + dm.W : v, f[2] ? v + dm.W * dm.H : v};
        const i32 b = bas[v], l = lab[v];
        const i32 cb[3] = {bas[u[0]], bas[u[1]], bas[u[2]]};
        const i32 cl[3] = {lab[u[0]], lab[u[1]], lab[u[2]]};
        if (b < 0) return;
        for (int d = 0; d < 3; ++d) {
            if (!f[d] || cl[d] != l || cb[d] == b) continue;
            const i32 c = cb[d];
            u64 key = b < c ? ((u64)(u32)b << 32) | (u32)c : ((u64)(u32)c << 32) | (u32)b;
            float w = ct.w(v, u[d]);
            u32 ek = 3u * (u32)v + d;
            u64 h = ulk_hash64(key) & T.mask;
            for (;;) {
                u64 old = ulk_atomic_cas(&T.key[h], ULK_NONE64, key);
                if (old == ULK_NONE64 || old == key) break;
                h = (h + 1) & T.mask;
            }
            ulk_atomic_min(&T.wk[h], ((u64)ulk_f2ord(w) << 32) | ek);
            ulk_atomic_min(&T.k3[h], ek);
            ulk_atomic_add(&T.cnt[h], 1u);
            i64 fx = (i64)((double)w * fixscale);
            ulk_atomic_add(&T.shi[h], (u64)(fx >> 32));
            ulk_atomic_add(&T.slo[h], (u64)(fx & 0xffffffffll));
        }
    }
};
struct K_RagFlag {
    const u64 *key; i32 *flag;
    ULK_D void operator()(i64 i) const { flag[i] = key[i] != ULK_NONE64 ? 1 : 0; }
};
struct K_RagGather {
    UlkRagTable T; const i32 *pre; const i32 *g_lab;
    i32 *ea, *eb; u64 *wk; u32 *k3, *cnt; u64 *shi, *slo; u32 *elab; u32 *eidx;
    ULK_D void operator()(i64 i) const
    {
        u64 k = T.key[i];
        if (k == ULK_NONE64) return;
        i32 j = pre[i];
        ea[j] = (i32)(k >> 32); eb[j] = (i32)(u32)k;
        wk[j] = T.wk[i]; k3[j] = T.k3[i]; cnt[j] = T.cnt[i]; shi[j] = T.shi[i]; slo[j] = T.slo[i];
        elab[j] = (u32)g_lab[(i32)(k >> 32)];
        eidx[j] = (u32)j;
    }
};
template <class T> struct K_Gather {
    const T *src; const u32 *idx; T *dst;
    ULK_D void operator()(i64 i) const { dst[i] = src[idx[i]]; }
};
struct K_Hist {
    const u32 *lab; i32 *hist;
    ULK_D void operator()(i64 i) const { ulk_atomic_add(&hist[lab[i] - 1], 1); }
};

// ------------------------------------------------------------------ raster-key bottleneck of a basin
// hg.filter_small_nodes_from_tree re-binarises the tree with bpt_canonical(leaf_graph,
// saliency); inside a basin every edge has saliency 0, so Kruskal runs there by edge index
// alone and the basin's altitude-0 node is created by the LAST edge it accepts: the
// bottleneck of the raster-key MST.  K0 = max over voxels of their smallest in-basin edge
// key is a lower bound that is tight whenever edges <= K0 already connect the basin.
// edges of a basin whose raster key is <= K0[basin].  Runs of x edges between voxels that are
// neighbours in the (raster-ordered) voxel list and sit in one warp are hooked to their first
// voxel at initialisation; K_K3Union joins the runs (same scheme as the component labelling).
struct K_K3Init {
    const i32 *bas; const u32 *K0; i32 *P;
    ULK_D ULK_INLINE bool xlink(i64 v, i32 x, i32 b) const
    {
        return b >= 0 && x > 0 && bas[v - 1] == b && 3u * (u32)(v - 1) <= K0[b];
    }
#ifdef ULK_EMU
    void operator()(i64 v, i32, i32, i32 x) const
    {
        i32 b = bas[v];
        P[v] = b < 0 ? -1 : ((x & 31) != 0 && xlink(v, x, b) ? P[v - 1] : (i32)v);
    }
#else
    ULK_D void warp(i64 v, i32, i32, i32 x, bool valid) const
    {
        const int lane = threadIdx.x & 31;
        i32 b = valid ? bas[v] : -1;
        i32 vprev = __shfl_up_sync(0xffffffffu, (i32)v, 1);
        u32 linked = __ballot_sync(0xffffffffu, lane > 0 && vprev == (i32)v - 1 && xlink(v, x, b));
        if (valid) P[v] = b < 0 ? -1 : (i32)v - (lane - ulk_run_start(linked, lane));
    }
#endif
};
struct K_K3Union {
    const i32 *bas; UlkDims dm; const u32 *K0; i32 *P;
    // `boundary`: the x edge to v-1 was not joined at initialisation
    ULK_D ULK_INLINE void join(i32 v, i32 z, i32 y, i32 x, bool boundary) const
    {
        const bool hx = x > 0, hy = y > 0, hz = z > 0;
        const i32 uy = hy ? v - dm.W : v, uz = hz ? v - dm.W * dm.H : v;
        const i32 b = bas[v];
        const i32 bl = bas[hx ? v - 1 : v], by = bas[uy], byl = bas[hx ? uy - 1 : uy], bz = bas[uz], bzl = bas[hx ? uz - 1 : uz];
        if (b < 0) return;
        const u32 kmax = K0[b];
        const bool left = hx && bl == b && 3u * (u32)(v - 1) <= kmax;
        if (left && boundary) ulk_vunion(P, v, v - 1);
        // (u-1,u), (u-1,v-1) carry smaller keys than (u,v): with (v-1,v) they already join u and v
        if (hy && by == b && 3u * (u32)uy + 1u <= kmax && !(left && byl == b)) ulk_vunion(P, v, uy);
        if (hz && bz == b && 3u * (u32)uz + 2u <= kmax && !(left && bzl == b)) ulk_vunion(P, v, uz);
    }
#ifdef ULK_EMU
    void operator()(i64 v, i32 z, i32 y, i32 x) const { join((i32)v, z, y, x, (x & 31) == 0); }
#else
    ULK_D void warp(i64 v, i32 z, i32 y, i32 x, bool valid) const
    {
        i32 vprev = __shfl_up_sync(0xffffffffu, (i32)v, 1);
        if (valid) join((i32)v, z, y, x, (threadIdx.x & 31) == 0 || vprev != (i32)v - 1);
    }
#endif
};
struct K_K3FlattenCount {
    const i32 *bas; i32 *P; i32 *nroot; i32 *nmulti;
    ULK_D void operator()(i64 v, i32, i32, i32) const
    {
        if (P[v] < 0) return;
        i32 r = ulk_vfind(P, (i32)v);
        P[v] = r;
        if (r != (i32)v) return;
        i32 old = ulk_atomic_add(&nroot[bas[v]], 1);
        if (old == 1) ulk_atomic_add(nmulti, 1);
    }
};
struct K_K3FindBest {     // Boruvka step on the basins that are still split
    const i32 *bas; UlkDims dm; const i32 *nroot; const i32 *P; u32 *best;
    ULK_D void operator()(i64 vv, i32 z, i32 y, i32 x) const
    {
        i32 v = (i32)vv;
        i32 b = bas[v];
        if (b < 0 || nroot[b] <= 1) return;
        i32 r = ulk_vfind(P, v);
        u32 k = ULK_NONE32;
        for (int d = 0; d < 3; ++d) {
            i32 st = dm.stride(d);
            if (dm.bwd(d, z, y, x)) { i32 u = v - st; if (bas[u] == b && ulk_vfind(P, u) != r) { u32 kk = 3u * (u32)u + d; if (kk < k) k = kk; } }
            if (dm.fwd(d, z, y, x)) { i32 u = v + st; if (bas[u] == b && ulk_vfind(P, u) != r) { u32 kk = 3u * (u32)v + d; if (kk < k) k = kk; } }
        }
        if (k != ULK_NONE32) ulk_atomic_min(&best[r], k);
    }
};
struct K_K3Hook {
    const i32 *bas; UlkDims dm; i32 *P; u32 *best; u32 *K0;
    ULK_D void operator()(i64 vv, i32, i32, i32) const
    {
        i32 v = (i32)vv;
        if (bas[v] < 0) return;
        u32 k = best[v];
        if (k == ULK_NONE32) return;
        best[v] = ULK_NONE32;
        i32 a = (i32)(k / 3u), c = a + dm.stride((int)(k % 3u));
        ulk_vunion(P, a, c);
        ulk_atomic_max(&K0[bas[v]], k);
    }
};

// ------------------------------------------------------------------ label writing (pass A -> pass B)
struct K_SmallCcRecords {   // one region record per component that is not oversegmented
    const i32 *cc_big; u32 *rec_cc; u64 *rec_w; u64 *rec_k;
    ULK_D void operator()(i64 c) const
    {
        rec_cc[c] = cc_big[c] >= 0 ? ULK_NONE32 : (u32)c;
        rec_w[c] = 0; rec_k[c] = 0;
    }
};
struct K_RegionRecords {    // one record per cut region of an oversegmented component
    const i32 *region_of; const u64 *rkey_w, *rkey_k; const i32 *g_lab; const i32 *big_cc; i32 ncc;
    u32 *rec_cc; u64 *rec_w; u64 *rec_k;
    ULK_D void operator()(i64 g) const
    {
        i32 i = ncc + (i32)g;
        if (region_of[g] != (i32)g) { rec_cc[i] = ULK_NONE32; rec_w[i] = 0; rec_k[i] = 0; return; }
        rec_cc[i] = (u32)big_cc[g_lab[g] - 1];
        rec_w[i] = rkey_w[g]; rec_k[i] = rkey_k[g];
    }
};
struct K_RankToLabel {
    const u32 *perm; const u32 *rec_cc_sorted; i32 *label_of_rec;
    ULK_D void operator()(i64 i) const
    {
        label_of_rec[perm[i]] = rec_cc_sorted[i] == ULK_NONE32 ? 0 : (i32)i + 1;
    }
};
// final region labels over the mode-1 component labels, in place
struct K_WriteLabels {
    i32 ncc; const i32 *basA; const i32 *region_of; const i32 *label_of_rec;
    i32 *lab;
    ULK_D void operator()(i64 v, i32, i32, i32) const
    {
        i32 l = lab[v];
        if (l > 0) lab[v] = label_of_rec[ncc + region_of[basA[v]]];
        else if (l <= -2) lab[v] = label_of_rec[-l - 2];
        else lab[v] = 0;
    }
};
struct K_CopyLabels {       // region labels as the caller sees them: 0 on the background
    const i32 *lab; i32 *out;
    ULK_D void operator()(i64 v) const { i32 l = lab[v]; out[v] = l > 0 ? l : 0; }
};
template <class T> struct K_FillAt {
    T *p; T val;
    ULK_D void operator()(i64 v, i32, i32, i32) const { p[v] = val; }
};
