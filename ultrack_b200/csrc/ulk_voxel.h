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
#define ULK_PK_OUT 0xFFFFFFFEu    // steepest-edge key of a list voxel outside every region (dropped component)

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
#ifndef ULK_EMU
// lanes hold consecutive entries of a raster-ordered list, so lanes that share a key are almost
// always neighbours: `heads` has a bit for every lane whose key differs from the lane below it;
// returns the lane mask of the caller's run (runs need not be maximal groups -- a key that
// comes back later in the warp simply forms a second run)
ULK_D ULK_INLINE u32 ulk_run_mask(u32 heads, int lane)
{
    const u32 upto = 0xffffffffu >> (31 - lane);
    const int start = 31 - __clz((heads | 1u) & upto);
    const u32 above = heads & ~upto;
    const int end = above ? __ffs(above) - 2 : 31;
    return (0xffffffffu >> (31 - end)) & (0xffffffffu << start);
}
// roots of ULK_MULTI chains walked in lockstep (x < 0: no chain); every round issues one load
// per unfinished chain, all independent
ULK_D ULK_INLINE void ulk_vfind_multi(const i32 *P, i32 (&x)[ULK_MULTI])
{
    bool more;
    do {
        more = false;
        i32 p[ULK_MULTI];
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) p[u] = x[u] >= 0 ? P[x[u]] : -1;
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u)
            if (p[u] != x[u]) { x[u] = p[u]; more = true; }
    } while (more);
}
#endif

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
    int skip_ct;      // float32 contours are read in place later: K_MinEdge finds max |ct| over the regions
    ULK_D ULK_INLINE bool stage(i64 v, u32 &bits) const
    {
        double f = ulk_load_as_double(fg, fgt, v);
        bits = 0;
        if (ctt == ULK_F32) {       // read in place, never staged
            if (!skip_ct) {
                float a = fabsf(((const float *)ct)[v]);
                if (a <= 3.0e38f) memcpy(&bits, &a, 4);
            }
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
    // every neighbour is loaded up front (clamped to v at the frame border) so the loads
    // overlap; only the sign is used, and the sign of an entry never changes
    struct S1 { i32 v, p0, pl, py, pyl, pz, pzl; u32 h; };
    typedef S1 S2;
    ULK_D ULK_INLINE S1 load1(i32 v, i32 z, i32 y, i32 x, bool ok) const
    {
        S1 s;
        s.v = v; s.p0 = -1; s.h = 0;
        if (!ok) return s;
        const bool hx = x > 0, hy = y > 0, hz = z > 0;
        const i32 uy = hy ? v - dm.W : v, uz = hz ? v - dm.W * dm.H : v;
        s.h = (hx ? 1u : 0u) | (hy ? 2u : 0u) | (hz ? 4u : 0u) | ((x & 31) == 0 ? 8u : 0u);
        s.p0 = P[v];
        s.pl = P[hx ? v - 1 : v]; s.py = P[uy]; s.pyl = P[hx ? uy - 1 : uy]; s.pz = P[uz]; s.pzl = P[hx ? uz - 1 : uz];
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &s, bool) const { return s; }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const
    {
        const i32 v = s.v;
        if (s.p0 < 0) return;
        if (s.p0 == v) cnt[v] = 0;      // a final root is never hooked: its size counter starts at 0
        const bool left = (s.h & 1u) && s.pl >= 0;
        if (left && (s.h & 8u)) ulk_vunion(P, v, v - 1);
        if ((s.h & 2u) && s.py >= 0 && !(left && s.pyl >= 0)) ulk_vunion(P, v, v - dm.W);
        if ((s.h & 4u) && s.pz >= 0 && !(left && s.pzl >= 0)) ulk_vunion(P, v, v - dm.W * dm.H);
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
    ULK_D void multi(const i32 (&v)[ULK_MULTI], const bool (&ok)[ULK_MULTI]) const
    {
        i32 r[ULK_MULTI];
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) r[u] = ok[u] ? P[v[u]] : -1;     // foreground only: P >= 0
        ulk_vfind_multi(P, r);
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) {
            if (r[u] >= 0) P[v[u]] = r[u];
            const int lane = threadIdx.x & 31;
            const i32 rp_ = __shfl_up_sync(0xffffffffu, r[u], 1);
            const u32 grp = ulk_run_mask(__ballot_sync(0xffffffffu, lane == 0 || rp_ != r[u]), lane);
            if (r[u] < 0) continue;
            if ((threadIdx.x & 31) == __ffs(grp) - 1) atomicAdd(&cnt[r[u]], __popc(grp));
            if (r[u] == v[u]) roots[atomicAdd(nroots, 1)] = r[u];
        }
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
    i32 *biglist; i32 *nbig;       // mode 1: the voxels of the components above max_area (32-entry chunks in raster order)
    struct S1 { i32 v, r; };
    struct S2 { i32 v, c; };
    ULK_D ULK_INLINE S1 load1(i32 v, i32, i32, i32, bool ok) const { S1 s; s.v = v; s.r = ok ? lab[v] : -1; return s; }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const { S2 s; s.v = a.v; s.c = a.r >= 0 ? code[a.r] : -1; return s; }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const
    {
        const i32 c = s.c;
        i32 g = -1;
        if (c >= 0) {
            if (mode == 0) lab[s.v] = c;
            else if (c == 0) lab[s.v] = -1;
            else {
                g = cc_big[c - 1];
                lab[s.v] = g >= 0 ? g + 1 : -(c - 1) - 2;
            }
        }
        if (mode == 1) {
#ifdef ULK_EMU
            if (g >= 0) biglist[(*nbig)++] = s.v;
#else
            ulk_warp_append(g >= 0, s.v, biglist, nbig);
#endif
        }
    }
#ifndef ULK_EMU
    // mode 1, ULK_MULTI entries per thread: the three dependent loads of all entries in lockstep
    // and ONE list reservation per warp for all of them
    ULK_D void multi(const i32 (&v)[ULK_MULTI], const bool (&ok)[ULK_MULTI]) const
    {
        i32 r[ULK_MULTI], c[ULK_MULTI], g[ULK_MULTI];
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) r[u] = ok[u] ? lab[v[u]] : -1;
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) c[u] = r[u] >= 0 ? code[r[u]] : -1;
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) g[u] = c[u] > 0 ? cc_big[c[u] - 1] : -1;
        const int lane = threadIdx.x & 31;
        u32 m[ULK_MULTI];
        i32 total = 0;
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) {
            if (c[u] >= 0) lab[v[u]] = c[u] == 0 ? -1 : (g[u] >= 0 ? g[u] + 1 : -(c[u] - 1) - 2);
            m[u] = __ballot_sync(0xffffffffu, g[u] >= 0);
            total += __popc(m[u]);
        }
        if (total == 0) return;
        i32 base = 0;
        if (lane == 0) base = atomicAdd(nbig, total);
        base = __shfl_sync(0xffffffffu, base, 0);
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) {
            if (g[u] >= 0) biglist[base + __popc(m[u] & ((1u << lane) - 1u))] = v[u];
            base += __popc(m[u]);
        }
    }
#endif
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
    u32 *cmax;        // when set (float32 contours): max |ct| over the voxels of the regions, as float bits
    // neighbour indices clamped to v at the frame border: all loads are issued together
    struct S1 { i32 v, l; u32 h; i32 ln[6]; float c, cn[6]; };
    typedef S1 S2;
    ULK_D ULK_INLINE i32 nb(i32 v, int i, u32 h) const
    {
        const i32 sy = dm.W, sz = dm.W * dm.H;
        const i32 off = i == 0 ? -sz : i == 1 ? -sy : i == 2 ? -1 : i == 3 ? 1 : i == 4 ? sy : sz;
        return ((h >> i) & 1u) ? v + off : v;
    }
    ULK_D ULK_INLINE S1 load1(i32 v, i32 z, i32 y, i32 x, bool ok) const
    {
        S1 s;
        s.v = v; s.l = 0; s.h = 0; s.c = 0.f;
        if (!ok) return s;
        s.h = (z > 0 ? 1u : 0u) | (y > 0 ? 2u : 0u) | (x > 0 ? 4u : 0u) | (x + 1 < dm.W ? 8u : 0u) | (y + 1 < dm.H ? 16u : 0u) | (z + 1 < dm.D ? 32u : 0u);
        s.l = lab[v];
#ifdef ULK_EMU
        for (int i = 0; i < 6; ++i) s.ln[i] = lab[nb(v, i, s.h)];
        if (ct.f) { s.c = ct.f[v]; for (int i = 0; i < 6; ++i) s.cn[i] = ct.f[nb(v, i, s.h)]; }
#else
#pragma unroll
        for (int i = 0; i < 6; ++i) s.ln[i] = lab[nb(v, i, s.h)];
        if (ct.f) {
            s.c = ct.f[v];
#pragma unroll
            for (int i = 0; i < 6; ++i) s.cn[i] = ct.f[nb(v, i, s.h)];
        }
#endif
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &s, bool) const { return s; }
    ULK_D ULK_INLINE void run(const S2 &s, bool ok) const
    {
        const i32 v = s.v, l = s.l;
        if (cmax) {
            u32 bits = 0;
            if (l > 0) { float a = fabsf(s.c); if (a <= 3.0e38f) memcpy(&bits, &a, 4); }
#ifdef ULK_EMU
            if (bits > *cmax) *cmax = bits;
#else
            const u32 m = __reduce_max_sync(0xffffffffu, bits);
            if ((threadIdx.x & 31) == 0 && m > *cmax) atomicMax(cmax, m);
#endif
        }
        if (l <= 0) {          // K_Hook1 reads the key instead of the label; lanes past the list end write nothing
            if (ok) pk[v] = ULK_PK_OUT;
            return;
        }
        u32 bw = ULK_NONE32, bk = ULK_NONE32;
#ifndef ULK_EMU
#pragma unroll
#endif
        for (int i = 0; i < 6; ++i) {
            if (!((s.h >> i) & 1u) || s.ln[i] != l) continue;
            const i32 u = nb(v, i, s.h);
            // candidates in increasing edge-key order: the first strictly smaller weight wins
            const i32 a = i < 3 ? u : v, b = i < 3 ? v : u;
            const float wf = ct.f ? (i < 3 ? ulk_wsum(s.cn[i] + s.c) : ulk_wsum(s.c + s.cn[i])) : ct.w(a, b);
            const u32 w = ulk_f2ord(wf);
            const u32 d = i == 0 || i == 5 ? 2u : (i == 1 || i == 4 ? 1u : 0u);
            if (w < bw) { bw = w; bk = 3u * (u32)a + d; }
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
// reads only the steepest-edge keys (its own and the other endpoint's); the level of a root is
// fetched by the roots alone, and `state` only exists at roots (nothing reads it elsewhere)
struct K_Hook1 {
    const u32 *mw, *pk; UlkDims dm;
    i32 *bp; u8 *state; u32 *Lr; u32 *best; i32 *nbasins;
    struct S1 { i32 v; u32 k; };
    struct S2 { i32 v, u; u32 k, ku; };
    ULK_D ULK_INLINE S1 load1(i32 v, i32, i32, i32, bool ok) const
    {
        S1 s;
        s.v = v; s.k = ULK_PK_OUT;
        if (ok) s.k = pk[v];
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool ok) const
    {
        S2 s;
        s.v = a.v; s.k = a.k; s.u = a.v; s.ku = ULK_NONE32;
        if (ok && a.k < ULK_PK_OUT) { s.u = ulk_edge_other(a.k, a.v, dm); s.ku = pk[s.u]; }
        return s;
    }
    ULK_D ULK_INLINE void run(const S2 &s, bool ok) const
    {
        if (!ok) return;
        const i32 v = s.v;
        if (s.k == ULK_PK_OUT) { bp[v] = -1; return; }
        if (s.k == ULK_NONE32) {
            bp[v] = v; state[v] = 1; Lr[v] = ULK_NONE32; best[v] = ULK_NONE32;
            ulk_atomic_add(nbasins, 1);
            return;
        }
        if (s.ku == s.k && v < s.u) { bp[v] = v; state[v] = 0; Lr[v] = mw[v]; best[v] = ULK_NONE32; }
        else bp[v] = s.u;
    }
};
struct K_Compress {
    i32 *bp;
    ULK_D void operator()(i64 v) const { if (bp[v] >= 0) bp[v] = ulk_vfind(bp, (i32)v); }
};
// path compression of the hook forest, fused with the lists of the flow rounds: growing
// roots, and candidates = voxels whose steepest weight equals their supervertex level
struct K_CompressCand {
    i32 *bp; const u32 *mw; const u32 *Lr; const u8 *state;
    i32 *cand; i32 *ncand; i32 *groots; i32 *ngroots; i32 *sroots; i32 *nsroots;
#ifdef ULK_EMU
    void operator()(i64 vv, i32, i32, i32) const
    {
        i32 v = (i32)vv;
        if (bp[v] < 0) return;
        i32 r = ulk_vfind(bp, v);
        bp[v] = r;
        if (r == v) sroots[(*nsroots)++] = v;
        if (r == v && state[v] == 0) groots[(*ngroots)++] = v;
        if (state[r] == 0 && mw[v] == Lr[r]) cand[(*ncand)++] = v;
    }
#else
    ULK_D void multi(const i32 (&v)[ULK_MULTI], const bool (&ok)[ULK_MULTI]) const
    {
        i32 r[ULK_MULTI];
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) r[u] = ok[u] ? bp[v[u]] : -1;    // -1 where the voxel is not in a region
        ulk_vfind_multi(bp, r);
        u8 st[ULK_MULTI]; u32 m[ULK_MULTI], L[ULK_MULTI];
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u)
            if (r[u] >= 0) { bp[v[u]] = r[u]; st[u] = state[r[u]]; m[u] = mw[v[u]]; L[u] = Lr[r[u]]; }
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) {
            const bool root = r[u] >= 0 && r[u] == v[u];
            const bool growing = r[u] >= 0 && st[u] == 0;
            ulk_warp_append(root, v[u], sroots, nsroots);
            ulk_warp_append(growing && root, v[u], groots, ngroots);
            ulk_warp_append(growing && m[u] == L[u], v[u], cand, ncand);
        }
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
#ifndef ULK_EMU
    ULK_D void multi(const i32 (&v)[ULK_MULTI], const bool (&ok)[ULK_MULTI]) const
    {
        i32 r[ULK_MULTI];
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) r[u] = ok[u] ? bp[v[u]] : -1;
        ulk_vfind_multi(bp, r);
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) {
            if (r[u] < 0) continue;
            bp[v[u]] = r[u];
            if (r[u] == v[u]) rootkeys[ulk_atomic_add(nroots, 1)] = ((u64)(u32)lab[v[u]] << 32) | (u32)v[u];
        }
    }
#endif
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
// bp[v] is the supervertex root of v (K_CompressCand) and bp[root] the basin root
// (K_CompressRoots on the root list): two hops, then the basin id
struct K_AssignBasin {     // background stays at the memset value -1
    const i32 *bp; const u32 *gid_at_root;
    i32 *bas;
    struct S1 { i32 v, r; };
    struct S2 { i32 v, rr; };
    ULK_D ULK_INLINE S1 load1(i32 v, i32, i32, i32, bool ok) const { S1 s; s.v = ok ? v : -1; s.r = ok ? bp[v] : -1; return s; }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const { S2 s; s.v = a.v; s.rr = a.r < 0 ? -1 : bp[a.r]; return s; }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const { if (s.v >= 0) bas[s.v] = s.rr < 0 ? -1 : (i32)gid_at_root[s.rr]; }
};
// instead of a memset of the whole basin-id volume: -1 on the neighbours of the list's voxels
// that are outside every region (the only entries outside the list a stencil kernel reads)
struct K_ClearHalo {
    const i32 *lab; UlkDims dm; i32 *bas;
    ULK_D void operator()(i64 vv, i32 z, i32 y, i32 x) const
    {
        const i32 v = (i32)vv;
        for (int d = 0; d < 3; ++d) {
            const i32 st = dm.stride(d);
            if (dm.bwd(d, z, y, x) && lab[v - st] <= 0) bas[v - st] = -1;
            if (dm.fwd(d, z, y, x) && lab[v + st] <= 0) bas[v + st] = -1;
        }
    }
};
// flow edges are spanning-tree edges of the basin of their supervertex
struct K_FlowTop {
    const i32 *flow_root; const u32 *flow_key; const i32 *bas; UlkCt ct; UlkDims dm; u64 *b_top; u32 *Tk;
    ULK_D void operator()(i64 j) const
    {
        u32 fk = flow_key[j];
        i32 a = (i32)(fk / 3u), e = a + dm.stride((int)(fk % 3u));
        i32 b = bas[flow_root[j]];
        ulk_atomic_max(&b_top[b], ((u64)ulk_f2ord(ct.w(a, e)) << 32) | fk);
        if (Tk) ulk_atomic_max(&Tk[b], fk);
    }
};
// A basin whose spanning-tree edges (steepest edges + flow edges) all carry raster keys <= K0 is
// connected by edges <= K0: its bottleneck IS K0 and it needs no union-find check.  need[b] = 2
// marks the others (the value the split-list selection tests for).
struct K_K3Need {
    const u32 *Tk, *K0; i32 *need; i32 *nneed; i32 *needlist;
    ULK_D void operator()(i64 b) const
    {
        bool n = Tk[b] > K0[b];
        need[b] = n ? 2 : 0;
        if (n) needlist[ulk_atomic_add(nneed, 1)] = (i32)b;
    }
};
#ifndef ULK_EMU
// the voxels of the basins that need the check, one warp per basin scanning its bounding box
// row by row (x-runs of a basin stay contiguous in the list, which is all K_K3Init relies on)
struct K_K3Collect {
    const i32 *needlist; i64 nneed; const i32 *bas; const i32 *bbmin, *bbmax; UlkDims dm;
    i32 *list; i32 *count;
    ULK_D void warp(i64 ii, bool) const
    {
        const int lane = threadIdx.x & 31;
        const i64 i = ii >> 5;
        if (i >= nneed) return;
        const i32 b = needlist[i];
        const i32 z0 = bbmin[3 * b], y0 = bbmin[3 * b + 1], x0 = bbmin[3 * b + 2];
        const i32 z1 = bbmax[3 * b], y1 = bbmax[3 * b + 1], x1 = bbmax[3 * b + 2];
        for (i32 z = z0; z <= z1; ++z)
            for (i32 y = y0; y <= y1; ++y) {
                const i32 row = (z * dm.H + y) * dm.W;
                for (i32 xb = x0; xb <= x1; xb += 32) {
                    const i32 x = xb + lane;
                    const bool hit = x <= x1 && bas[row + x] == b;
                    ulk_warp_append(hit, row + x, list, count);
                }
            }
    }
};
#endif
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
    int want_k0; u32 *K0; u32 *Tk;      // Tk: largest raster key among the basin's steepest edges
    i32 *xl; i32 *nxl;                  // voxels with a forward edge into another basin of their region
    const i32 *g_lab;                   // region label of a basin (the voxel's own label is not read)
    // phase 1: the voxel's own entries and its neighbours' basins (indices clamped at the border)
    struct S1 { i32 v, z, y, x, b, l; u32 m, k; i32 a[3], c[3]; };
    // phase 2: what depends on them -- the basin level, the weight across the steepest edge, the
    // labels of forward neighbours in another basin
    struct S2 { i32 v, z, y, x, b; u64 top; u32 nc, kmin, jkv; bool plateau; };
    ULK_D ULK_INLINE S1 load1(i32 v, i32 z, i32 y, i32 x, bool ok) const
    {
        S1 s;
        s.v = v; s.z = z; s.y = y; s.x = x; s.b = -1;
        if (!ok) return s;
        const i32 sy = dm.W, sz = dm.W * dm.H;
        s.b = bas[v];
        s.a[0] = bas[x + 1 < dm.W ? v + 1 : v]; s.a[1] = bas[y + 1 < dm.H ? v + sy : v]; s.a[2] = bas[z + 1 < dm.D ? v + sz : v];
        s.c[0] = s.c[1] = s.c[2] = -1;
        if (want_k0) { s.c[0] = bas[x > 0 ? v - 1 : v]; s.c[1] = bas[y > 0 ? v - sy : v]; s.c[2] = bas[z > 0 ? v - sz : v]; }
        s.m = mw[v]; s.k = pk[v]; s.l = 0;
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const
    {
        S2 o;
        o.v = a.v; o.z = a.z; o.y = a.y; o.x = a.x; o.b = a.b;
        o.top = 0; o.nc = 0; o.kmin = ULK_NONE32; o.jkv = ULK_JK_CORE; o.plateau = false;
        const i32 b = a.b, v = a.v;
        if (b < 0) return o;
        const i32 sy = dm.W, sz = dm.W * dm.H;
        const bool f0 = a.x + 1 < dm.W, f1 = a.y + 1 < dm.H, f2 = a.z + 1 < dm.D;
        // issue the dependent loads first
        const bool q0 = f0 && a.a[0] != b, q1 = f1 && a.a[1] != b, q2 = f2 && a.a[2] != b;
        const i32 l0 = q0 ? lab[v + 1] : 0, l1 = q1 ? lab[v + sy] : 0, l2 = q2 ? lab[v + sz] : 0;
        const i32 lown = (q0 || q1 || q2) ? g_lab[b] : 0;
        u32 bl = 0, mo = 0;
        const bool has_edge = a.k != ULK_NONE32;
        if (want_join) { bl = b_L[b]; if (has_edge) mo = mw[ulk_edge_other(a.k, v, dm)]; }
        o.top = has_edge ? ((u64)a.m << 32) | a.k : 0;
        if (q0 && l0 == lown) ++o.nc;
        if (q1 && l1 == lown) ++o.nc;
        if (q2 && l2 == lown) ++o.nc;
        u32 kk;
        if (f0 && a.a[0] == b) { kk = 3u * (u32)v; if (kk < o.kmin) o.kmin = kk; }
        if (f1 && a.a[1] == b) { kk = 3u * (u32)v + 1u; if (kk < o.kmin) o.kmin = kk; }
        if (f2 && a.a[2] == b) { kk = 3u * (u32)v + 2u; if (kk < o.kmin) o.kmin = kk; }
        if (want_k0) {
            if (a.x > 0 && a.c[0] == b) { kk = 3u * (u32)(v - 1); if (kk < o.kmin) o.kmin = kk; }
            if (a.y > 0 && a.c[1] == b) { kk = 3u * (u32)(v - sy) + 1u; if (kk < o.kmin) o.kmin = kk; }
            if (a.z > 0 && a.c[2] == b) { kk = 3u * (u32)(v - sz) + 2u; if (kk < o.kmin) o.kmin = kk; }
        }
        if (want_join) {
            if (a.m == bl || !has_edge) o.jkv = ULK_JK_CORE;
            else if (mo < a.m) o.jkv = a.k;
            else { o.jkv = ULK_NONE32; o.plateau = true; }
        }
        return o;
    }
#ifdef ULK_EMU
    void run(const S2 &o, bool) const
    {
        const i32 b = o.b;
        if (b < 0) return;
        if (want_join) jk[o.v] = o.jkv;
        b_size[b] += 1;
        if (o.top > b_top[b]) b_top[b] = o.top;
        *ncross += o.nc;
        if (o.nc) xl[(*nxl)++] = o.v;
        if (o.plateau) hp[(*nhp)++] = o.v;
        if (want_k0 && o.kmin != ULK_NONE32 && o.kmin > K0[b]) K0[b] = o.kmin;
        if (want_k0 && o.top != 0 && (u32)o.top > Tk[b]) Tk[b] = (u32)o.top;
        if (want_stats) {
            const i32 c[3] = {o.z, o.y, o.x};
            for (int d = 0; d < 3; ++d) {
                ulk_atomic_min(&bbmin[3 * b + d], c[d]);
                ulk_atomic_max(&bbmax[3 * b + d], c[d]);
                csum[3 * b + d] += c[d];
            }
        }
    }
#else
    // lanes of a warp that sit in the same basin AND the same row form a group (usually one or
    // two per warp): z and y are uniform in it and x grows with the lane, so the bounding box and
    // the coordinate sums need no reduction; one lane per group issues the atomics, and only
    // those that can still change the accumulator (a stale cached value only costs an atomic)
    // Every lane stays in the collectives (full mask): a reduction over a run is a segmented
    // shuffle reduction towards the run's first lane (partial-mask __reduce_*_sync serialises
    // over the distinct masks of a warp).
    ULK_D ULK_INLINE void run(const S2 &o, bool) const
    {
        const int lane = threadIdx.x & 31;
        const i32 b = o.b, v = o.v, x = o.x;
        u32 nct = __reduce_add_sync(0xffffffffu, o.nc);
        if (lane == 0 && nct) atomicAdd((unsigned long long *)ncross, (unsigned long long)nct);
        if (nct) ulk_warp_append(o.nc > 0, v, xl, nxl);
        if (want_join) {
            if (b >= 0) jk[v] = o.jkv;
            ulk_warp_append(o.plateau, v, hp, nhp);
        }
        const i32 row = v - x;
        const i32 bp_ = __shfl_up_sync(0xffffffffu, b, 1), rp_ = __shfl_up_sync(0xffffffffu, row, 1);
        const u32 grp = ulk_run_mask(__ballot_sync(0xffffffffu, lane == 0 || bp_ != b || rp_ != row), lane);
        const int first = __ffs(grp) - 1, last = 31 - __clz(grp);
        const bool lead = lane == first && b >= 0;
        const i32 n = last - first + 1;
        // the leader reads the accumulators it may have to update before the reductions start
        u64 cur_top = 0; u32 cur_k0 = 0, cur_tk = 0;
        i32 clo[3] = {0, 0, 0}, chi[3] = {0, 0, 0};
        if (lead) {
            cur_top = b_top[b];
            if (want_k0) { cur_k0 = K0[b]; cur_tk = Tk[b]; }
            if (want_stats) {
                clo[0] = bbmin[3 * b]; clo[1] = bbmin[3 * b + 1]; clo[2] = bbmin[3 * b + 2];
                chi[0] = bbmax[3 * b]; chi[1] = bbmax[3 * b + 1]; chi[2] = bbmax[3 * b + 2];
            }
        }
        u64 t = o.top;
        u32 km = o.kmin == ULK_NONE32 ? 0u : o.kmin, tk = o.top != 0 ? (u32)o.top : 0u;
        const i32 x1 = __shfl_sync(0xffffffffu, x, last);      // x of the run's last lane
        u32 sx = (u32)x;
        const bool contig = __shfl_sync(0xffffffffu, x1 - x + 1 == n ? 1 : 0, first) != 0;   // the leader's view
        const bool any_gap = __any_sync(0xffffffffu, b >= 0 && want_stats && !contig);
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const u64 t2 = __shfl_down_sync(0xffffffffu, t, off);
            const u32 km2 = __shfl_down_sync(0xffffffffu, km, off), tk2 = __shfl_down_sync(0xffffffffu, tk, off);
            u32 sx2 = 0;
            if (any_gap) sx2 = __shfl_down_sync(0xffffffffu, sx, off);
            if (lane + off <= last) {
                t = t2 > t ? t2 : t;
                km = km2 > km ? km2 : km;
                tk = tk2 > tk ? tk2 : tk;
                sx += sx2;
            }
        }
        if (!lead) return;
        if (contig) sx = (u32)(((i64)(x + x1) * n) >> 1);
        atomicAdd(&b_size[b], n);
        if (t > cur_top) atomicMax((unsigned long long *)&b_top[b], (unsigned long long)t);
        if (want_k0 && km > cur_k0) atomicMax(&K0[b], km);
        if (want_k0 && tk > cur_tk) atomicMax(&Tk[b], tk);
        if (want_stats) {
            const i32 z = o.z, y = o.y;
            i32 *lo = bbmin + 3 * b, *hi = bbmax + 3 * b;
            i64 *cs = csum + 3 * b;
            if (z < clo[0]) atomicMin(&lo[0], z);
            if (z > chi[0]) atomicMax(&hi[0], z);
            if (y < clo[1]) atomicMin(&lo[1], y);
            if (y > chi[1]) atomicMax(&hi[1], y);
            if (x < clo[2]) atomicMin(&lo[2], x);
            if (x1 > chi[2]) atomicMax(&hi[2], x1);
            atomicAdd((unsigned long long *)&cs[0], (unsigned long long)((i64)z * n));
            atomicAdd((unsigned long long *)&cs[1], (unsigned long long)((i64)y * n));
            atomicAdd((unsigned long long *)&cs[2], (unsigned long long)sx);
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
    struct S1 { i32 b; u32 j, m; };
    struct S2 { i32 o, node, p; u64 J; };
    ULK_D ULK_INLINE S1 load1(i32 v, i32, i32, i32, bool ok) const
    {
        S1 s;
        s.b = -1; s.j = ULK_JK_CORE; s.m = 0;
        if (ok) { s.b = bas[v]; s.j = jk[v]; s.m = mw[v]; }
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const
    {
        S2 s;
        s.o = 0; s.node = -1; s.p = -1; s.J = 0;
        if (a.b < 0 || a.j == ULK_JK_CORE) return s;
        s.J = ((u64)a.m << 32) | a.j;
        const i32 st = bs[b_lab[a.b] - 1];
        s.o = 2 * st;
        s.node = a.b - st;
        s.p = Ipar[s.o + s.node];
        return s;
    }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const
    {
        if (s.node < 0) return;
        const i32 o = s.o, start = s.node;
        i32 node = start, p = s.p;
        // the parent's merge key and the grandparent index are loaded together: one memory
        // round trip per level of the climb
        while (p != node) {
            u64 w = Iwk[o + p];
            i32 pp = Ipar[o + p];
            if (!(w < s.J)) break;
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
    struct S1 { i32 v, b, l; u32 f; i32 cb[3], cl[3]; };
    typedef S1 S2;
    ULK_D ULK_INLINE S1 load1(i32 v, i32 z, i32 y, i32 x, bool ok) const
    {
        S1 s;
        s.v = v; s.b = -1; s.l = 0; s.f = 0;
        if (!ok) return s;
        s.f = (x + 1 < dm.W ? 1u : 0u) | (y + 1 < dm.H ? 2u : 0u) | (z + 1 < dm.D ? 4u : 0u);
        const i32 u0 = (s.f & 1u) ? v + 1 : v, u1 = (s.f & 2u) ? v + dm.W : v, u2 = (s.f & 4u) ? v + dm.W * dm.H : v;
        s.b = bas[v]; s.l = lab[v];
        s.cb[0] = bas[u0]; s.cb[1] = bas[u1]; s.cb[2] = bas[u2];
        s.cl[0] = lab[u0]; s.cl[1] = lab[u1]; s.cl[2] = lab[u2];
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &s, bool) const { return s; }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const
    {
        const i32 v = s.v, b = s.b, l = s.l;
        if (b < 0) return;
        for (int d = 0; d < 3; ++d) {
            if (!((s.f >> d) & 1u) || s.cl[d] != l || s.cb[d] == b) continue;
            const i32 c = s.cb[d];
            const i32 u = v + dm.stride(d);
            u64 key = b < c ? ((u64)(u32)b << 32) | (u32)c : ((u64)(u32)c << 32) | (u32)b;
            float w = ct.w(v, u);
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
    struct S1 { i32 v, b, bl; bool ok, cons; };     // cons: the previous lane holds voxel v-1 of the same row
    struct S2 { i32 v, b; bool ok, link; };
    ULK_D ULK_INLINE S1 load1(i32 v, i32, i32, i32 x, bool ok) const
    {
        S1 s;
        s.v = v; s.ok = ok; s.b = -1; s.bl = -1;
#ifdef ULK_EMU
        s.cons = x > 0 && (x & 31) != 0;
#else
        i32 vprev = __shfl_up_sync(0xffffffffu, v, 1);
        s.cons = (threadIdx.x & 31) > 0 && x > 0 && vprev == v - 1;
#endif
        if (ok) { s.b = bas[v]; s.bl = bas[x > 0 ? v - 1 : v]; }
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const
    {
        S2 s;
        s.v = a.v; s.b = a.b; s.ok = a.ok;
        s.link = a.ok && a.cons && a.b >= 0 && a.bl == a.b && 3u * (u32)(a.v - 1) <= K0[a.b];
        return s;
    }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const
    {
#ifdef ULK_EMU
        P[s.v] = s.b < 0 ? -1 : (s.link ? P[s.v - 1] : s.v);
#else
        const int lane = threadIdx.x & 31;
        u32 linked = __ballot_sync(0xffffffffu, s.link);
        if (s.ok) P[s.v] = s.b < 0 ? -1 : s.v - (lane - ulk_run_start(linked, lane));
#endif
    }
};
struct K_K3Union {
    const i32 *bas; UlkDims dm; const u32 *K0; i32 *P;
    struct S1 { i32 v, b, bl, by, byl, bz, bzl; u32 h; };     // h: x>0, y>0, z>0, run boundary
    struct S2 { i32 v; u32 todo; };                             // todo: unions with -x, -y, -z
    ULK_D ULK_INLINE S1 load1(i32 v, i32 z, i32 y, i32 x, bool ok) const
    {
        S1 s;
        s.v = v; s.b = -1; s.h = 0;
#ifdef ULK_EMU
        const bool boundary = (x & 31) == 0;
#else
        i32 vprev = __shfl_up_sync(0xffffffffu, v, 1);
        const bool boundary = (threadIdx.x & 31) == 0 || vprev != v - 1;
#endif
        if (!ok) return s;
        const bool hx = x > 0, hy = y > 0, hz = z > 0;
        const i32 uy = hy ? v - dm.W : v, uz = hz ? v - dm.W * dm.H : v;
        s.h = (hx ? 1u : 0u) | (hy ? 2u : 0u) | (hz ? 4u : 0u) | (boundary ? 8u : 0u);
        s.b = bas[v];
        s.bl = bas[hx ? v - 1 : v]; s.by = bas[uy]; s.byl = bas[hx ? uy - 1 : uy]; s.bz = bas[uz]; s.bzl = bas[hx ? uz - 1 : uz];
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const
    {
        S2 s;
        s.v = a.v; s.todo = 0;
        const i32 b = a.b, v = a.v;
        if (b < 0) return s;
        const u32 kmax = K0[b];
        const i32 uy = v - dm.W, uz = v - dm.W * dm.H;
        const bool left = (a.h & 1u) && a.bl == b && 3u * (u32)(v - 1) <= kmax;
        if (left && (a.h & 8u)) s.todo |= 1u;
        // (u-1,u), (u-1,v-1) carry smaller keys than (u,v): with (v-1,v) they already join u and v
        if ((a.h & 2u) && a.by == b && 3u * (u32)uy + 1u <= kmax && !(left && a.byl == b)) s.todo |= 2u;
        if ((a.h & 4u) && a.bz == b && 3u * (u32)uz + 2u <= kmax && !(left && a.bzl == b)) s.todo |= 4u;
        return s;
    }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const
    {
        if (s.todo & 1u) ulk_vunion(P, s.v, s.v - 1);
        if (s.todo & 2u) ulk_vunion(P, s.v, s.v - dm.W);
        if (s.todo & 4u) ulk_vunion(P, s.v, s.v - dm.W * dm.H);
    }
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
#ifndef ULK_EMU
    ULK_D void multi(const i32 (&v)[ULK_MULTI], const bool (&ok)[ULK_MULTI]) const
    {
        i32 r[ULK_MULTI];
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) r[u] = ok[u] ? P[v[u]] : -1;
        ulk_vfind_multi(P, r);
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) {
            if (r[u] < 0) continue;
            P[v[u]] = r[u];
            if (r[u] != v[u]) continue;
            i32 old = ulk_atomic_add(&nroot[bas[v[u]]], 1);
            if (old == 1) ulk_atomic_add(nmulti, 1);
        }
    }
#endif
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
    struct S1 { i32 v, l, ba; };
    struct S2 { i32 v, rec; };       // rec: index into label_of_rec, -1: label 0, -2: nothing to write
    ULK_D ULK_INLINE S1 load1(i32 v, i32, i32, i32, bool ok) const
    {
        S1 s;
        s.v = v; s.l = 0; s.ba = 0;
        if (ok) { s.l = lab[v]; if (s.l > 0) s.ba = basA[v]; }
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool ok) const
    {
        S2 s;
        s.v = a.v;
        if (!ok) s.rec = -2;
        else if (a.l > 0) s.rec = ncc + region_of[a.ba];
        else if (a.l <= -2) s.rec = -a.l - 2;
        else s.rec = -1;
        return s;
    }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const
    {
        if (s.rec == -2) return;
        lab[s.v] = s.rec >= 0 ? label_of_rec[s.rec] : 0;
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
