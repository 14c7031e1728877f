// ulk_voxel.h -- voxel-scale kernels of the segmentation path (functor bodies).
//
// Vocabulary follows ultrack / higra: a *region* is one hierarchy's support
// (core/segmentation/hierarchy.py:89-90), a *basin* is a catchment basin of the watershed cut
// of a region (altitude-0 node of hg.watershed_hierarchy_by_*), the *RAG* is the basin
// adjacency graph.
//
// RANK SPACE.  After the threshold the frame is only ever touched through its foreground
// voxels.  Every per-voxel array of this stage is indexed by the voxel's RANK in the raster-
// ordered foreground list (`fgl[rank] = voxel`), never by the voxel index: a kernel's own
// accesses are dense and coalesced whatever the foreground fraction, the arrays are n_fg long
// instead of V, and nothing has to be cleared on the background.  Neighbours are found through
// the foreground map `UlkFg`: one 8-byte record per 32-voxel row segment (foreground bits +
// rank of the segment's first foreground voxel); the rank of a neighbour is one record load and
// a population count, and the x-neighbours of rank r are simply r-1 / r+1 when fgl says so.
// Edge keys stay voxel based (3*voxel + direction): they reproduce the enumeration order of
// edges_from_mask (vendored/graph.py:41-57), i.e. higra's stable tie-break; rank order equals
// voxel order, so "smaller voxel" and "smaller rank" are interchangeable.
#pragma once
#include "ulk_common.h"

#define ULK_JK_CORE 0xFFFFFFFEu   // join key of voxels inside the regional minimum
#define ULK_PO_NONE (-1)          // steepest-edge endpoint: no in-region neighbour (a one-voxel basin seed)
#define ULK_PO_OUT (-2)           // steepest-edge endpoint: the voxel is outside every region (dropped component)

#ifdef ULK_EMU
ULK_INLINE int ulk_ffs(u32 x) { return __builtin_ffs((int)x); }
#else
ULK_D ULK_INLINE int ulk_ffs(u32 x) { return __ffs((int)x); }
#endif

// ------------------------------------------------------------------ foreground map
struct alignas(8) UlkOW { u32 w; i32 o; };     // bits of a 32-voxel row segment, rank of its first foreground voxel
struct UlkNbRaw { UlkOW e[4]; i32 vl, vr; };   // segment records at -z, -y, +y, +z; voxels of ranks r-1, r+1
struct UlkFg {
    const UlkOW *ow;      // [D*H*S]
    const i32 *fgl;       // [n] voxel of rank r
    i32 n;                // foreground voxels
    i32 S, SH;            // segments per row, per slice
    UlkDims dm;
    ULK_D ULK_INLINE UlkOW load(i32 t) const
    {
#ifdef ULK_EMU
        return ow[t];
#else
        const uint2 q = __ldg((const uint2 *)ow + t);
        UlkOW e;
        e.w = q.x; e.o = (i32)q.y;
        return e;
#endif
    }
    static ULK_D ULK_INLINE i32 rank_in(const UlkOW &e, int b)
    {
        return ((e.w >> b) & 1u) ? e.o + ulk_popc(e.w & ((1u << b) - 1u)) : -1;
    }
    ULK_D ULK_INLINE i32 seg(i32 z, i32 y, i32 x) const { return (z * dm.H + y) * S + (x >> 5); }
    // rank of an arbitrary voxel (-1: background); two multiply-shift divisions + one record load
    ULK_D ULK_INLINE i32 rank_of(i32 v) const
    {
        i32 z, y, x;
        dm.coords(v, z, y, x);
        return rank_in(load(seg(z, y, x)), x & 31);
    }
    // everything the six neighbour ranks of (rank r, segment t, row y, slice z) depend on; the
    // loads are independent of each other (issue them together, convert in the next phase)
    ULK_D ULK_INLINE UlkNbRaw nb_load(i32 r, i32 t, i32 z, i32 y, bool back, bool fwd) const
    {
        UlkNbRaw q;
        for (int i = 0; i < 4; ++i) { q.e[i].w = 0; q.e[i].o = 0; }
        q.vl = -1; q.vr = -1;
        if (back) {
            if (z > 0) q.e[0] = load(t - SH);
            if (y > 0) q.e[1] = load(t - S);
            if (r > 0) q.vl = fgl[r - 1];
        }
        if (fwd) {
            if (y + 1 < dm.H) q.e[2] = load(t + S);
            if (z + 1 < dm.D) q.e[3] = load(t + SH);
            if (r + 1 < n) q.vr = fgl[r + 1];
        }
        return q;
    }
    // neighbour ranks in increasing edge-key order: -z, -y, -x, +x, +y, +z (-1: border or background)
    ULK_D ULK_INLINE void nb_ranks(const UlkNbRaw &q, i32 r, i32 v, i32 x, i32 (&out)[6]) const
    {
        const int b = x & 31;
        out[0] = rank_in(q.e[0], b);
        out[1] = rank_in(q.e[1], b);
        out[2] = (x > 0 && q.vl == v - 1) ? r - 1 : -1;
        out[3] = (x + 1 < dm.W && q.vr == v + 1) ? r + 1 : -1;
        out[4] = rank_in(q.e[2], b);
        out[5] = rank_in(q.e[3], b);
    }
    // voxel offset of neighbour slot i (same order)
    ULK_D ULK_INLINE i32 nb_off(int i) const
    {
        const i32 sy = dm.W, sz = dm.W * dm.H;
        return i == 0 ? -sz : i == 1 ? -sy : i == 2 ? -1 : i == 3 ? 1 : i == 4 ? sy : sz;
    }
};
// direction (0 = x, 1 = y, 2 = z) of neighbour slot i
ULK_HD ULK_INLINE u32 ulk_nb_dir(int i) { return i == 0 || i == 5 ? 2u : (i == 1 || i == 4 ? 1u : 0u); }

// ------------------------------------------------------------------ union-find on ranks
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
// first lane of the chain of `linked` lanes that ends at `lane` (bit l of linked: lane l is
// joined to lane l-1)
ULK_D ULK_INLINE int ulk_run_start(u32 linked, int lane)
{
    u32 stop = ~linked & (0xffffffffu >> (31 - lane));   // lanes <= lane that start a run
    return 31 - __clz(stop);                               // lane 0 never has its link bit set
}
#endif
// same for bit positions of a 32-voxel segment word (host + device)
ULK_D ULK_INLINE int ulk_bit_run_start(u32 word, int b)
{
    const u32 linked = word & (word << 1);
    u32 stop = ~linked & (0xffffffffu >> (31 - b));
    int p = 31;
    while (!((stop >> p) & 1u)) --p;
    return p;
}

// ------------------------------------------------------------------ K1: threshold -> foreground words
// processing.py:212 `foreground[time] > threshold`, one bit per voxel, one 32-bit word per
// 32-voxel row segment.  Nothing else is written for the frame's V voxels.  (The functor form
// is what the emulation and unusual dtypes run; ulk_seg.cu has the vectorised streaming kernels.)
struct K_Threshold {
    const void *fg; int fgt; double thr; UlkDims dm; i32 S; u32 *words;
    ULK_D void operator()(i64 t) const
    {
        const i32 row = (i32)(t / S), sg = (i32)(t - (i64)row * S);
        const i64 base = (i64)row * dm.W + (i64)sg * 32;
        u32 w = 0;
        for (int l = 0; l < 32 && sg * 32 + l < dm.W; ++l)
            if (ulk_load_as_double(fg, fgt, base + l) > thr) w |= 1u << l;
        words[t] = w;
    }
};
// Expansion of the words into rank space: the foreground list, the contour value of every
// foreground voxel (float32, or float64 for wide dtypes: vendored/hierarchy.py:158-159,350 up-
// casts), the component forest with every x-run of a segment already hooked to its first voxel,
// the interleaved (bits, first rank) records, and max |contour| over the foreground.
struct K_Expand {
    const u32 *words; const i32 *off; UlkDims dm; i32 S;
    const void *ct; int ctt;
    UlkOW *ow; i32 *fgl; float *ctf; double *ctd; i32 *P; u32 *cmax;
    ULK_D ULK_INLINE u32 emit(i32 r, i32 v, i32 head) const
    {
        fgl[r] = v;
        P[r] = head;
        u32 bits = 0;
        if (ctt == ULK_F32) {
            const float c = ((const float *)ct)[v];
            ctf[r] = c;
            const float a = fabsf(c);
            if (a <= 3.0e38f) memcpy(&bits, &a, 4);
        } else {
            const double c = ulk_load_as_double(ct, ctt, v);
            if (ctf) ctf[r] = (float)c; else ctd[r] = c;
            const float a = (float)(c < 0 ? -c : c);
            if (a == a && a <= 3.0e38f) memcpy(&bits, &a, 4);
        }
        return bits;
    }
    ULK_D void operator()(i64 t) const     // one segment per thread (emulation)
    {
        const u32 w = words[t];
        const i32 o = off[t];
        UlkOW e;
        e.w = w; e.o = o;
        ow[t] = e;
        const i32 row = (i32)(t / S), sg = (i32)(t - (i64)row * S);
        const i32 base = row * dm.W + sg * 32;
        u32 mx = 0;
        for (u32 m = w; m; m &= m - 1) {
            const int b = ulk_ffs(m) - 1;
            const i32 r = o + ulk_popc(w & ((1u << b) - 1u));
            const int b0 = ulk_bit_run_start(w, b);
            const u32 bits = emit(r, base + b, o + ulk_popc(w & ((1u << b0) - 1u)));
            mx = bits > mx ? bits : mx;
        }
        if (mx) ulk_atomic_max(cmax, mx);
    }
};

// ------------------------------------------------------------------ connected components (ndi.label)
// One thread per 32-voxel row segment.  Every x-run inside a segment is already hooked to its
// first voxel (K_Expand), so what is left is one union per run of overlap with the row above /
// the slice above (start bits of w & w_up) and one across the segment boundary.  Size counters
// of all run starts (the only possible roots) are zeroed.
struct K_CclWords {
    UlkFg fg; UlkFastDiv fS, fH; i32 *P; i32 *cnt;
    ULK_D void operator()(i64 tt) const
    {
        const i32 t = (i32)tt;
        const UlkOW e = fg.load(t);
        const u32 w = e.w;
        if (w == 0) return;
        const u32 r = fS.div((u32)t), sg = (u32)t - r * (u32)fg.S;
        const u32 zz = fH.div(r), y = r - zz * (u32)fg.dm.H;
        for (u32 m = w & ~(w << 1); m; m &= m - 1) cnt[e.o + ulk_popc(w & ((m & (0u - m)) - 1u))] = 0;
        if (sg > 0 && (w & 1u)) {
            const UlkOW l = fg.load(t - 1);
            if (l.w >> 31) ulk_vunion(P, e.o, e.o - 1);
        }
        if (y > 0) {
            const UlkOW u = fg.load(t - fg.S);
            const u32 o = w & u.w;
            for (u32 m = o & ~(o << 1); m; m &= m - 1) {
                const u32 below = (m & (0u - m)) - 1u;
                ulk_vunion(P, e.o + ulk_popc(w & below), u.o + ulk_popc(u.w & below));
            }
        }
        if (zz > 0) {
            const UlkOW u = fg.load(t - fg.SH);
            const u32 o = w & u.w;
            for (u32 m = o & ~(o << 1); m; m &= m - 1) {
                const u32 below = (m & (0u - m)) - 1u;
                ulk_vunion(P, e.o + ulk_popc(w & below), u.o + ulk_popc(u.w & below));
            }
        }
    }
};
// flatten, component sizes (one atomic per group of lanes with the same root) and root list
struct K_FlattenCount {
    i32 *P; i32 *cnt; i32 *roots; i32 *nroots;
#ifdef ULK_EMU
    void operator()(i32 r) const
    {
        i32 root = ulk_vfind(P, r);
        P[r] = root;
        cnt[root] += 1;
        if (root == r) roots[(*nroots)++] = r;
    }
#else
    ULK_D void multi(const i32 (&v)[ULK_MULTI], const bool (&ok)[ULK_MULTI]) const
    {
        i32 r[ULK_MULTI];
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) r[u] = ok[u] ? P[v[u]] : -1;
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
// code at the root of every component.  final == 0: the 1-based index among the kept components
// (0: dropped).  final == 1 (some component is above max_area): the label K_LabelCc mode 1 writes,
// so that the voxel pass needs two dependent loads instead of three.
struct K_RootCode {
    const u32 *roots; const i32 *keep, *keeppre, *big, *bigpre; i32 *code; i32 *cc_big; i32 *big_cc; int final;
    ULK_D void operator()(i64 i) const
    {
        u32 r = roots[i];
        if (!keep[i]) { code[r] = 0; return; }
        i32 c = keeppre[i];
        cc_big[c] = big[i] ? bigpre[i] : -1;
        if (big[i]) big_cc[bigpre[i]] = c;
        code[r] = !final ? c + 1 : (big[i] ? bigpre[i] + 1 : -c - 2);
    }
};
// Component labels, written in place over the flattened component forest.
// mode 0 (no component above max_area): the 1-based component index, 0 if dropped.
// mode 1: components above max_area get their 1-based index among themselves (the region labels
// of the oversegmentation pass), the other kept components -(index + 2), dropped ones -1.
// (mode 1 reads the final label from `code`, see K_RootCode; -1 never is one)
struct K_LabelCc {
    const i32 *code; int mode; i32 *lab;
    i32 *biglist; i32 *nbig;       // mode 1: the ranks of the voxels of the components above max_area (32-entry chunks in raster order)
    i32 *po;                       // steepest-edge endpoints: voxels of dropped components leave the graph (ULK_PO_OUT)
    struct S1 { i32 r, root; };
    struct S2 { i32 r, c; };
    ULK_D ULK_INLINE S1 load1(i32 r, bool ok) const { S1 s; s.r = r; s.root = ok ? lab[r] : -1; return s; }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const { S2 s; s.r = a.r; s.c = a.root >= 0 ? code[a.root] : -1; return s; }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const
    {
        const i32 c = s.c;
        i32 g = -1;
        if (c == 0) po[s.r] = ULK_PO_OUT;
        if (c != -1) {
            if (mode == 0) lab[s.r] = c;
            else {
                lab[s.r] = c == 0 ? -1 : c;
                g = c > 0 ? c - 1 : -1;
            }
        }
        if (mode == 1) {
#ifdef ULK_EMU
            if (g >= 0) biglist[(*nbig)++] = s.r;
#else
            ulk_warp_append(g >= 0, s.r, biglist, nbig);
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
        for (int u = 0; u < ULK_MULTI; ++u) g[u] = c[u] > 0 ? c[u] - 1 : -1;
        const int lane = threadIdx.x & 31;
        u32 m[ULK_MULTI];
        i32 total = 0;
#pragma unroll
        for (int u = 0; u < ULK_MULTI; ++u) {
            if (c[u] != -1) lab[v[u]] = c[u] == 0 ? -1 : c[u];
            if (c[u] == 0) po[v[u]] = ULK_PO_OUT;
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

// ------------------------------------------------------------------ voxel stage: steepest edges
// For every voxel of the list the minimum incident edge by (weight, edge key); edge key =
// 3*source voxel + direction reproduces the enumeration order of edges_from_mask
// (vendored/graph.py:41-57), i.e. higra's stable tie-break.  Candidates are visited in
// increasing edge-key order (-z, -y, -x, +x, +y, +z), so a strict comparison of the weights alone
// yields the minimum of (weight, key).  Outputs: the weight (order-preserving bits), the key and
// the RANK of the other endpoint (ULK_PO_NONE / ULK_PO_OUT otherwise).
struct K_MinEdge {
    const i32 *lab; UlkCt ct; UlkFg fg;
    u32 *mw, *pk; i32 *po;
    struct S1 { i32 r, v, x; bool ok; UlkNbRaw q; };
    struct S2 { i32 r, v, l; bool ok; i32 rn[6], ln[6]; float c, cn[6]; };
    ULK_D ULK_INLINE S1 load1(i32 r, i32 v, i32 z, i32 y, i32 x, bool ok) const
    {
        S1 s;
        s.r = r; s.v = v; s.x = x; s.ok = ok;
        if (ok) s.q = fg.nb_load(r, fg.seg(z, y, x), z, y, true, true);
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool ok) const
    {
        S2 s;
        s.r = a.r; s.v = a.v; s.ok = ok; s.l = 0; s.c = 0.f;
        if (!ok) return s;
        fg.nb_ranks(a.q, a.r, a.v, a.x, s.rn);
        // lab == nullptr: component space (every foreground neighbour is in the same component)
        s.l = lab ? lab[a.r] : 1;
#ifndef ULK_EMU
#pragma unroll
#endif
        for (int i = 0; i < 6; ++i) s.ln[i] = s.rn[i] >= 0 ? (lab ? lab[s.rn[i]] : 1) : 0;
        if (ct.f) {
            s.c = ct.f[a.r];
#ifndef ULK_EMU
#pragma unroll
#endif
            for (int i = 0; i < 6; ++i) s.cn[i] = s.rn[i] >= 0 ? ct.f[s.rn[i]] : 0.f;
        }
        return s;
    }
    ULK_D ULK_INLINE void run(const S2 &s, bool ok) const
    {
        if (!ok) return;
        const i32 r = s.r, v = s.v, l = s.l;
        if (l <= 0) { po[r] = ULK_PO_OUT; return; }
        u32 bw = ULK_NONE32, bk = ULK_NONE32;
        i32 bo = ULK_PO_NONE;
#ifndef ULK_EMU
#pragma unroll
#endif
        for (int i = 0; i < 6; ++i) {
            if (s.rn[i] < 0 || s.ln[i] != l) continue;
            // candidates in increasing edge-key order: the first strictly smaller weight wins
            const float wf = ct.f ? ulk_wsum(i < 3 ? s.cn[i] + s.c : s.c + s.cn[i]) : ct.w(r, s.rn[i]);
            const u32 w = ulk_f2ord(wf);
            const i32 a = i < 3 ? v + fg.nb_off(i) : v;
            if (w < bw) { bw = w; bk = 3u * (u32)a + ulk_nb_dir(i); bo = s.rn[i]; }
        }
        mw[r] = bw;
        pk[r] = bk;
        po[r] = bo;
    }
};
// first Boruvka round: every voxel hooks along its steepest edge; mutual pairs (each is the
// other's steepest endpoint: one edge joins two neighbours, so that is the same edge) keep the
// smaller voxel as root.  state: 0 growing, 1 minimum (basin seed); the level of a root and its
// state only exist at roots.  Rank space only: no coordinates, no neighbour lookups.
struct K_Hook1 {
    const u32 *mw; const i32 *po;
    i32 *bp; u8 *state; u32 *Lr; u32 *best; i32 *nbasins;
    struct S1 { i32 r, o; };
    struct S2 { i32 r, o, oo; };
    ULK_D ULK_INLINE S1 load1(i32 r, bool ok) const { S1 s; s.r = r; s.o = ok ? po[r] : ULK_PO_OUT; return s; }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const { S2 s; s.r = a.r; s.o = a.o; s.oo = a.o >= 0 ? po[a.o] : -3; return s; }
    ULK_D ULK_INLINE void run(const S2 &s, bool ok) const
    {
        if (!ok) return;
        const i32 r = s.r;
        if (s.o == ULK_PO_OUT) { bp[r] = -1; return; }
        if (s.o == ULK_PO_NONE) {
            bp[r] = r; state[r] = 1; Lr[r] = ULK_NONE32; best[r] = ULK_NONE32;
            ulk_atomic_add(nbasins, 1);
            return;
        }
        if (s.oo == r && r < s.o) { bp[r] = r; state[r] = 0; Lr[r] = mw[r]; best[r] = ULK_NONE32; }
        else bp[r] = s.o;
    }
};
// path compression of the hook forest, fused with the lists of the flow rounds: growing
// roots, and candidates = voxels whose steepest weight equals their supervertex level
struct K_CompressCand {
    i32 *bp; const u32 *mw; const u32 *Lr; const u8 *state;
    i32 *cand; i32 *ncand; i32 *groots; i32 *ngroots; i32 *sroots; i32 *nsroots;
#ifdef ULK_EMU
    void operator()(i32 v) const
    {
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
    UlkCt ct; UlkFg fg;
    u32 *best; i32 *cand2; i32 *ncand2;
    ULK_D void operator()(i64 i) const
    {
        const i32 r = cand[i];
        const i32 root = ulk_vfind(bp, r);
        if (state[root] != 0 || mw[r] != Lr[root]) return;
        cand2[ulk_atomic_add(ncand2, 1)] = r;
        const i32 l = lab[r];
        const u32 lv = Lr[root];
        const i32 v = fg.fgl[r];
        i32 z, y, x, rn[6];
        fg.dm.coords(v, z, y, x);
        fg.nb_ranks(fg.nb_load(r, fg.seg(z, y, x), z, y, true, true), r, v, x, rn);
        u32 b = ULK_NONE32;
        for (int j = 0; j < 6; ++j) {
            const i32 u = rn[j];
            if (u < 0 || lab[u] != l || ulk_f2ord(ct.w(r, u)) != lv || ulk_vfind(bp, u) == root) continue;
            const u32 k = 3u * (u32)(j < 3 ? v + fg.nb_off(j) : v) + ulk_nb_dir(j);
            if (k < b) b = k;
        }
        if (b != ULK_NONE32) ulk_atomic_min(&best[root], b);
    }
};
// decide, read-only part: target root of every growing supervertex (-1: becomes a minimum,
// -2: stays root as the winner of a mutual pair)
struct K_Decide1 {
    const i32 *groots; const i32 *bp; const u32 *best; UlkFg fg;
    i32 *tgt; i32 *src_rank;       // src_rank: rank of the source voxel of the chosen edge
    ULK_D void operator()(i64 i) const
    {
        const i32 r = groots[i];
        const u32 k = best[r];
        if (k == ULK_NONE32) { tgt[i] = -1; return; }
        const i32 a = (i32)(k / 3u), b = a + fg.dm.stride((int)(k % 3u));
        const i32 qa = fg.rank_of(a), qb = fg.rank_of(b);
        src_rank[i] = qa;
        const i32 ra = ulk_vfind(bp, qa), rb = ulk_vfind(bp, qb);
        const i32 t = ra == r ? rb : ra;
        if (best[t] == k && r < t) { tgt[i] = -2; return; }
        tgt[i] = t;
    }
};
struct K_Decide2 {
    const i32 *groots; const i32 *tgt; const i32 *src_rank;
    i32 *bp; u8 *state; u32 *best; u8 *fmark;
    i32 *groots2; i32 *ngroots2;
    i32 *flow_root; u32 *flow_key; i32 *flow_src; i32 *nflow; i32 *nbasins;
    ULK_D void operator()(i64 i) const
    {
        const i32 r = groots[i];
        const i32 t = tgt[i];
        const u32 k = best[r];
        best[r] = ULK_NONE32;
        if (t == -1) { state[r] = 1; ulk_atomic_add(nbasins, 1); return; }
        if (t == -2) { groots2[ulk_atomic_add(ngroots2, 1)] = r; return; }
        bp[r] = t;
        const i32 j = ulk_atomic_add(nflow, 1);
        const u32 a = (u32)src_rank[i], d = k % 3u;
        flow_root[j] = r; flow_key[j] = k; flow_src[j] = (i32)a;
        // spanning-tree mark of the flow edge at its source voxel, one bit per direction
        ulk_atomic_or((u32 *)(fmark + (a & ~3u)), (1u << d) << (8u * (a & 3u)));
    }
};

// ------------------------------------------------------------------ basins
// second path compression; every basin root appends (region label, root rank): sorting
// these keys numbers the basins grouped by region, in raster order of their roots
struct K_CompressRoots {
    i32 *bp; const i32 *lab; u64 *rootkeys; i32 *nroots;
    ULK_D void operator()(i32 v) const
    {
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
// bp[r] is the supervertex root of r (K_CompressCand) and bp[root] the basin root
// (K_CompressRoots on the root list): two hops, then the basin id (-1 outside the regions)
struct K_AssignBasin {
    const i32 *bp; const u32 *gid_at_root;
    i32 *bas;
    struct S1 { i32 r, p; };
    struct S2 { i32 r, rr; };
    ULK_D ULK_INLINE S1 load1(i32 r, bool ok) const { S1 s; s.r = ok ? r : -1; s.p = ok ? bp[r] : -1; return s; }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const { S2 s; s.r = a.r; s.rr = a.p < 0 ? -1 : bp[a.p]; return s; }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const { if (s.r >= 0) bas[s.r] = s.rr < 0 ? -1 : (i32)gid_at_root[s.rr]; }
};
// flow edges are spanning-tree edges of the basin of their supervertex
struct K_FlowTop {
    const i32 *flow_root; const u32 *flow_key; const i32 *flow_src; const i32 *bas; UlkCt ct; UlkFg fg; u64 *b_top; u32 *Tk;
    ULK_D void operator()(i64 j) const
    {
        const u32 fk = flow_key[j];
        const i32 a = (i32)(fk / 3u), e = a + fg.dm.stride((int)(fk % 3u));
        const i32 b = bas[flow_root[j]];
        ulk_atomic_max(&b_top[b], ((u64)ulk_f2ord(ct.w(flow_src[j], fg.rank_of(e))) << 32) | fk);
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
#define ULK_K3_WARPS 8
#ifndef ULK_EMU
// the voxels (ranks) of the basins that need the check, one warp per basin scanning its bounding
// box row by row (x-runs of a basin stay contiguous in the list, which is all K_K3Init relies on)
struct K_K3Collect {
    const i32 *needlist; i64 nneed; const i32 *bas; const i32 *bbmin, *bbmax; UlkFg fg;
    i32 *list; i32 *count;
    // ULK_K3_WARPS warps share a basin (slices z0 + w, z0 + w + ULK_K3_WARPS, ...): the basins that
    // need the check are few (2 % at C5), one warp each leaves most of the GPU idle
    ULK_D void warp(i64 ii, bool) const
    {
        const int lane = threadIdx.x & 31;
        const i64 i = ii / (32 * ULK_K3_WARPS);
        if (i >= nneed) return;
        const i32 sub = (i32)((ii >> 5) % ULK_K3_WARPS);
        const i32 b = needlist[i];
        const i32 z0 = bbmin[3 * b], y0 = bbmin[3 * b + 1], x0 = bbmin[3 * b + 2];
        const i32 z1 = bbmax[3 * b], y1 = bbmax[3 * b + 1], x1 = bbmax[3 * b + 2];
        for (i32 z = z0 + sub; z <= z1; z += ULK_K3_WARPS)
            for (i32 y = y0; y <= y1; ++y) {
                const i32 trow = (z * fg.dm.H + y) * fg.S;
                for (i32 xb = x0; xb <= x1; xb += 32) {
                    const i32 x = xb + lane;
                    i32 r = -1;
                    if (x <= x1) r = UlkFg::rank_in(fg.load(trow + (x >> 5)), x & 31);
                    const bool hit = r >= 0 && bas[r] == b;
                    ulk_warp_append(hit, r, list, count);
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
// "Same region" is decided on basins (g_lab[basin] = region label): no label volume is read.
struct K_BasinAccum {
    const i32 *bas; const u32 *mw, *pk; const i32 *po; UlkFg fg;
    i32 *b_size; u64 *b_top; u64 *ncross;
    int want_stats; i32 *bbmin, *bbmax; i64 *csum;
    int want_join; const u32 *b_L; u32 *jk; i32 *hp; i32 *nhp;
    int want_k0; u32 *K0; u32 *Tk;      // Tk: largest raster key among the basin's steepest edges
    i32 *xl; i32 *nxl;                  // ranks of voxels with a forward edge into another basin of their region
    const i32 *g_lab;                   // region label of a basin
    // phase 1: the voxel's own entries and what its neighbours' ranks depend on
    struct S1 { i32 r, v, z, y, x, b, o; u32 m, k; UlkNbRaw q; };
    // phase 2: the neighbours' basins, the basin level, the weight across the steepest edge
    struct S2 { i32 r, v, z, y, x, b; u64 top; u32 nc, kmin, jkv; bool plateau; };
    // phase 0: what the rank alone addresses (requested together with the voxel index)
    struct S0 { i32 b, o; u32 m, k; };
    ULK_D ULK_INLINE S0 load0(i32 r, bool ok) const
    {
        S0 p;
        p.b = bas[r]; p.m = mw[r]; p.k = pk[r]; p.o = po[r];
        if (!ok) { p.b = -1; p.o = ULK_PO_OUT; p.m = 0; p.k = ULK_NONE32; }
        return p;
    }
    ULK_D ULK_INLINE S1 load1(const S0 &p, i32 r, i32 v, i32 z, i32 y, i32 x, bool ok) const
    {
        S1 s;
        s.r = r; s.v = v; s.z = z; s.y = y; s.x = x; s.b = p.b; s.o = p.o; s.m = p.m; s.k = p.k;
        if (ok) s.q = fg.nb_load(r, fg.seg(z, y, x), z, y, want_k0 != 0, true);
        else {
            for (int i = 0; i < 4; ++i) { s.q.e[i].w = 0; s.q.e[i].o = 0; }
            s.q.vl = -1; s.q.vr = -1;
        }
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const
    {
        S2 o;
        o.r = a.r; o.v = a.v; o.z = a.z; o.y = a.y; o.x = a.x; o.b = a.b;
        o.top = 0; o.nc = 0; o.kmin = ULK_NONE32; o.jkv = ULK_JK_CORE; o.plateau = false;
        const i32 b = a.b, v = a.v;
        if (b < 0) return o;
        i32 rn[6], nb[6];
        fg.nb_ranks(a.q, a.r, v, a.x, rn);
#ifndef ULK_EMU
#pragma unroll
#endif
        for (int i = 0; i < 6; ++i) nb[i] = rn[i] >= 0 ? bas[rn[i]] : -1;
        u32 bl = 0, mo = 0;
        const bool has_edge = a.o >= 0;
        if (want_join) { bl = b_L[b]; if (has_edge) mo = mw[a.o]; }
        const bool q0 = nb[3] >= 0 && nb[3] != b, q1 = nb[4] >= 0 && nb[4] != b, q2 = nb[5] >= 0 && nb[5] != b;
        if (q0 || q1 || q2) {
            const i32 lown = g_lab[b];
            if (q0 && g_lab[nb[3]] == lown) ++o.nc;
            if (q1 && g_lab[nb[4]] == lown) ++o.nc;
            if (q2 && g_lab[nb[5]] == lown) ++o.nc;
        }
        o.top = has_edge ? ((u64)a.m << 32) | a.k : 0;
        const i32 sy = fg.dm.W, sz = fg.dm.W * fg.dm.H;
        u32 kk;
        if (nb[3] == b) { kk = 3u * (u32)v; if (kk < o.kmin) o.kmin = kk; }
        if (nb[4] == b) { kk = 3u * (u32)v + 1u; if (kk < o.kmin) o.kmin = kk; }
        if (nb[5] == b) { kk = 3u * (u32)v + 2u; if (kk < o.kmin) o.kmin = kk; }
        if (want_k0) {
            if (nb[2] == b) { kk = 3u * (u32)(v - 1); if (kk < o.kmin) o.kmin = kk; }
            if (nb[1] == b) { kk = 3u * (u32)(v - sy) + 1u; if (kk < o.kmin) o.kmin = kk; }
            if (nb[0] == b) { kk = 3u * (u32)(v - sz) + 2u; if (kk < o.kmin) o.kmin = kk; }
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
        if (want_join) jk[o.r] = o.jkv;
        b_size[b] += 1;
        if (o.top > b_top[b]) b_top[b] = o.top;
        *ncross += o.nc;
        if (o.nc) xl[(*nxl)++] = o.r;
        if (o.plateau) hp[(*nhp)++] = o.r;
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
        if (nct) ulk_warp_append(o.nc > 0, o.r, xl, nxl);
        if (want_join) {
            if (b >= 0) jk[o.r] = o.jkv;
            ulk_warp_append(o.plateau, o.r, hp, nhp);
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
    const i32 *hp; const i32 *lab; const u32 *mw, *pk; const u8 *fmark; UlkCt ct; UlkFg fg;
    u32 *jk; i32 *changed;
    ULK_D void operator()(i64 i) const
    {
        const i32 r = hp[i];
        const i32 l = lab[r];
        const u32 lv = mw[r];
        const u32 cur = jk[r];
        const i32 v = fg.fgl[r];
        i32 z, y, x, rn[6];
        fg.dm.coords(v, z, y, x);
        fg.nb_ranks(fg.nb_load(r, fg.seg(z, y, x), z, y, true, true), r, v, x, rn);
        const u32 pkv = pk[r];
        u32 bestc = cur;
        for (int j = 0; j < 6; ++j) {
            const i32 u = rn[j];
            if (u < 0 || lab[u] != l) continue;
            const u32 d = ulk_nb_dir(j);
            const i32 srcv = j < 3 ? v + fg.nb_off(j) : v, srcr = j < 3 ? u : r;
            const u32 k = 3u * (u32)srcv + d;
            const bool tree = pkv == k || pk[u] == k || ((fmark[srcr] >> d) & 1);
            if (!tree) continue;
            if (ulk_f2ord(ct.w(r, u)) != lv) continue;
            u32 c;
            if (mw[u] < lv) c = k;
            else { u32 ju = jk[u]; c = ju > k ? ju : k; }
            if (c < bestc) bestc = c;
        }
        if (bestc < cur) { jk[r] = bestc; *changed = 1; }
    }
};
// late joiners climb the basin BPT: a voxel attaches to the highest ancestor of its basin
// whose merge precedes its join key
// one 16-byte record per basin: where its leaf sits in the region's tree and the first climb step
// (parent p, its merge key and its own parent pp: most voxels stop at that first comparison)
struct alignas(16) UlkAttachInfo { i32 o, node, p, pp; u64 pkey; u64 pad; };
struct K_AttachInfo {
    const i32 *b_lab; const i32 *bs; const i32 *Ipar; const u64 *Iwk; UlkAttachInfo *info;
    ULK_D void operator()(i64 b) const
    {
        const i32 st = bs[b_lab[b] - 1];
        UlkAttachInfo r;
        r.o = 2 * st; r.node = (i32)b - st; r.p = Ipar[r.o + r.node];
        r.pp = Ipar[r.o + r.p]; r.pkey = Iwk[r.o + r.p]; r.pad = 0;
        info[b] = r;
    }
};
struct K_Attach {
    const i32 *bas; const u32 *mw, *jk; const UlkAttachInfo *info;
    const i32 *Ipar; const u64 *Iwk; i32 *attach;
    u64 *awhi, *awlo; double fixscale;      // volume attribute only (else nullptr): weight sums per attach node
    struct S1 { i32 b; u32 j, m; };
    struct S2 { i32 o, node, p, pp; u64 J, pkey; u32 m; bool climb; };
    ULK_D ULK_INLINE S1 load1(i32 r, bool ok) const
    {
        S1 s;
        s.b = -1; s.j = ULK_JK_CORE; s.m = 0;
        if (ok) { s.b = bas[r]; s.j = jk[r]; s.m = mw[r]; }
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const
    {
        S2 s;
        s.o = 0; s.node = -1; s.p = -1; s.pp = -1; s.J = 0; s.pkey = 0; s.m = a.m; s.climb = false;
        if (a.b < 0) return s;
        const bool late = a.j != ULK_JK_CORE;
        if (!late && !awhi) return s;
#ifdef ULK_EMU
        const UlkAttachInfo q = info[a.b];
#else
        const int4 raw = __ldg((const int4 *)info + 2 * (i64)a.b);
        UlkAttachInfo q;
        q.o = raw.x; q.node = raw.y; q.p = raw.z; q.pp = raw.w; q.pkey = 0;
        if (late) q.pkey = __ldg((const unsigned long long *)info + 4 * (i64)a.b + 2);
#endif
        s.o = q.o;
        s.node = q.node;
        if (late) {
            s.climb = true;
            s.J = ((u64)a.m << 32) | a.j;
            s.p = q.p; s.pp = q.pp; s.pkey = q.pkey;
        }
        return s;
    }
    // the node a late joiner attaches to: the highest ancestor of its basin whose merge precedes J
    ULK_D ULK_INLINE i32 climb(const S2 &s) const
    {
        const i32 o = s.o;
        i32 node = s.node, p = s.p;
        // the first step comes with the basin record; further up the parent's merge key and the
        // grandparent index are loaded together: one memory round trip per level of the climb
        if (p != node && s.pkey < s.J) {
            node = p;
            p = s.pp;
            while (p != node) {
                u64 w = Iwk[o + p];
                i32 pp = Ipar[o + p];
                if (!(w < s.J)) break;
                node = p;
                p = pp;
            }
        }
        return node;
    }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const
    {
        const i32 o = s.o, start = s.node;
        i32 node = start;
        if (s.climb) node = climb(s);
#ifdef ULK_EMU
        if (node != start) { attach[o + start] -= 1; attach[o + node] += 1; }
#else
        {   // neighbouring lanes (consecutive voxels of one basin) move between the same two counters:
            // one pair of atomics per run of equal (from, to); every thread of the grid gets here
            const int ln = threadIdx.x & 31;
            const i32 from = node != start ? o + start : -1 - ln, to = node != start ? o + node : -1;
            i32 c = node != start ? 1 : 0;
            const i32 fprev = __shfl_up_sync(0xffffffffu, from, 1), tprev = __shfl_up_sync(0xffffffffu, to, 1);
            const u32 rn = ulk_run_mask(__ballot_sync(0xffffffffu, ln == 0 || fprev != from || tprev != to), ln);
            const int f0 = __ffs(rn) - 1, l0 = 31 - __clz(rn);
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const i32 c2 = __shfl_down_sync(0xffffffffu, c, off);
                if (ln + off <= l0) c += c2;
            }
            if (c && ln == f0) { atomicAdd(&attach[from], -c); atomicAdd(&attach[to], c); }
        }
#endif
        if (!awhi) return;          // uniform over the grid
        // volume: the voxel's steepest-edge weight goes to the node it is attached to
        const bool has = node >= 0 && s.m != ULK_NONE32;
        const i64 fx = has ? (i64)((double)ulk_ord2f(s.m) * fixscale) : 0;
        u64 hi = (u64)(fx >> 32), lo = (u64)(fx & 0xffffffffll);
#ifdef ULK_EMU
        if (has) { awhi[o + node] += hi; awlo[o + node] += lo; }
#else
        // lanes of one basin's core share the node: segmented shuffle sum towards the run's first lane
        const int lane = threadIdx.x & 31;
        const i32 key = has ? o + node : -1 - lane;
        const i32 kprev = __shfl_up_sync(0xffffffffu, key, 1);
        const u32 run = ulk_run_mask(__ballot_sync(0xffffffffu, lane == 0 || kprev != key), lane);
        const int first = __ffs(run) - 1, last = 31 - __clz(run);
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const u64 h2 = __shfl_down_sync(0xffffffffu, hi, off), l2 = __shfl_down_sync(0xffffffffu, lo, off);
            if (lane + off <= last) { hi += h2; lo += l2; }
        }
        if (has && lane == first) {
            atomicAdd((unsigned long long *)&awhi[key], (unsigned long long)hi);
            atomicAdd((unsigned long long *)&awlo[key], (unsigned long long)lo);
        }
#endif
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
    const i32 *bas; const i32 *g_lab; UlkCt ct; UlkFg fg; UlkRagTable T; double fixscale;
    struct S1 { i32 r, v, x, b; UlkNbRaw q; };
    struct S2 { i32 r, v, b; i32 cr[3], cb[3]; };
    ULK_D ULK_INLINE S1 load1(i32 r, i32 v, i32 z, i32 y, i32 x, bool ok) const
    {
        S1 s;
        s.r = r; s.v = v; s.x = x; s.b = -1;
        if (!ok) return s;
        s.b = bas[r];
        s.q = fg.nb_load(r, fg.seg(z, y, x), z, y, false, true);
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const
    {
        S2 s;
        s.r = a.r; s.v = a.v; s.b = a.b;
        s.cr[0] = s.cr[1] = s.cr[2] = -1; s.cb[0] = s.cb[1] = s.cb[2] = -1;
        if (a.b < 0) return s;
        i32 rn[6];
        fg.nb_ranks(a.q, a.r, a.v, a.x, rn);
        for (int d = 0; d < 3; ++d) { s.cr[d] = rn[3 + d]; s.cb[d] = rn[3 + d] >= 0 ? bas[rn[3 + d]] : -1; }
        return s;
    }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const
    {
        const i32 v = s.v, b = s.b;
        if (b < 0) return;
        const i32 l = g_lab[b];
        for (int d = 0; d < 3; ++d) {
            const i32 c = s.cb[d];
            if (c < 0 || c == b || g_lab[c] != l) continue;
            u64 key = b < c ? ((u64)(u32)b << 32) | (u32)c : ((u64)(u32)c << 32) | (u32)b;
            float w = ct.w(s.r, s.cr[d]);
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
// Runs of x edges (key <= K0) between voxels that are neighbours in the list and sit in one warp
// are hooked to their first voxel at initialisation; K_K3Union joins the runs (same scheme as
// the component labelling).
struct K_K3Init {
    const i32 *bas; const u32 *K0; const i32 *fgl; i32 *P;
    struct S1 { i32 r, v, b, bl; bool ok, cons; };     // cons: the previous lane holds voxel v-1 of the same row
    struct S2 { i32 r, b; bool ok, link; };
    ULK_D ULK_INLINE S1 load1(i32 r, i32 v, i32, i32, i32 x, bool ok) const
    {
        S1 s;
        s.r = r; s.v = v; s.ok = ok; s.b = -1; s.bl = -1;
#ifdef ULK_EMU
        s.cons = x > 0 && (x & 31) != 0 && r > 0 && fgl[r - 1] == v - 1;
#else
        i32 vprev = __shfl_up_sync(0xffffffffu, v, 1);
        s.cons = (threadIdx.x & 31) > 0 && x > 0 && vprev == v - 1;
#endif
        if (ok) { s.b = bas[r]; s.bl = s.cons ? bas[r - 1] : -1; }
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const
    {
        S2 s;
        s.r = a.r; s.b = a.b; s.ok = a.ok;
        s.link = a.ok && a.cons && a.b >= 0 && a.bl == a.b && 3u * (u32)(a.v - 1) <= K0[a.b];
        return s;
    }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const
    {
#ifdef ULK_EMU
        P[s.r] = s.b < 0 ? -1 : (s.link ? P[s.r - 1] : s.r);
#else
        const int lane = threadIdx.x & 31;
        u32 linked = __ballot_sync(0xffffffffu, s.link);
        if (s.ok) P[s.r] = s.b < 0 ? -1 : s.r - (lane - ulk_run_start(linked, lane));
#endif
    }
};
struct K_K3Union {
    const i32 *bas; UlkFg fg; const u32 *K0; i32 *P;
    struct S1 { i32 r, v, x, b; bool boundary; UlkNbRaw q; };
    struct S2 { i32 r; i32 ul, uy, uz; };      // ranks to union with (-1: none)
    ULK_D ULK_INLINE S1 load1(i32 r, i32 v, i32 z, i32 y, i32 x, bool ok) const
    {
        S1 s;
        s.r = r; s.v = v; s.x = x; s.b = -1;
#ifdef ULK_EMU
        s.boundary = (x & 31) == 0;
#else
        i32 vprev = __shfl_up_sync(0xffffffffu, v, 1);
        s.boundary = (threadIdx.x & 31) == 0 || vprev != v - 1;
#endif
        if (!ok) return s;
        s.b = bas[r];
        s.q = fg.nb_load(r, fg.seg(z, y, x), z, y, true, false);
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool) const
    {
        S2 s;
        s.r = a.r; s.ul = s.uy = s.uz = -1;
        const i32 b = a.b, v = a.v;
        if (b < 0) return s;
        i32 rn[6];
        fg.nb_ranks(a.q, a.r, v, a.x, rn);
        const u32 kmax = K0[b];
        const i32 uy = v - fg.dm.W, uz = v - fg.dm.W * fg.dm.H;
        const bool left = rn[2] >= 0 && bas[rn[2]] == b && 3u * (u32)(v - 1) <= kmax;
        if (left && a.boundary) s.ul = rn[2];
        // (u-1,u), (u-1,v-1) carry smaller keys than (u,v): with (v-1,v) they already join u and v.
        // u-1 is the voxel of rank(u)-1 when bit (x-1) of u's segment is set
        const int bb = a.x & 31;
        if (rn[1] >= 0 && bas[rn[1]] == b && 3u * (u32)uy + 1u <= kmax) {
            const bool upleft = left && bb > 0 && ((a.q.e[1].w >> (bb - 1)) & 1u) && bas[rn[1] - 1] == b;
            if (!upleft) s.uy = rn[1];
        }
        if (rn[0] >= 0 && bas[rn[0]] == b && 3u * (u32)uz + 2u <= kmax) {
            const bool upleft = left && bb > 0 && ((a.q.e[0].w >> (bb - 1)) & 1u) && bas[rn[0] - 1] == b;
            if (!upleft) s.uz = rn[0];
        }
        return s;
    }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const
    {
        if (s.ul >= 0) ulk_vunion(P, s.r, s.ul);
        if (s.uy >= 0) ulk_vunion(P, s.r, s.uy);
        if (s.uz >= 0) ulk_vunion(P, s.r, s.uz);
    }
};
struct K_K3FlattenCount {
    const i32 *bas; i32 *P; i32 *nroot; i32 *nmulti;
    ULK_D void operator()(i32 v) const
    {
        if (P[v] < 0) return;
        i32 r = ulk_vfind(P, v);
        P[v] = r;
        if (r != v) return;
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
    const i32 *bas; UlkFg fg; const i32 *nroot; const i32 *P; u32 *best;
    ULK_D void operator()(i32 r, i32 v, i32 z, i32 y, i32 x) const
    {
        const i32 b = bas[r];
        if (b < 0 || nroot[b] <= 1) return;
        const i32 root = ulk_vfind(P, r);
        i32 rn[6];
        fg.nb_ranks(fg.nb_load(r, fg.seg(z, y, x), z, y, true, true), r, v, x, rn);
        u32 k = ULK_NONE32;
        for (int j = 0; j < 6; ++j) {
            const i32 u = rn[j];
            if (u < 0 || bas[u] != b || ulk_vfind(P, u) == root) continue;
            const u32 kk = 3u * (u32)(j < 3 ? v + fg.nb_off(j) : v) + ulk_nb_dir(j);
            if (kk < k) k = kk;
        }
        if (k != ULK_NONE32) ulk_atomic_min(&best[root], k);
    }
};
struct K_K3Hook {
    const i32 *bas; UlkFg fg; i32 *P; u32 *best; u32 *K0;
    ULK_D void operator()(i32 r, i32, i32, i32, i32) const
    {
        if (bas[r] < 0) return;
        u32 k = best[r];
        if (k == ULK_NONE32) return;
        best[r] = ULK_NONE32;
        const i32 a = (i32)(k / 3u), c = a + fg.dm.stride((int)(k % 3u));
        ulk_vunion(P, fg.rank_of(a), fg.rank_of(c));
        ulk_atomic_max(&K0[bas[r]], k);
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
    i32 *lab; i32 *po;
    struct S1 { i32 r, l, ba; };
    struct S2 { i32 r, rec; };       // rec: index into label_of_rec, -1: label 0, -2: nothing to write
    ULK_D ULK_INLINE S1 load1(i32 r, bool ok) const
    {
        S1 s;
        s.r = r; s.l = 0; s.ba = 0;
        // basA is read whatever the label says (it is only used for l > 0): a load that waits for
        // the label would serialise the entries of a thread
        if (ok) { s.l = lab[r]; s.ba = basA[r]; }
        return s;
    }
    ULK_D ULK_INLINE S2 load2(const S1 &a, bool ok) const
    {
        S2 s;
        s.r = a.r;
        if (!ok) s.rec = -2;
        else if (a.l > 0) s.rec = ncc + region_of[a.ba];
        else if (a.l <= -2) s.rec = -a.l - 2;
        else s.rec = -1;
        return s;
    }
    ULK_D ULK_INLINE void run(const S2 &s, bool) const
    {
        if (s.rec == -2) return;
        const i32 l = s.rec >= 0 ? label_of_rec[s.rec] : 0;
        lab[s.r] = l;
        if (l <= 0) po[s.r] = ULK_PO_OUT;
    }
};
struct K_ScatterLabels {    // region labels as the caller sees them (dense volume, 0 on the background: cleared before)
    const i32 *lab; const i32 *fgl; i32 *out;
    ULK_D void operator()(i64 r) const { i32 l = lab[r]; out[fgl[r]] = l > 0 ? l : 0; }
};
template <class T> struct K_FillAt {
    const i32 *list; T *p; T val;
    ULK_D void operator()(i64 i) const { p[list[i]] = val; }
};
