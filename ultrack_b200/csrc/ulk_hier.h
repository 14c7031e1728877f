// ulk_hier.h -- hierarchy construction at watershed-basin scale.
//
// After the voxel-scale kernels have collapsed every region (a "hierarchy" in
// ultrack's vocabulary, core/segmentation/hierarchy.py:89-90) into its watershed
// basins and their region-adjacency graph (RAG), the tie-sensitive part of higra's
// pipeline runs here on b basins / r RAG edges instead of n voxels / 3n edges:
//
//   Kruskal #1  (w, edge key)        -> BPT over basins (the "interesting" merges of
//                                       hg.bpt_canonical, vendored/hierarchy.py:169)
//   attribute   area | dynamics      -> new MST weights (hg.watershed_hierarchy_by_*)
//   Kruskal #2  (w', rank1)          -> watershed hierarchy T
//   saliency + Kruskal #3 (sal, key) -> re-binarised tree (hg.filter_small_nodes_from_tree)
//   Kruskal #4  (wb, rank3) + canonize, contour strength, frontier / max_area
//   simplification, height, node emission (vendored/hierarchy.py:172-218,273-309)
//
// One thread walks one hierarchy; the parallelism is across the (typically 1e4-1e6)
// hierarchies of a frame.  Everything is integer / exact arithmetic except the
// final float features.
#pragma once
#include "ulk_rt.h"

typedef unsigned __int128 u128;

// ------------------------------------------------------------------ small helpers
template <class Less> ULK_HD void ulk_heapsort(i32 *a, i32 n, Less less)
{
    // in-place heapsort of an index array (no recursion, O(1) extra space)
    for (i32 start = n / 2 - 1; start >= 0; --start) {
        i32 root = start;
        for (;;) {
            i32 child = 2 * root + 1;
            if (child >= n) break;
            if (child + 1 < n && less(a[child], a[child + 1])) ++child;
            if (!less(a[root], a[child])) break;
            i32 t = a[root]; a[root] = a[child]; a[child] = t;
            root = child;
        }
    }
    for (i32 end = n - 1; end > 0; --end) {
        i32 t = a[0]; a[0] = a[end]; a[end] = t;
        i32 root = 0;
        for (;;) {
            i32 child = 2 * root + 1;
            if (child >= end) break;
            if (child + 1 < end && less(a[child], a[child + 1])) ++child;
            if (!less(a[root], a[child])) break;
            i32 t2 = a[root]; a[root] = a[child]; a[child] = t2;
            root = child;
        }
    }
}

ULK_HD ULK_INLINE i32 ulk_uf_find(i32 *uf, i32 x)
{
    while (uf[x] != x) { uf[x] = uf[uf[x]]; x = uf[x]; }
    return x;
}

// exact u128 -> double, round to nearest even (matches gcc's __floattidf)
ULK_HD ULK_INLINE double ulk_u128_to_double(u128 x)
{
    if (x == 0) return 0.0;
    u64 hi = (u64)(x >> 64), lo = (u64)x;
    int top;  // index of the highest set bit
    if (hi) { top = 127; while (!((hi >> (top - 64)) & 1)) --top; }
    else { top = 63; while (!((lo >> top) & 1)) --top; }
    if (top <= 52) return (double)lo;
    int sh = top - 52;                 // keep 53 bits
    u128 m = x >> sh;
    u128 rem = x & ((((u128)1) << sh) - 1);
    u128 half = ((u128)1) << (sh - 1);
    if (rem > half || (rem == half && (m & 1))) ++m;
    double r = (double)(u64)m;         // m <= 2^53, exact
    // scale by 2^sh exactly
    for (int i = 0; i < sh / 30; ++i) r *= 1073741824.0;
    r *= (double)(1u << (sh % 30));
    return r;
}

// ------------------------------------------------------------------ data views
struct UlkBasins {          // [B], grouped by hierarchy
    const i32 *size;        // voxels
    const u32 *L;           // lowest edge weight of the basin (order-preserving float bits)
    const u64 *top;         // max (w, key) over the basin's spanning-tree edges
    const u32 *k3;          // bottleneck edge key of the basin under raster-key Kruskal
    const i32 *bbmin, *bbmax;   // [3B] z,y,x
    const i64 *csum;        // [3B] coordinate sums
};
struct UlkEdges {           // [Eg], grouped by hierarchy, ascending (w, key) inside a group
    const i32 *a, *b;       // global basin ids
    const u64 *wk;          // min (w, key) over the voxel edges between the two basins
    const u32 *k3;          // min edge key
    const u32 *cnt;         // number of voxel edges
    const u64 *shi, *slo;   // exact fixed-point sum of their weights (hi/lo 32-bit lanes)
};
struct UlkHierScratch {     // node space [2B] unless noted
    i32 *Ipar, *Ic0, *Ic1;  // BPT of Kruskal #1 over basins
    u64 *Iwk;               // key of the merge
    u32 *IminL;             // lowest basin level below the node
    i32 *attach;            // voxels that join the node before its parent forms (area)
    u64 *awhi, *awlo;       // fixed-point sum of those voxels' steepest-edge weights (volume; hi/lo 32-bit lanes)
    i32 *acc1;              // [B] RAG edge accepted as k-th merge (edge space offset)
    double *wprime;         // new MST weight per inner node of I
    i32 *par, *c0, *c1;     // working binary tree (T1, Tb, T2 in turn)
    double *alt;            // altitudes of the working tree
    double *alt1;           // altitudes of T1 (kept for saliency)
    i32 *par1;              // parents of T1
    i32 *acc;               // [B] accepted edge per merge of the working Kruskal
    i32 *uf, *rootnode;     // [B]
    i32 *ord;               // [max(B,Eg)] sort permutation (edge space offset es, size r or b)
    double *sal;            // [Eg]
    i32 *area;              // node areas
    u64 *Xhi, *Xlo;         // u128 crossing-weight sums
    i64 *Xc;                // crossing-edge counts
    i32 *fp;                // final parent (nearest surviving ancestor)
    u8 *surv;               // surviving flag
    i32 *fin;               // surviving nodes in final index order
    double *fr, *low;       // frontier, lowest inner altitude
    i32 *emit_idx;          // node -> emitted slot (hierarchy local) or -1
    i32 *nbbmin, *nbbmax;   // [3*2B]
    i64 *ncsum;             // [3*2B]
};
struct UlkHierParams {
    int kind;               // 0 area, 1 dynamics, 2 volume
    double min_area, max_area, min_frontier;
    int fix_shift;
};
// emitted node records, hierarchy-local slots at node-space offset 2*bs
struct UlkNodeOut {
    i32 *count;             // [R] nodes emitted per hierarchy
    i32 *area;              // [2B]
    double *frontier, *height;
    i32 *bbmin, *bbmax;     // [3*2B]
    i64 *csum;              // [3*2B]
    i32 *parent;            // emitted slot of the tree parent, -1 if none
    i32 *hidx;              // voxel-level tree index minus n_leaves (position among surviving nodes)
    i32 *basin_node;        // [B] lowest emitted node above the basin (local slot) or -1
    i32 *nsurv;             // [R] surviving inner-node count
    i32 *rsize;             // [R] voxels of the region (= leaves of higra's tree)
};

// Kruskal over `m` edges visited in the order ord[0..m): endpoints via functor.
// Builds a binary tree over b leaves in par/c0/c1 (local node ids); acc[k] = edge of
// the k-th merge.  Returns the number of merges.
template <class EP>
ULK_HD i32 ulk_kruskal(i32 b, i32 m, const i32 *ord, EP ep, i32 *uf, i32 *rootnode,
                       i32 *par, i32 *c0, i32 *c1, i32 *acc)
{
    for (i32 i = 0; i < b; ++i) { uf[i] = i; rootnode[i] = i; }
    for (i32 i = 0; i < 2 * b - 1; ++i) { par[i] = i; c0[i] = -1; c1[i] = -1; }
    i32 k = 0;
    for (i32 t = 0; t < m && k < b - 1; ++t) {
        i32 e = ord[t], x, y;
        ep(e, x, y);
        i32 rx = ulk_uf_find(uf, x), ry = ulk_uf_find(uf, y);
        if (rx == ry) continue;
        i32 nn = b + k;
        i32 n0 = rootnode[rx], n1 = rootnode[ry];
        par[n0] = nn; par[n1] = nn;
        c0[nn] = n0 < n1 ? n0 : n1;
        c1[nn] = n0 < n1 ? n1 : n0;
        uf[ry] = rx;
        rootnode[rx] = nn;
        acc[k++] = e;
    }
    return k;
}

ULK_HD ULK_INLINE i32 ulk_lca(const i32 *par, i32 a, i32 b)
{
    while (a != b) { if (a < b) a = par[a]; else b = par[b]; }
    return a;
}

// ------------------------------------------------------------------ stage 1
// Kruskal #1 on the RAG, edges already sorted by (w, key).  Fills the I tree.
ULK_HD inline void ulk_hier_stage1(i32 h, const i32 *bs_, const i32 *es_, const UlkBasins &B,
                                   const UlkEdges &E, const UlkHierScratch &S)
{
    i32 bs = bs_[h], b = bs_[h + 1] - bs;
    i32 es = es_[h], r = es_[h + 1] - es;
    i32 o = 2 * bs;
    i32 *ord = S.ord + es;
    for (i32 j = 0; j < r; ++j) ord[j] = j;
    const i32 *ea = E.a + es, *eb = E.b + es;
    auto ep = [&](i32 e, i32 &x, i32 &y) { x = ea[e] - bs; y = eb[e] - bs; };
    ulk_kruskal(b, r, ord, ep, S.uf + bs, S.rootnode + bs, S.Ipar + o, S.Ic0 + o, S.Ic1 + o, S.acc1 + bs);
    for (i32 i = 0; i < b; ++i) { S.IminL[o + i] = B.L[bs + i]; S.Iwk[o + i] = 0; S.attach[o + i] = B.size[bs + i]; }
    for (i32 i = 0; i < 2 * b - 1; ++i) { S.awhi[o + i] = 0; S.awlo[o + i] = 0; }
    for (i32 k = 0; k < b - 1; ++k) {
        i32 x = o + b + k;
        S.Iwk[x] = E.wk[es + S.acc1[bs + k]];
        u32 l0 = S.IminL[o + S.Ic0[x]], l1 = S.IminL[o + S.Ic1[x]];
        S.IminL[x] = l0 < l1 ? l0 : l1;
        S.attach[x] = 0;
    }
}

// new MST weights w'(x) = min over the two children of the attribute
// fixed-point image of a float32 edge weight (the same map for voxel sums and altitudes)
ULK_HD ULK_INLINE i64 ulk_w_fixed(float w, double fixscale) { return (i64)((double)w * fixscale); }

ULK_HD inline void ulk_hier_wprime(i32 bs, i32 b, const UlkHierScratch &S, int kind, int fix_shift = 0)
{
    i32 o = 2 * bs;
    if (kind == 2) {
        // hg.attribute_volume on the voxel-level BPT: V(c) = sum over the inner nodes d below (and
        // at) c of area(d) * (alt(parent d) - alt(d)).  For one voxel v the terms of its ancestors
        // telescope to alt(parent c) - alt(first inner ancestor of v), and that first ancestor is
        // created by v's steepest edge, so V(c) = area(c) * alt(x) - sum_{v in c} mw(v) with x the
        // merge above c.  Exact in 2^-fix_shift fixed point, rounded to double once.
        double fixscale = 1.0;
        for (int k = 0; k < fix_shift; ++k) fixscale *= 2.0;
        for (i32 i = 0; i < 2 * b - 1; ++i) {
            S.area[o + i] = S.attach[o + i];
            u128 x = ((u128)S.awhi[o + i] << 32) + (u128)S.awlo[o + i];
            S.Xhi[o + i] = (u64)(x >> 64); S.Xlo[o + i] = (u64)x;
        }
        for (i32 i = 0; i < 2 * b - 2; ++i) {
            i32 p = o + S.Ipar[o + i];
            S.area[p] += S.area[o + i];
            u128 x = (((u128)S.Xhi[p] << 64) | S.Xlo[p]) + (((u128)S.Xhi[o + i] << 64) | S.Xlo[o + i]);
            S.Xhi[p] = (u64)(x >> 64); S.Xlo[p] = (u64)x;
        }
        double inv = 1.0;
        for (int k = 0; k < fix_shift; ++k) inv *= 0.5;
        for (i32 x = b; x < 2 * b - 1; ++x) {
            const i64 ax = ulk_w_fixed(ulk_ord2f((u32)(S.Iwk[o + x] >> 32)), fixscale);
            u128 vmin = 0;
            for (int c = 0; c < 2; ++c) {
                const i32 ch = o + (c ? S.Ic1[o + x] : S.Ic0[o + x]);
                const u128 full = (u128)(u64)S.area[ch] * (u128)(u64)ax;
                const u128 sub = ((u128)S.Xhi[ch] << 64) | S.Xlo[ch];
                const u128 v = full > sub ? full - sub : 0;
                if (c == 0 || v < vmin) vmin = v;
            }
            S.wprime[o + x] = ulk_u128_to_double(vmin) * inv;
        }
    } else if (kind == 0) {
        // area of each I node = voxels attached below it
        for (i32 i = 0; i < 2 * b - 1; ++i) S.area[o + i] = S.attach[o + i];
        for (i32 i = 0; i < 2 * b - 2; ++i) S.area[o + S.Ipar[o + i]] += S.area[o + i];
        for (i32 x = b; x < 2 * b - 1; ++x) {
            // area(child) as it is when the merge happens: without the voxels that
            // attach to x itself
            i32 a0 = S.area[o + S.Ic0[o + x]], a1 = S.area[o + S.Ic1[o + x]];
            S.wprime[o + x] = (double)(a0 < a1 ? a0 : a1);
        }
    } else {
        for (i32 x = b; x < 2 * b - 1; ++x) {
            float wx = ulk_ord2f((u32)(S.Iwk[o + x] >> 32));
            float h0 = wx - ulk_ord2f(S.IminL[o + S.Ic0[o + x]]);
            float h1 = wx - ulk_ord2f(S.IminL[o + S.Ic1[o + x]]);
            S.wprime[o + x] = (double)(h0 < h1 ? h0 : h1);
        }
    }
}

// ------------------------------------------------------------------ stage 2A (oversegmentation cut)
// hg.labelisation_horizontal_cut_from_threshold on the area hierarchy
// (vendored/hierarchy.py:355-360): regions = basins joined by merges with w' <= max_area.
// region_of[basin] = global basin id of the region representative; for representatives
// rkey_w / rkey_k hold the ordering key (= index order of the cut node in higra's tree).
ULK_HD inline void ulk_hier_cut(i32 h, const i32 *bs_, const i32 *es_, const UlkBasins &B,
                                const UlkEdges &E, const UlkHierScratch &S, double max_area,
                                i32 *region_of, u64 *rkey_w, u64 *rkey_k)
{
    i32 bs = bs_[h], b = bs_[h + 1] - bs;
    i32 es = es_[h];
    i32 o = 2 * bs;
    ulk_hier_wprime(bs, b, S, 0);
    i32 *uf = S.uf + bs;
    for (i32 i = 0; i < b; ++i) { uf[i] = i; rkey_w[bs + i] = ulk_d2ord(0.0); rkey_k[bs + i] = B.top[bs + i]; }
    for (i32 k = 0; k < b - 1; ++k) {
        double wp = S.wprime[o + b + k];
        if (!(wp <= max_area)) continue;
        i32 e = S.acc1[bs + k];
        i32 x = ulk_uf_find(uf, E.a[es + e] - bs), y = ulk_uf_find(uf, E.b[es + e] - bs);
        // merges are visited in rank1 order, so the last one with the largest w' wins
        u64 kw = rkey_w[bs + x] > rkey_w[bs + y] ? rkey_w[bs + x] : rkey_w[bs + y];
        u64 kk = 0;
        u64 mine = ulk_d2ord(wp);
        if (mine >= kw) { kw = mine; kk = S.Iwk[o + b + k]; }
        else kk = rkey_w[bs + x] > rkey_w[bs + y] ? rkey_k[bs + x]
                  : (rkey_w[bs + y] > rkey_w[bs + x] ? rkey_k[bs + y]
                     : (rkey_k[bs + x] > rkey_k[bs + y] ? rkey_k[bs + x] : rkey_k[bs + y]));
        uf[y] = x;
        rkey_w[bs + x] = kw;
        rkey_k[bs + x] = kk;
    }
    for (i32 i = 0; i < b; ++i) region_of[bs + i] = bs + ulk_uf_find(uf, i);
}

// ------------------------------------------------------------------ stage 2B (full hierarchy)
ULK_HD inline void ulk_hier_full(i32 h, const i32 *bs_, const i32 *es_, const UlkBasins &B,
                                 const UlkEdges &E, const UlkHierScratch &S,
                                 const UlkHierParams &P, const UlkNodeOut &O)
{
    const i32 bs = bs_[h], b = bs_[h + 1] - bs;
    const i32 es = es_[h], r = es_[h + 1] - es;
    const i32 o = 2 * bs;
    const i32 N = 2 * b - 1;
    i32 *par = S.par + o, *c0 = S.c0 + o, *c1 = S.c1 + o;
    double *alt = S.alt + o;
    i32 *uf = S.uf + bs, *rootnode = S.rootnode + bs, *acc = S.acc + bs;
    const i32 *ea = E.a + es, *eb = E.b + es;
    i32 *ord = S.ord + es;   // r >= b-1 whenever b > 1
    i32 *area = S.area + o;

    if (b > 1) {
        // ---- Kruskal #2: I's merges re-weighted, ties by rank1
        ulk_hier_wprime(bs, b, S, P.kind, P.fix_shift);
        const double *wp = S.wprime + o;
        for (i32 k = 0; k < b - 1; ++k) ord[k] = k;
        ulk_heapsort(ord, b - 1, [&](i32 x, i32 y) {
            double wx = wp[b + x], wy = wp[b + y];
            return wx < wy || (wx == wy && x < y);
        });
        const i32 *acc1 = S.acc1 + bs;
        auto ep1 = [&](i32 k, i32 &x, i32 &y) { i32 e = acc1[k]; x = ea[e] - bs; y = eb[e] - bs; };
        i32 *par1 = S.par1 + o;
        double *alt1 = S.alt1 + o;
        ulk_kruskal(b, b - 1, ord, ep1, uf, rootnode, par1, c0, c1, acc);
        for (i32 i = 0; i < b; ++i) alt1[i] = 0.0;
        for (i32 k = 0; k < b - 1; ++k) alt1[b + k] = wp[b + acc[k]];
        // ---- saliency of every RAG edge, Kruskal #3 by (sal, min edge key)
        double *sal = S.sal + es;
        for (i32 j = 0; j < r; ++j) sal[j] = alt1[ulk_lca(par1, ea[j] - bs, eb[j] - bs)];
        const u32 *k3 = E.k3 + es;
        for (i32 j = 0; j < r; ++j) ord[j] = j;
        ulk_heapsort(ord, r, [&](i32 x, i32 y) {
            return sal[x] < sal[y] || (sal[x] == sal[y] && k3[x] < k3[y]);
        });
        auto ep3 = [&](i32 e, i32 &x, i32 &y) { x = ea[e] - bs; y = eb[e] - bs; };
        ulk_kruskal(b, r, ord, ep3, uf, rootnode, par, c0, c1, acc);
        // areas of the re-binarised tree; non-relevant merges get weight 0
        for (i32 i = 0; i < N; ++i) area[i] = i < b ? B.size[bs + i] : 0;
        for (i32 i = 0; i < N - 1; ++i) area[par[i]] += area[i];
        double *wb = S.wprime + o;   // reuse: indexed by merge k
        for (i32 k = 0; k < b - 1; ++k) {
            i32 y = b + k;
            i32 a0 = area[c0[y]], a1 = area[c1[y]];
            double m = (double)(a0 < a1 ? a0 : a1);
            wb[k] = m < P.min_area ? 0.0 : sal[acc[k]];
        }
        // ---- Kruskal #4 on the MST of #3, ties by rank3
        i32 *acc3 = S.acc1 + bs;     // reuse: RAG edge of the k-th merge of #3
        for (i32 k = 0; k < b - 1; ++k) acc3[k] = acc[k];
        for (i32 k = 0; k < b - 1; ++k) ord[k] = k;
        ulk_heapsort(ord, b - 1, [&](i32 x, i32 y) { return wb[x] < wb[y] || (wb[x] == wb[y] && x < y); });
        auto ep4 = [&](i32 k, i32 &x, i32 &y) { i32 e = acc3[k]; x = ea[e] - bs; y = eb[e] - bs; };
        ulk_kruskal(b, b - 1, ord, ep4, uf, rootnode, par, c0, c1, acc);
        for (i32 i = 0; i < b; ++i) alt[i] = 0.0;
        for (i32 k = 0; k < b - 1; ++k) alt[b + k] = wb[acc[k]];
    } else {
        par[0] = 0; c0[0] = c1[0] = -1; alt[0] = 0.0;
    }
    const i32 root = N - 1;

    // ---- canonize at basin scale: which nodes of the voxel-level tree survive
    u8 *surv = S.surv + o;
    for (i32 y = 0; y < N; ++y) {
        if (y == root) surv[y] = 1;
        else if (y < b) surv[y] = alt[par[y]] > 0.0;        // the basin is an altitude-0 node
        else surv[y] = alt[y] != alt[par[y]];
    }
    // ---- areas, node statistics (bottom-up over the binary tree)
    i32 *nbmin = S.nbbmin + 3 * o, *nbmax = S.nbbmax + 3 * o;
    i64 *ncs = S.ncsum + 3 * o;
    for (i32 i = 0; i < N; ++i) {
        if (i < b) {
            area[i] = B.size[bs + i];
            for (int d = 0; d < 3; ++d) {
                nbmin[3 * i + d] = B.bbmin[3 * (bs + i) + d];
                nbmax[3 * i + d] = B.bbmax[3 * (bs + i) + d];
                ncs[3 * i + d] = B.csum[3 * (bs + i) + d];
            }
        } else {
            area[i] = 0;
            for (int d = 0; d < 3; ++d) { nbmin[3 * i + d] = 0x7fffffff; nbmax[3 * i + d] = -0x7fffffff; ncs[3 * i + d] = 0; }
        }
    }
    for (i32 i = 0; i < N - 1; ++i) {
        i32 p = par[i];
        area[p] += area[i];
        for (int d = 0; d < 3; ++d) {
            if (nbmin[3 * i + d] < nbmin[3 * p + d]) nbmin[3 * p + d] = nbmin[3 * i + d];
            if (nbmax[3 * i + d] > nbmax[3 * p + d]) nbmax[3 * p + d] = nbmax[3 * i + d];
            ncs[3 * p + d] += ncs[3 * i + d];
        }
    }
    // ---- contour strength (hg.attribute_contour_strength): crossing sums, exact
    u64 *Xhi = S.Xhi + o, *Xlo = S.Xlo + o;
    i64 *Xc = S.Xc + o;
    for (i32 i = 0; i < N; ++i) { Xhi[i] = 0; Xlo[i] = 0; Xc[i] = 0; }
    // signed 128-bit accumulate in (hi,lo): use two's complement on u128
    auto addx = [&](i32 y, u128 v, i64 c, bool neg) {
        u128 cur = ((u128)Xhi[y] << 64) | Xlo[y];
        cur = neg ? cur - v : cur + v;
        Xhi[y] = (u64)(cur >> 64); Xlo[y] = (u64)cur;
        Xc[y] += neg ? -c : c;
    };
    for (i32 j = 0; j < r; ++j) {
        u128 ws = ((u128)E.shi[es + j] << 32) + (u128)E.slo[es + j];
        i64 c = (i64)E.cnt[es + j];
        i32 x = ea[j] - bs, y = eb[j] - bs;
        addx(x, ws, c, false);
        addx(y, ws, c, false);
        i32 l = ulk_lca(par, x, y);
        addx(l, 2 * ws, 2 * c, true);
    }
    for (i32 i = 0; i < N - 1; ++i) {
        u128 ci = ((u128)Xhi[i] << 64) | Xlo[i];
        u128 cp = ((u128)Xhi[par[i]] << 64) | Xlo[par[i]];
        cp += ci;
        Xhi[par[i]] = (u64)(cp >> 64); Xlo[par[i]] = (u64)cp;
        Xc[par[i]] += Xc[i];
    }
    double *fr = S.fr + o;
    const double scale = 1.0 / (double)(1ull << 31) / (double)(1ull << 31);   // 2^-62
    for (i32 i = 0; i < N; ++i) {
        if (i == root || Xc[i] <= 0) { fr[i] = 0.0; continue; }
        u128 x = ((u128)Xhi[i] << 64) | Xlo[i];
        double s = ulk_u128_to_double(x);
        // 2^-shift, exact
        double sc = scale;
        for (int k = P.fix_shift; k < 62; ++k) sc *= 2.0;
        fr[i] = (s * sc) / (double)Xc[i];
    }
    // ---- final parents (nearest surviving proper ancestor)
    i32 *fp = S.fp + o;
    auto refresh_fp = [&]() {
        for (i32 y = N - 1; y >= 0; --y) {
            if (y == root) { fp[y] = root; continue; }
            i32 p = par[y];
            fp[y] = surv[p] ? p : fp[p];
        }
    };
    refresh_fp();
    // ---- _filter_contour_strength (vendored/hierarchy.py:125-143)
    if (P.min_frontier > 0.0) {
        u8 *del = (u8 *)S.emit_idx + (size_t)4 * o;   // borrow as byte flags for this step
        for (i32 y = 0; y < N; ++y) {
            del[y] = 0;
            if (!surv[y] || y == root) continue;
            if (fr[y] < P.min_frontier && !((double)area[fp[y]] > P.max_area)) del[y] = 1;
        }
        for (i32 y = 0; y < N; ++y) if (del[y]) surv[y] = 0;
        refresh_fp();
    }
    // ---- simplify_tree(area > max_area) (vendored/hierarchy.py:187-192)
    for (i32 y = 0; y < N; ++y)
        if (surv[y] && y != root && (double)area[y] > P.max_area) surv[y] = 0;
    refresh_fp();
    // ---- final index order: altitude-0 single basins by their raster-key bottleneck,
    //      then the inner nodes of the binary tree by index
    i32 *fin = S.fin + o;
    i32 nf = 0;
    for (i32 y = 0; y < b; ++y) if (surv[y] && !(b == 1)) fin[nf++] = y;
    const u32 *bk3 = B.k3 + bs;
    ulk_heapsort(fin, nf, [&](i32 x, i32 y) { return bk3[x] < bk3[y]; });
    if (b == 1) fin[nf++] = 0;
    else for (i32 y = b; y < N; ++y) if (surv[y]) fin[nf++] = y;
    // ---- height (hg.attribute_height): alt[parent] - min inner altitude below
    double *low = S.low + o;
    for (i32 t = 0; t < nf; ++t) low[fin[t]] = alt[fin[t]];
    for (i32 t = 0; t < nf; ++t) {
        i32 y = fin[t];
        if (y == root) continue;
        if (low[y] < low[fp[y]]) low[fp[y]] = low[y];
    }
    // ---- emission (vendored/hierarchy.py:273-309)
    i32 *emit = S.emit_idx + o;
    for (i32 y = 0; y < N; ++y) emit[y] = -1;
    i32 ne = 0;
    for (i32 t = 0; t < nf; ++t) {
        i32 y = fin[t];
        if ((double)area[y] <= P.max_area) emit[y] = ne++;
    }
    bool fallback = ne == 0;
    if (fallback) { emit[root] = 0; ne = 1; }
    O.count[h] = ne;
    O.nsurv[h] = nf;
    O.rsize[h] = area[root];
    for (i32 t = 0; t < nf; ++t) {
        i32 y = fin[t];
        i32 s = emit[y];
        if (s < 0) continue;
        i32 g = o + s;
        O.area[g] = area[y];
        if (fallback) { O.frontier[g] = -1.0; O.height[g] = -1.0; }
        else {
            O.frontier[g] = fr[y];
            if (P.kind == 1) O.height[g] = (double)((float)alt[fp[y]] - (float)low[y]);
            else O.height[g] = alt[fp[y]] - low[y];
        }
        for (int d = 0; d < 3; ++d) {
            O.bbmin[3 * g + d] = nbmin[3 * y + d];
            O.bbmax[3 * g + d] = nbmax[3 * y + d];
            O.csum[3 * g + d] = ncs[3 * y + d];
        }
        O.parent[g] = (y != root && emit[fp[y]] >= 0) ? emit[fp[y]] : -1;
        O.hidx[g] = t;
    }
    // lowest emitted node above (or at) each basin
    for (i32 i = 0; i < b; ++i) {
        i32 y = i;
        i32 res = -1;
        if (b == 1) res = emit[0];
        else {
            if (surv[y] && emit[y] >= 0) res = emit[y];
            else {
                for (;;) {
                    i32 p = fp[y];
                    if (emit[p] >= 0) { res = emit[p]; break; }
                    if (p == y) break;
                    y = p;
                }
            }
        }
        O.basin_node[bs + i] = res;
    }
}
