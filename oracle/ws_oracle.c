/*
 * ws_oracle.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement (plain C) of the
 * hierarchical-watershed arithmetic that ultrack's segmentation path obtains
 * from numba + higra.  Nothing under ultrack_b200/ may link or call this; it is
 * the checker for tests/, __graft_entry__.smoke() and bench.py's cpu_baseline.
 *
 * Reference call sites (relative to /root/reference/ultrack):
 *   core/segmentation/vendored/graph.py:9-59      edges_from_mask (numba)
 *   core/segmentation/vendored/hierarchy.py:145-194  Hierarchy.watershed_hierarchy
 *   core/segmentation/vendored/hierarchy.py:125-143  _filter_contour_strength
 *   core/segmentation/vendored/hierarchy.py:327-365  oversegment_components
 *   core/segmentation/vendored/hierarchy.py:196-218  area / height attributes
 *
 * The higra (0.6.12, pixi.lock:137) source is NOT in /root/reference and higra is
 * not installable here, so every hg.* call is restated from higra's published
 * algorithms (Najman, Cousty, Perret, "Playing with Kruskal", ISMM 2013; higra
 * docs).  PARITY vs REAL higra IS UNPINNED; the pieces that
 * could be pinned (edges_from_mask, the IoU kernels) are pinned against the
 * reference's numba code in tests/golden/.
 *
 * Conventions: a tree is a parent array, leaves first (0..n_leaves-1), the
 * root last, parent[root] == root and parent[i] > i otherwise (higra's layout).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef int64_t i64;

#define ULK_O_EXPORT __attribute__((visibility("default")))

/* ------------------------------------------------------------------ graph */

/* vendored/graph.py:9-59.  mask: uint8 [D,H,W]; img: double [D,H,W] holding the
 * exactly-representable image values; f32add != 0 means the image dtype is
 * float32 (or was up-cast to it), so `image[p] + image[q]` rounds to float32
 * before the float64 `/ 2.0 + 1e-8` (numba typing).
 * Vertex id = raster rank of the voxel inside the mask; edges are emitted per
 * voxel in raster order as +x, +y, +z.  src/dst/w need room for 3*count. */
ULK_O_EXPORT i64 ulk_o_edges_from_mask(const uint8_t *mask, const double *img,
                                       i64 D, i64 H, i64 W, int f32add,
                                       i64 *ref, i64 *src, i64 *dst, float *w,
                                       i64 *n_vertices)
{
    i64 count = 0, V = D * H * W;
    for (i64 i = 0; i < V; ++i) ref[i] = mask[i] ? count++ : -1;
    *n_vertices = count;
    i64 e = 0;
    const i64 off[3] = {1, W, W * H};
    for (i64 z = 0; z < D; ++z)
        for (i64 y = 0; y < H; ++y)
            for (i64 x = 0; x < W; ++x) {
                i64 p = (z * H + y) * W + x;
                if (!mask[p]) continue;
                int ok[3] = {x + 1 < W, y + 1 < H, z + 1 < D};
                for (int d = 0; d < 3; ++d) {
                    if (!ok[d]) continue;
                    i64 q = p + off[d];
                    if (!mask[q]) continue;
                    double s;
                    if (f32add) s = (double)((float)img[p] + (float)img[q]);
                    else s = img[p] + img[q];
                    src[e] = ref[p];
                    dst[e] = ref[q];
                    w[e] = (float)(s / 2.0 + 1e-8);
                    ++e;
                }
            }
    return e;
}

/* ------------------------------------------------------------ union-find */

static i64 uf_find(i64 *uf, i64 x)
{
    i64 r = x;
    while (uf[r] != r) r = uf[r];
    while (uf[x] != r) { i64 n = uf[x]; uf[x] = r; x = n; }
    return r;
}

typedef struct { double w; i64 i; } srec;

static int srec_cmp(const void *a, const void *b)
{
    const srec *x = (const srec *)a, *y = (const srec *)b;
    if (x->w < y->w) return -1;
    if (x->w > y->w) return 1;
    return (x->i > y->i) - (x->i < y->i);
}

/* hg.bpt_canonical [UPSTREAM-RECALL]: stable sort of the edges by weight (ties
 * by edge index), Kruskal; the k-th accepted edge creates node n+k whose
 * altitude is that edge's weight and whose children are the two current roots.
 * Returns the number of accepted edges (n-1 iff connected). */
ULK_O_EXPORT i64 ulk_o_bpt_canonical(i64 n, i64 E, const i64 *src, const i64 *dst,
                                     const double *w, i64 *parents, double *alt,
                                     i64 *mst_edge_map)
{
    srec *order = (srec *)malloc(sizeof(srec) * (size_t)(E > 0 ? E : 1));
    for (i64 i = 0; i < E; ++i) { order[i].w = w[i]; order[i].i = i; }
    qsort(order, (size_t)E, sizeof(srec), srec_cmp);
    i64 *uf = (i64 *)malloc(sizeof(i64) * (size_t)n);
    i64 *root = (i64 *)malloc(sizeof(i64) * (size_t)n);
    for (i64 i = 0; i < n; ++i) { uf[i] = i; root[i] = i; }
    for (i64 i = 0; i < 2 * n - 1; ++i) { parents[i] = i; alt[i] = 0.0; }
    i64 found = 0, nn = n;
    for (i64 k = 0; k < E && found < n - 1; ++k) {
        i64 ei = order[k].i;
        i64 c1 = uf_find(uf, src[ei]), c2 = uf_find(uf, dst[ei]);
        if (c1 == c2) continue;
        alt[nn] = w[ei];
        parents[root[c1]] = nn;
        parents[root[c2]] = nn;
        uf[c2] = c1;
        root[c1] = nn;
        mst_edge_map[found++] = ei;
        ++nn;
    }
    free(order); free(uf); free(root);
    return found;
}

/* ---------------------------------------------------------- tree helpers */

/* hg.simplify_tree [UPSTREAM-RECALL]: drop the flagged inner nodes (never the
 * root, never leaves); children re-attach to the nearest kept ancestor;
 * survivors keep their relative order.  node_map[new] = old.  Returns the new
 * node count. */
static i64 simplify_tree(i64 nl, i64 nn, const i64 *par, const uint8_t *del,
                         i64 *npar, i64 *node_map)
{
    i64 *newidx = (i64 *)malloc(sizeof(i64) * (size_t)nn);
    i64 c = 0;
    for (i64 i = 0; i < nn; ++i) {
        int drop = (i >= nl) && (i != nn - 1) && del[i];
        if (drop) newidx[i] = -1;
        else { newidx[i] = c; node_map[c] = i; ++c; }
    }
    for (i64 k = 0; k < c; ++k) {
        i64 i = node_map[k];
        i64 p = par[i];
        while (newidx[p] < 0) p = par[p];
        npar[k] = newidx[p];
    }
    free(newidx);
    return c;
}

/* hg.canonize_hierarchy: delete inner nodes whose altitude equals the parent's */
static i64 canonize(i64 nl, i64 nn, const i64 *par, const double *alt,
                    i64 *npar, double *nalt)
{
    uint8_t *del = (uint8_t *)calloc((size_t)nn, 1);
    i64 *map = (i64 *)malloc(sizeof(i64) * (size_t)nn);
    for (i64 i = nl; i < nn - 1; ++i) del[i] = alt[i] == alt[par[i]];
    i64 c = simplify_tree(nl, nn, par, del, npar, map);
    for (i64 k = 0; k < c; ++k) nalt[k] = alt[map[k]];
    free(del); free(map);
    return c;
}

static void tree_area(i64 nl, i64 nn, const i64 *par, double *area)
{
    for (i64 i = 0; i < nn; ++i) area[i] = i < nl ? 1.0 : 0.0;
    for (i64 i = 0; i < nn - 1; ++i) area[par[i]] += area[i];
}

/* min altitude over the non-leaf nodes of each subtree (+inf for leaves) */
static void tree_lowest_inner(i64 nl, i64 nn, const i64 *par, const double *alt,
                              double *low)
{
    for (i64 i = 0; i < nn; ++i) low[i] = i < nl ? INFINITY : alt[i];
    for (i64 i = nl; i < nn - 1; ++i)
        if (low[i] < low[par[i]]) low[par[i]] = low[i];
}

static i64 tree_lca(const i64 *par, i64 a, i64 b)
{
    while (a != b) { if (a < b) a = par[a]; else b = par[b]; }
    return a;
}

/* --------------------------------------------------- watershed hierarchies */

/* hg.watershed_hierarchy_by_{area,dynamics,volume} [UPSTREAM-RECALL]
 * (config/segmentationconfig.py:12-16): BPT by altitude ordering; increasing
 * attribute; extinction values (a child that holds no regional minimum has
 * extinction 0; among the children holding minima the one with the largest
 * attribute inherits the parent's extinction, the others take their own
 * attribute); every MST edge is re-weighted with the min over the two children
 * of its BPT node; BPT of the MST under the new weights; canonize.
 * kind: 0 area, 1 dynamics (attribute_height, float32), 2 volume.
 * f32 altitudes: the weights are float32, area/volume attributes float64.
 * Outputs a canonized tree (capacity 2n-1) and returns its node count. */
ULK_O_EXPORT i64 ulk_o_watershed_hierarchy(i64 n, i64 E, const i64 *src,
                                           const i64 *dst, const float *w, int kind,
                                           i64 *parents_out, double *alt_out)
{
    if (n == 1) { parents_out[0] = 0; alt_out[0] = 0.0; return 1; }
    i64 N = 2 * n - 1;
    double *wd = (double *)malloc(sizeof(double) * (size_t)(E > 0 ? E : 1));
    for (i64 i = 0; i < E; ++i) wd[i] = (double)w[i];
    i64 *p0 = (i64 *)malloc(sizeof(i64) * (size_t)N);
    double *a0 = (double *)malloc(sizeof(double) * (size_t)N);
    i64 *mst = (i64 *)malloc(sizeof(i64) * (size_t)n);
    i64 found = ulk_o_bpt_canonical(n, E, src, dst, wd, p0, a0, mst);
    if (found != n - 1) { free(wd); free(p0); free(a0); free(mst); return -1; }

    double *area = (double *)malloc(sizeof(double) * (size_t)N);
    double *low = (double *)malloc(sizeof(double) * (size_t)N);
    double *attr = (double *)malloc(sizeof(double) * (size_t)N);
    tree_area(n, N, p0, area);
    tree_lowest_inner(n, N, p0, a0, low);
    if (kind == 0) {
        memcpy(attr, area, sizeof(double) * (size_t)N);
    } else if (kind == 1) {
        /* attribute_height: alt[parent] - min inner altitude of the subtree,
         * float32 arithmetic because the altitudes are float32 */
        for (i64 i = 0; i < N; ++i)
            attr[i] = i < n ? 0.0 : (double)((float)a0[p0[i]] - (float)low[i]);
    } else {
        /* attribute_volume: leaves 0; inner = sum(children) + area*|alt[parent]-alt| */
        for (i64 i = 0; i < N; ++i) attr[i] = 0.0;
        for (i64 i = n; i < N; ++i) {
            attr[i] += area[i] * fabs(a0[p0[i]] - a0[i]);
            if (i != N - 1) attr[p0[i]] += attr[i];
        }
    }
    /* regional minima (attribute_extrema): inner node whose subtree is flat at
     * its altitude and whose parent is at another altitude (a flat root counts) */
    uint8_t *hasmin = (uint8_t *)calloc((size_t)N, 1);
    for (i64 i = n; i < N; ++i) {
        int flat = low[i] == a0[i];
        if (flat && (i == N - 1 || a0[p0[i]] != a0[i])) hasmin[i] = 1;
        if (hasmin[i] && i != N - 1) hasmin[p0[i]] = 1;
    }
    /* children of the binary tree */
    i64 *c0 = (i64 *)malloc(sizeof(i64) * (size_t)N), *c1 = (i64 *)malloc(sizeof(i64) * (size_t)N);
    for (i64 i = 0; i < N; ++i) c0[i] = c1[i] = -1;
    for (i64 i = 0; i < N - 1; ++i) { if (c0[p0[i]] < 0) c0[p0[i]] = i; else c1[p0[i]] = i; }
    double *ext = (double *)calloc((size_t)N, sizeof(double));
    ext[N - 1] = attr[N - 1];
    for (i64 i = N - 1; i >= n; --i) {
        i64 ch[2] = {c0[i], c1[i]};
        i64 best = -1;
        for (int k = 0; k < 2; ++k)
            if (hasmin[ch[k]] && (best < 0 || attr[ch[k]] > attr[best])) best = ch[k];
        for (int k = 0; k < 2; ++k) {
            i64 c = ch[k];
            if (!hasmin[c]) ext[c] = 0.0;
            else if (c == best) ext[c] = ext[i];
            else ext[c] = attr[c];
        }
    }
    i64 *ms = (i64 *)malloc(sizeof(i64) * (size_t)n), *md = (i64 *)malloc(sizeof(i64) * (size_t)n);
    double *nw = (double *)malloc(sizeof(double) * (size_t)n);
    for (i64 k = 0; k < n - 1; ++k) {
        i64 x = n + k;
        double a = ext[c0[x]], b = ext[c1[x]];
        nw[k] = a < b ? a : b;
        ms[k] = src[mst[k]];
        md[k] = dst[mst[k]];
    }
    i64 *p1 = (i64 *)malloc(sizeof(i64) * (size_t)N);
    double *a1 = (double *)malloc(sizeof(double) * (size_t)N);
    i64 *mst1 = (i64 *)malloc(sizeof(i64) * (size_t)n);
    ulk_o_bpt_canonical(n, n - 1, ms, md, nw, p1, a1, mst1);
    i64 out = canonize(n, N, p1, a1, parents_out, alt_out);
    free(wd); free(p0); free(a0); free(mst); free(area); free(low); free(attr);
    free(hasmin); free(c0); free(c1); free(ext); free(ms); free(md); free(nw);
    free(p1); free(a1); free(mst1);
    return out;
}

/* hg.filter_small_nodes_from_tree [UPSTREAM-RECALL] (vendored/hierarchy.py:172):
 * the canonized input tree does not carry an MST, so it is re-binarised through
 * its saliency map: sal(e) = alt[lca(u,v)] on every edge of the leaf graph,
 * bpt_canonical(leaf_graph, sal).  An inner node of that binary tree is
 * non-relevant iff min(area(children)) < size_threshold; its MST edge gets
 * weight 0; bpt_canonical(mst, weights); canonize. */
static i64 filter_small_nodes(i64 n, i64 E, const i64 *src, const i64 *dst,
                              i64 nn, const i64 *par, const double *alt,
                              double size_threshold, i64 *parents_out, double *alt_out)
{
    (void)nn;
    if (n == 1) { parents_out[0] = 0; alt_out[0] = 0.0; return 1; }
    i64 N = 2 * n - 1;
    double *sal = (double *)malloc(sizeof(double) * (size_t)(E > 0 ? E : 1));
    for (i64 e = 0; e < E; ++e) sal[e] = alt[tree_lca(par, src[e], dst[e])];
    i64 *pb = (i64 *)malloc(sizeof(i64) * (size_t)N);
    double *ab = (double *)malloc(sizeof(double) * (size_t)N);
    i64 *mstb = (i64 *)malloc(sizeof(i64) * (size_t)n);
    ulk_o_bpt_canonical(n, E, src, dst, sal, pb, ab, mstb);
    double *area = (double *)malloc(sizeof(double) * (size_t)N);
    double *minc = (double *)malloc(sizeof(double) * (size_t)N);
    tree_area(n, N, pb, area);
    for (i64 i = 0; i < N; ++i) minc[i] = INFINITY;
    for (i64 i = 0; i < N - 1; ++i) if (area[i] < minc[pb[i]]) minc[pb[i]] = area[i];
    i64 *ms = (i64 *)malloc(sizeof(i64) * (size_t)n), *md = (i64 *)malloc(sizeof(i64) * (size_t)n);
    double *nw = (double *)malloc(sizeof(double) * (size_t)n);
    for (i64 k = 0; k < n - 1; ++k) {
        nw[k] = minc[n + k] < size_threshold ? 0.0 : ab[n + k];
        ms[k] = src[mstb[k]];
        md[k] = dst[mstb[k]];
    }
    i64 *p2 = (i64 *)malloc(sizeof(i64) * (size_t)N);
    double *a2 = (double *)malloc(sizeof(double) * (size_t)N);
    i64 *mst2 = (i64 *)malloc(sizeof(i64) * (size_t)n);
    ulk_o_bpt_canonical(n, n - 1, ms, md, nw, p2, a2, mst2);
    i64 out = canonize(n, N, p2, a2, parents_out, alt_out);
    free(sal); free(pb); free(ab); free(mstb); free(area); free(minc);
    free(ms); free(md); free(nw); free(p2); free(a2); free(mst2);
    return out;
}

/* exact, order-independent accumulation of float32 weights: fixed point with
 * 2^shift fractional bits in a 128-bit integer (the CUDA path splits the same
 * integer over 64-bit atomics) */
static __int128 fix_weight(float w, int shift)
{
    return (__int128)(long long)((double)w * ldexp(1.0, shift));
}

/* hg.attribute_contour_strength with no_border_vertex_out_degree=None
 * [UPSTREAM-RECALL] (vendored/hierarchy.py:174-176): mean weight of the leaf
 * graph edges with exactly one end under the node; 0 for the root. */
static void contour_strength(i64 n, i64 E, const i64 *src, const i64 *dst,
                             const float *w, i64 nn, const i64 *par, int shift,
                             double *frontier)
{
    __int128 *sum = (__int128 *)calloc((size_t)nn, sizeof(__int128));
    i64 *cnt = (i64 *)calloc((size_t)nn, sizeof(i64));
    (void)n;
    for (i64 e = 0; e < E; ++e) {
        i64 l = tree_lca(par, src[e], dst[e]);
        __int128 f = fix_weight(w[e], shift);
        for (i64 a = src[e]; a != l; a = par[a]) { sum[a] += f; cnt[a]++; }
        for (i64 a = dst[e]; a != l; a = par[a]) { sum[a] += f; cnt[a]++; }
    }
    for (i64 i = 0; i < nn; ++i) {
        if (cnt[i] == 0) { frontier[i] = 0.0; continue; }
        double s = ldexp((double)sum[i], -shift);
        frontier[i] = s / (double)cnt[i];
    }
    frontier[nn - 1] = 0.0;
    free(sum); free(cnt);
}

/* Hierarchy.watershed_hierarchy + area + height (vendored/hierarchy.py:145-218).
 * Inputs: the leaf graph of one region (from ulk_o_edges_from_mask).
 * Outputs (capacity 2n-1 each): parents, alt, frontier, area, height.
 * Returns the node count of the final tree, -1 if the graph is not connected. */
ULK_O_EXPORT i64 ulk_o_hierarchy(i64 n, i64 E, const i64 *src, const i64 *dst,
                                 const float *w, int kind, double min_area,
                                 double max_area, double min_frontier, int fix_shift,
                                 i64 *parents, double *alt, double *frontier,
                                 double *area, double *height)
{
    i64 N = 2 * n - 1;
    i64 *pa = (i64 *)malloc(sizeof(i64) * (size_t)N), *pb = (i64 *)malloc(sizeof(i64) * (size_t)N);
    double *aa = (double *)malloc(sizeof(double) * (size_t)N), *ab = (double *)malloc(sizeof(double) * (size_t)N);
    i64 nn = ulk_o_watershed_hierarchy(n, E, src, dst, w, kind, pa, aa);
    if (nn < 0) { free(pa); free(pb); free(aa); free(ab); return -1; }
    nn = filter_small_nodes(n, E, src, dst, nn, pa, aa, min_area, pb, ab);
    double *fr = (double *)malloc(sizeof(double) * (size_t)N);
    contour_strength(n, E, src, dst, w, nn, pb, fix_shift, fr);
    uint8_t *del = (uint8_t *)calloc((size_t)N, 1);
    i64 *map = (i64 *)malloc(sizeof(i64) * (size_t)N);
    double *ar = (double *)malloc(sizeof(double) * (size_t)N);
    if (min_frontier > 0.0) {
        /* _filter_contour_strength, vendored/hierarchy.py:125-143 */
        tree_area(n, nn, pb, ar);
        for (i64 i = 0; i < nn; ++i) {
            del[i] = fr[i] < min_frontier;
            if (ar[pb[i]] > max_area) del[i] = 0;
        }
        i64 c = simplify_tree(n, nn, pb, del, pa, map);
        for (i64 k = 0; k < c; ++k) { aa[k] = ab[map[k]]; frontier[k] = fr[map[k]]; }
        nn = c;
        memcpy(pb, pa, sizeof(i64) * (size_t)nn);
        memcpy(ab, aa, sizeof(double) * (size_t)nn);
        memcpy(fr, frontier, sizeof(double) * (size_t)nn);
    }
    /* simplify_tree(tree, area > max_area), vendored/hierarchy.py:187-192 */
    tree_area(n, nn, pb, ar);
    for (i64 i = 0; i < nn; ++i) del[i] = ar[i] > max_area;
    i64 c = simplify_tree(n, nn, pb, del, parents, map);
    for (i64 k = 0; k < c; ++k) { alt[k] = ab[map[k]]; frontier[k] = fr[map[k]]; }
    nn = c;
    tree_area(n, nn, parents, area);
    /* hg.attribute_height(tree, alt): alt[parent] - min inner altitude below.
     * float32 arithmetic for the dynamics hierarchy (float32 altitudes). */
    double *low = (double *)malloc(sizeof(double) * (size_t)nn);
    tree_lowest_inner(n, nn, parents, alt, low);
    for (i64 i = 0; i < nn; ++i) {
        if (i < n) { height[i] = 0.0; continue; }
        if (kind == 1) height[i] = (double)((float)alt[parents[i]] - (float)low[i]);
        else height[i] = alt[parents[i]] - low[i];
    }
    free(pa); free(pb); free(aa); free(ab); free(fr); free(del); free(map); free(ar); free(low);
    return nn;
}

/* oversegment_components body for one oversize component
 * (vendored/hierarchy.py:355-360): area hierarchy, then
 * hg.labelisation_horizontal_cut_from_threshold(tree, alt, max_area)
 * [UPSTREAM-RECALL]: a leaf takes the index of its highest ancestor reachable
 * through parents of altitude <= threshold.  The caller applies
 * relabel_sequential (sorted unique values -> consecutive labels). */
ULK_O_EXPORT i64 ulk_o_cut_labels(i64 n, i64 E, const i64 *src, const i64 *dst,
                                  const float *w, double threshold, i64 *leaf_label)
{
    i64 N = 2 * n - 1;
    i64 *par = (i64 *)malloc(sizeof(i64) * (size_t)N);
    double *alt = (double *)malloc(sizeof(double) * (size_t)N);
    i64 nn = ulk_o_watershed_hierarchy(n, E, src, dst, w, 0, par, alt);
    if (nn < 0) { free(par); free(alt); return -1; }
    i64 *lab = (i64 *)malloc(sizeof(i64) * (size_t)nn);
    lab[nn - 1] = nn - 1;
    for (i64 i = nn - 2; i >= 0; --i) lab[i] = alt[par[i]] <= threshold ? lab[par[i]] : i;
    for (i64 i = 0; i < n; ++i) leaf_label[i] = lab[i];
    free(par); free(alt); free(lab);
    return nn;
}

/* For every leaf, the list of flagged ancestors (used to paint node masks:
 * vendored/node.py:59-82 yields exactly "the leaves under the node").
 * out_node/out_leaf receive (node, leaf) pairs; returns the pair count, or the
 * required capacity (negated) when cap is too small. */
ULK_O_EXPORT i64 ulk_o_leaf_memberships(i64 n, i64 nn, const i64 *par,
                                        const uint8_t *emit, i64 cap,
                                        i64 *out_node, i64 *out_leaf)
{
    i64 c = 0;
    for (i64 i = 0; i < n; ++i) {
        i64 a = i;
        for (;;) {
            i64 p = par[a];
            if (p == a) break;
            a = p;
            if (emit[a]) { if (c < cap) { out_node[c] = a; out_leaf[c] = i; } ++c; }
        }
    }
    (void)nn;
    return c <= cap ? c : -c;
}

/* ----------------------------------------------------------------- linking */

/* core/segmentation/node.py:23-118 (_fast_iou_with_bbox and friends), restated
 * for dense uint8 masks laid out [bd,bh,bw]; 2-D masks use bd == 1 and bboxes
 * with a leading zero/one z range. bbox = (z0,y0,x0,z1,y1,x1), max exclusive. */
ULK_O_EXPORT double ulk_o_iou(const i64 *b1, const uint8_t *m1, const i64 *b2,
                              const uint8_t *m2)
{
    for (int i = 0; i < 3; ++i) {
        int k = i + 3;
        int ok = (b1[i] <= b2[i] && b2[i] < b1[k]) || (b1[i] < b2[k] && b2[k] <= b1[k]) ||
                 (b2[i] <= b1[i] && b1[i] < b2[k]) || (b2[i] < b1[k] && b1[k] <= b2[k]);
        if (!ok) return 0.0;
    }
    i64 lo[3], hi[3];
    for (int i = 0; i < 3; ++i) {
        lo[i] = b1[i] > b2[i] ? b1[i] : b2[i];
        hi[i] = b1[i + 3] < b2[i + 3] ? b1[i + 3] : b2[i + 3];
    }
    i64 s1[3] = {b1[3] - b1[0], b1[4] - b1[1], b1[5] - b1[2]};
    i64 s2[3] = {b2[3] - b2[0], b2[4] - b2[1], b2[5] - b2[2]};
    i64 inter = 0;
    for (i64 z = lo[0]; z < hi[0]; ++z)
        for (i64 y = lo[1]; y < hi[1]; ++y)
            for (i64 x = lo[2]; x < hi[2]; ++x) {
                i64 i1 = ((z - b1[0]) * s1[1] + (y - b1[1])) * s1[2] + (x - b1[2]);
                i64 i2 = ((z - b2[0]) * s2[1] + (y - b2[1])) * s2[2] + (x - b2[2]);
                inter += (m1[i1] != 0) & (m2[i2] != 0);
            }
    if (inter == 0) return 0.0;
    i64 a1 = 0, a2 = 0;
    for (i64 i = 0; i < s1[0] * s1[1] * s1[2]; ++i) a1 += m1[i] != 0;
    for (i64 i = 0; i < s2[0] * s2[1] * s2[2]; ++i) a2 += m2[i] != 0;
    return (double)inter / (double)(a1 + a2 - inter);
}
