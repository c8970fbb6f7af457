// pts_bvh.cpp — buildBVH / buildBVHwithSAH with the reference's exact output, built differently.
//
// Reference: ImportanceSampling_LowDiscrepancySequence/main.cpp:377-428 (median split) and :431-552
// (full-sweep SAH), comparators :153-167.  The reference std::sort()s 144-byte Triangle structs and
// recomputes both centroids inside every comparison; each SAH node does four such sorts.  The node
// array and the final triangle order are required to be BIT-EXACT (they decide traversal order and
// therefore which triangle wins a tie), and they depend on std::sort's behaviour for equal keys.
//
// Design here: sort 48-byte records {centroid xyz, id, bounds, left-sweep z-bounds} instead.
// std::sort's sequence of comparisons and moves depends only on comparison RESULTS, so sorting
// records whose key compares exactly like the reference comparator (same centroid expression,
// `(p1+p2+p3)/3` evaluated component-wise in binary32) yields the same permutation as sorting the
// triangles themselves — provided every sort the reference performs is performed, in the same
// order, on the same range (x, y, z, then the winning axis again; M:466-546).  The triangles are
// gathered once at the end.  Subtrees are independent, so they can be built as parallel tasks and
// stitched in preorder (node ids are assigned in preorder by the reference's push_back, M:434-435).
//
// Preserved source quirks (they are the spec): the left sweep uses p2.x in the z extent (M:481, 485);
// node 0 is the "testNode" dummy (M:780-786); a child index 0 means "no child".
#include "pts_scene.h"

#include <algorithm>
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace pts {

namespace {

const float kInf = 2147483648.0f;                    // float(INF), INF = 2147483647.0 (M:19)

struct Rec {
    float cx, cy, cz;                                 // centroid, as cmpx/cmpy/cmpz compute it
    int id;                                           // index into the input triangle array
    float mnx, mny, mnz, mxx, mxy, mxz;               // true bounds
    float lmnz, lmxz;                                 // the left sweep's z "bounds" (p2.x instead of p2.z)
};
static_assert(sizeof(Rec) == 48, "Rec is 48 bytes");

inline float mx(float a, float b) { return (a < b) ? b : a; }     // glm::max
inline float mn(float a, float b) { return (b < a) ? b : a; }     // glm::min

inline bool cmp_x(const Rec& a, const Rec& b) { return a.cx < b.cx; }
inline bool cmp_y(const Rec& a, const Rec& b) { return a.cy < b.cy; }
inline bool cmp_z(const Rec& a, const Rec& b) { return a.cz < b.cz; }

inline void sort_axis(Rec* first, Rec* last, int axis) {
    if (axis == 0) std::sort(first, last, cmp_x);
    else if (axis == 1) std::sort(first, last, cmp_y);
    else std::sort(first, last, cmp_z);
}

Rec make_rec(const Tri& t, int id) {
    Rec r;
    // (t.p1 + t.p2 + t.p3) / vec3(3,3,3): two adds then a divide, per component
    r.cx = ((t.p1.x + t.p2.x) + t.p3.x) / 3.0f;
    r.cy = ((t.p1.y + t.p2.y) + t.p3.y) / 3.0f;
    r.cz = ((t.p1.z + t.p2.z) + t.p3.z) / 3.0f;
    r.id = id;
    r.mnx = mn(t.p1.x, mn(t.p2.x, t.p3.x)); r.mny = mn(t.p1.y, mn(t.p2.y, t.p3.y)); r.mnz = mn(t.p1.z, mn(t.p2.z, t.p3.z));
    r.mxx = mx(t.p1.x, mx(t.p2.x, t.p3.x)); r.mxy = mx(t.p1.y, mx(t.p2.y, t.p3.y)); r.mxz = mx(t.p1.z, mx(t.p2.z, t.p3.z));
    r.lmnz = mn(t.p1.z, mn(t.p2.x, t.p3.z));
    r.lmxz = mx(t.p1.z, mx(t.p2.x, t.p3.z));
    return r;
}

struct Builder {
    Rec* recs;
    int leaf_n;
    bool sah;
    int task_cutoff;                                  // ranges larger than this spawn tasks (0 = never)

    // Appends the subtree over recs[l..r] to `out` with ids relative to out's own start + base, where
    // `base` is the absolute id the subtree root will have.  Returns the root's absolute id.
    int build(std::vector<Node>& out, int base, int l, int r) {
        if (l > r) return 0;
        const size_t me = out.size();
        out.push_back(Node());
        Node nd;
        nd.left = nd.right = nd.n = nd.index = 0;
        nd.AA = V3{kInf, kInf, kInf};
        nd.BB = V3{-kInf, -kInf, -kInf};
        for (int i = l; i <= r; i++) {
            const Rec& t = recs[i];
            nd.AA.x = mn(nd.AA.x, t.mnx); nd.AA.y = mn(nd.AA.y, t.mny); nd.AA.z = mn(nd.AA.z, t.mnz);
            nd.BB.x = mx(nd.BB.x, t.mxx); nd.BB.y = mx(nd.BB.y, t.mxy); nd.BB.z = mx(nd.BB.z, t.mxz);
        }
        const int count = r - l + 1;
        if (count <= leaf_n) {
            nd.n = count;
            nd.index = l;
            out[me] = nd;
            return base;
        }
        int split;
        if (sah) split = sah_split(l, r);
        else {
            float lenx = nd.BB.x - nd.AA.x, leny = nd.BB.y - nd.AA.y, lenz = nd.BB.z - nd.AA.z;
            // three independent ifs, not else-ifs (M:412-420): ties sort more than once
            if (lenx >= leny && lenx >= lenz) sort_axis(recs + l, recs + r + 1, 0);
            if (leny >= lenx && leny >= lenz) sort_axis(recs + l, recs + r + 1, 1);
            if (lenz >= lenx && lenz >= leny) sort_axis(recs + l, recs + r + 1, 2);
            split = (l + r) / 2;
        }
        if (task_cutoff > 0 && count > task_cutoff) {
            // children as independent tasks into private arrays, then stitched in preorder
            std::vector<Node> lsub, rsub;
            #pragma omp task shared(lsub) firstprivate(l, split)
            { build(lsub, 0, l, split); }
            #pragma omp task shared(rsub) firstprivate(split, r)
            { build(rsub, 0, split + 1, r); }
            #pragma omp taskwait
            const int lbase = base + 1, rbase = base + 1 + (int)lsub.size();
            append_rebased(out, lsub, lbase);
            append_rebased(out, rsub, rbase);
            nd.left = lsub.empty() ? 0 : lbase;
            nd.right = rsub.empty() ? 0 : rbase;
        } else {
            nd.left = build(out, base + 1, l, split);
            const int used = (int)(out.size() - me);  // self + left subtree
            nd.right = build(out, base + used, split + 1, r);
        }
        out[me] = nd;
        return base;
    }

    static void append_rebased(std::vector<Node>& out, const std::vector<Node>& sub, int new_base) {
        // sub was built with base 0: child ids are offsets from sub's root; 0 still means "none"
        // (a subtree root is never anyone's child inside its own subtree, so offset 0 is unambiguous)
        for (const Node& s : sub) {
            Node n = s;
            if (n.left) n.left += new_base;
            if (n.right) n.right += new_base;
            out.push_back(n);
        }
    }

    // M:463-546: for each axis sort, sweep prefix/suffix bounds, take the cheapest split; then sort by
    // the winning axis again.  Returns Split.
    int sah_split(int l, int r) {
        const int cnt = r - l + 1;
        float Cost = kInf;
        int Axis = 0;
        int Split = (l + r) / 2;
        std::vector<float> rmn(3 * (size_t)cnt), rmx(3 * (size_t)cnt);
        for (int axis = 0; axis < 3; axis++) {
            sort_axis(recs + l, recs + r + 1, axis);
            // suffix bounds [i, r] (M:490-503), true z
            float ax = kInf, ay = kInf, az = kInf, bx = -kInf, by = -kInf, bz = -kInf;
            for (int i = r; i >= l; i--) {
                const Rec& t = recs[i];
                bx = mx(bx, t.mxx); by = mx(by, t.mxy); bz = mx(bz, t.mxz);
                ax = mn(ax, t.mnx); ay = mn(ay, t.mny); az = mn(az, t.mnz);
                size_t k = 3 * (size_t)(i - l);
                rmn[k] = ax; rmn[k + 1] = ay; rmn[k + 2] = az;
                rmx[k] = bx; rmx[k + 1] = by; rmx[k + 2] = bz;
            }
            // prefix bounds [l, i] (M:474-486, with the p2.x quirk in z) fused with the cost loop (M:505-531)
            float cost = kInf;
            int split = l;
            ax = ay = az = kInf; bx = by = bz = -kInf;
            for (int i = l; i <= r - 1; i++) {
                const Rec& t = recs[i];
                bx = mx(bx, t.mxx); by = mx(by, t.mxy); bz = mx(bz, t.lmxz);
                ax = mn(ax, t.mnx); ay = mn(ay, t.mny); az = mn(az, t.lmnz);
                float lenx = bx - ax, leny = by - ay, lenz = bz - az;
                float leftS = 2.0f * ((lenx * leny) + (lenx * lenz) + (leny * lenz));
                float leftCost = leftS * (float)(i - l + 1);
                size_t k = 3 * (size_t)(i + 1 - l);
                lenx = rmx[k] - rmn[k]; leny = rmx[k + 1] - rmn[k + 1]; lenz = rmx[k + 2] - rmn[k + 2];
                float rightS = 2.0f * ((lenx * leny) + (lenx * lenz) + (leny * lenz));
                float rightCost = rightS * (float)(r - i);
                float totalCost = leftCost + rightCost;
                if (totalCost < cost) { cost = totalCost; split = i; }
            }
            if (cost < Cost) { Cost = cost; Axis = axis; Split = split; }
        }
        sort_axis(recs + l, recs + r + 1, Axis);
        return Split;
    }
};

}  // namespace

int build_bvh(std::vector<Tri>& tris, std::vector<Node>& nodes, int leaf_n, bool sah, int n_threads) {
    const int T = (int)tris.size();
    std::vector<Rec> recs((size_t)T);
    for (int i = 0; i < T; i++) recs[i] = make_rec(tris[i], i);

    nodes.clear();
    Node dummy;                                       // M:780-786
    dummy.left = 255; dummy.right = 128; dummy.n = 30; dummy.index = 0;
    dummy.AA = V3{1, 1, 0}; dummy.BB = V3{0, 1, 0};
    nodes.push_back(dummy);

    Builder b{recs.data(), leaf_n, sah, 0};
#ifdef _OPENMP
    if (n_threads > 1 && T > 4096) {
        b.task_cutoff = std::max(2048, T / (8 * n_threads));
        #pragma omp parallel num_threads(n_threads)
        #pragma omp single
        { b.build(nodes, 1, 0, T - 1); }
    } else
#endif
    {
        (void)n_threads;
        b.build(nodes, 1, 0, T - 1);
    }

    // gather the triangles into their final order (the reference moved them during every sort)
    std::vector<Tri> sorted((size_t)T);
    for (int i = 0; i < T; i++) sorted[i] = tris[recs[i].id];
    tris.swap(sorted);
    return (int)nodes.size();
}

}  // namespace pts
