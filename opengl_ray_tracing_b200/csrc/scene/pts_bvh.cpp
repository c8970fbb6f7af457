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
#include <deque>
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

    // bounds of recs[l..r] into nd (M:381-393 / M:436-448)
    void init_node(Node& nd, int l, int r) const {
        nd.left = nd.right = nd.n = nd.index = 0;
        nd.AA = V3{kInf, kInf, kInf};
        nd.BB = V3{-kInf, -kInf, -kInf};
        for (int i = l; i <= r; i++) {
            const Rec& t = recs[i];
            nd.AA.x = mn(nd.AA.x, t.mnx); nd.AA.y = mn(nd.AA.y, t.mny); nd.AA.z = mn(nd.AA.z, t.mnz);
            nd.BB.x = mx(nd.BB.x, t.mxx); nd.BB.y = mx(nd.BB.y, t.mxy); nd.BB.z = mx(nd.BB.z, t.mxz);
        }
    }

    // sorts recs[l..r] the way the reference does at this node and returns the split position
    int split_node(const Node& nd, int l, int r) {
        if (sah) return sah_split(l, r);
        float lenx = nd.BB.x - nd.AA.x, leny = nd.BB.y - nd.AA.y, lenz = nd.BB.z - nd.AA.z;
        // three independent ifs, not else-ifs (M:412-420): ties sort more than once
        if (lenx >= leny && lenx >= lenz) sort_axis(recs + l, recs + r + 1, 0);
        if (leny >= lenx && leny >= lenz) sort_axis(recs + l, recs + r + 1, 1);
        if (lenz >= lenx && lenz >= leny) sort_axis(recs + l, recs + r + 1, 2);
        return (l + r) / 2;
    }

    // The reference recurses (left subtree, then right subtree, node ids handed out in preorder by push_back).  Its own
    // SAH trees are thousands of levels deep on large meshes (6781 at 1 M triangles), so nothing here recurses: this
    // builds the subtree over recs[l..r] into `out` with an explicit stack.  Child ids are relative to the subtree
    // root's id `base` (id of a node = base + its position in `out` - position of the root).
    void build_seq(std::vector<Node>& out, int base, int l, int r) {
        if (l > r) return;
        struct Frame { size_t me; int l, r, split, state; };
        std::vector<Frame> st;
        const size_t root = out.size();
        out.push_back(Node());
        st.push_back(Frame{root, l, r, 0, 0});
        while (!st.empty()) {
            Frame& f = st.back();
            if (f.state == 0) {
                Node nd;
                init_node(nd, f.l, f.r);
                const int count = f.r - f.l + 1;
                if (count <= leaf_n) {
                    nd.n = count;
                    nd.index = f.l;
                    out[f.me] = nd;
                    st.pop_back();
                    continue;
                }
                f.split = split_node(nd, f.l, f.r);
                f.state = 1;
                out[f.me] = nd;
                if (f.l <= f.split) {                                  // left child = the next node in preorder
                    const size_t c = out.size();
                    out[f.me].left = base + (int)(c - root);
                    out.push_back(Node());
                    const Frame child{c, f.l, f.split, 0, 0};
                    st.push_back(child);                               // invalidates f
                }
            } else if (f.state == 1) {
                f.state = 2;
                if (f.split + 1 <= f.r) {
                    const size_t c = out.size();
                    out[f.me].right = base + (int)(c - root);
                    out.push_back(Node());
                    const Frame child{c, f.split + 1, f.r, 0, 0};
                    st.push_back(child);
                }
            } else {
                st.pop_back();
            }
        }
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

    Builder b{recs.data(), leaf_n, sah};
    bool done = false;
#ifdef _OPENMP
    if (n_threads > 1 && T > 4096) {
        // Ranges larger than `cutoff` ("big" nodes) are split by ONE thread walking the top of the tree in preorder with
        // an explicit stack — each level needs the previous level's sort order, there is nothing to parallelise on that
        // spine — and every range at or below the cutoff becomes a task that builds its subtree into a private array.
        // The arrays are stitched in preorder afterwards.
        const int cutoff = std::max(2048, T / (8 * n_threads));
        struct Big { Node nd; int child[2]; bool child_is_job[2]; };
        std::vector<Big> bigs;
        std::deque<std::vector<Node>> jobs;                              // deque: references stay valid while tasks run
        struct Walk { int big, l, r, split, state; };
        #pragma omp parallel num_threads(n_threads)
        #pragma omp single
        {
            std::vector<Walk> st;
            // returns (index, is_job) of the node built over [l, r]; big nodes are pushed onto the walk stack
            auto open_range = [&](int l, int r, int& idx, bool& is_job) {
                if (r - l + 1 <= cutoff) {
                    jobs.emplace_back();
                    std::vector<Node>* dst = &jobs.back();
                    idx = (int)jobs.size() - 1; is_job = true;
                    Builder* bp = &b;
                    #pragma omp task firstprivate(dst, bp, l, r)
                    { bp->build_seq(*dst, 0, l, r); }
                } else {
                    Big g;
                    b.init_node(g.nd, l, r);
                    g.child[0] = g.child[1] = -1; g.child_is_job[0] = g.child_is_job[1] = false;
                    const int split = b.split_node(g.nd, l, r);
                    bigs.push_back(g);
                    idx = (int)bigs.size() - 1; is_job = false;
                    st.push_back(Walk{idx, l, r, split, 0});
                }
            };
            int root_idx; bool root_job;
            open_range(0, T - 1, root_idx, root_job);
            while (!st.empty()) {
                Walk w = st.back();
                if (w.state == 2) { st.pop_back(); continue; }
                st.back().state = w.state + 1;
                const int lo = w.state == 0 ? w.l : w.split + 1, hi = w.state == 0 ? w.split : w.r;
                if (lo > hi) continue;
                int idx; bool is_job;
                open_range(lo, hi, idx, is_job);                        // may grow `bigs` and `st`
                bigs[w.big].child[w.state] = idx;
                bigs[w.big].child_is_job[w.state] = is_job;
            }
            #pragma omp taskwait
            // stitch in preorder: id of a node = its position in `nodes` (the dummy is 0, the root 1)
            if (root_job) {
                Builder::append_rebased(nodes, jobs[root_idx], 1);
            } else {
                struct Emit { int big; int at; int state; };
                std::vector<Emit> es;
                es.push_back(Emit{root_idx, (int)nodes.size(), 0});
                nodes.push_back(bigs[root_idx].nd);
                while (!es.empty()) {
                    Emit e = es.back();
                    if (e.state == 2) { es.pop_back(); continue; }
                    es.back().state = e.state + 1;
                    const Big& g = bigs[e.big];
                    const int c = g.child[e.state];
                    if (c < 0) continue;
                    const int id = (int)nodes.size();
                    if (e.state == 0) nodes[e.at].left = id; else nodes[e.at].right = id;
                    if (g.child_is_job[e.state]) {
                        Builder::append_rebased(nodes, jobs[c], id);
                    } else {
                        nodes.push_back(bigs[c].nd);
                        es.push_back(Emit{c, id, 0});
                    }
                }
            }
        }
        done = true;
    }
#endif
    (void)n_threads;
    if (!done) b.build_seq(nodes, 1, 0, T - 1);

    // gather the triangles into their final order (the reference moved them during every sort)
    std::vector<Tri> sorted((size_t)T);
    for (int i = 0; i < T; i++) sorted[i] = tris[recs[i].id];
    tris.swap(sorted);
    return (int)nodes.size();
}

}  // namespace pts
