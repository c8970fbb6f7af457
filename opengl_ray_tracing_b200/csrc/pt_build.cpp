// pt_build.cpp — host half of pt_upload_scene: the TRAVERSAL topology.
//
// What the reference's node array (buildBVHwithSAH, IS/main.cpp:431-552) decides about the RESULT of hitBVH:
//   (1) a triangle counts only if the ray passes hitAABB() > 0 on the box of the reference LEAF that holds it
//       (pass1.fsh:356-379; a leaf's box being hit implies every ancestor's box is hit: the slab arithmetic is
//       monotone in the box corners and ancestors are exact unions);
//   (2) among counted triangles the smallest t wins, and an exact-distance tie goes to the one visited first.
// Everything else about that tree — its depth, which boxes overlap — only decides how much work a ray costs.
// On large meshes the reference's tree degenerates (its left sweep mixes p2.x into the z extent,
// IS/main.cpp:481-485): the 1M-triangle scene gets a tree 6781 levels deep whose leaves are long slivers; rays
// through the image centre pierce thousands of leaf boxes (measured: one pixel = 16 ms of a 19 ms launch when
// the traversal tree was built over the reference's leaves, profiles/r01_cost_map.md).
//
// So the kernels traverse a tree built HERE over the individual triangles: a binned-SAH binary BVH (16 bins x 3
// axes, leaves of <= PT_LEAF_MAX triangles), triangles re-ordered so that a leaf is a contiguous range.  Its
// boxes are padded by 2^-17 of the scene radius, which makes "the slab test passes" a strict superset of "the
// reference's hitTriangle accepts" (the accepted point lies within a few ulps of the ray and of the triangle).
// Rule (1) is then applied exactly: a triangle that passes hitTriangle is accepted only if the ray also passes
// the reference's own hitAABB on the box of the reference leaf holding it (leaf id stored per triangle).  Rule (2)
// can only differ on an exact tie between two triangles: the kernel detects that and re-traces the ray on the
// reference topology (pt_trace.cuh, trace_wide_fallback).
//
// Output: 64-byte wide-node records (both child boxes in the parent; packed child refs: count << 24 | index, where
// a leaf's index is the first triangle IN THE NEW ORDER), the new order (new -> uploaded index), and per triangle
// the reference leaf node that holds it.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace pt {

struct FastTree {
    std::vector<float> wnodes;       // 16 floats per inner node
    std::vector<int> order;          // new position -> uploaded triangle index
    std::vector<int> leaf_node;      // new position -> reference node index of the leaf holding the triangle
    uint32_t root_ref = 0;
    int depth = 0;
    int n_leaves = 0;
    float pad = 0.0f;
};

namespace {

struct Box {
    float mn[3], mx[3];
    void reset() { mn[0] = mn[1] = mn[2] = INFINITY; mx[0] = mx[1] = mx[2] = -INFINITY; }
    void grow(const Box& b) {
        for (int a = 0; a < 3; a++) { mn[a] = std::min(mn[a], b.mn[a]); mx[a] = std::max(mx[a], b.mx[a]); }
    }
    void grow_pt(const float* p) {
        for (int a = 0; a < 3; a++) { mn[a] = std::min(mn[a], p[a]); mx[a] = std::max(mx[a], p[a]); }
    }
    float half_area() const {
        float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
        return dx * dy + dy * dz + dz * dx;
    }
};

struct Prim {
    Box box;               // exact box of the triangle's three vertices
    float c[3];            // box centre
    int tri;               // uploaded triangle index
};

const int NBINS = 16;

struct Node {              // build-time node; child < 0 => leaf: ~(first | count << 24) packed below
    Box lbox, rbox;
    int64_t left, right;   // >= 0: node index; < 0: ~packed leaf ref
};

inline int64_t leaf_code(int first, int count) { return ~(int64_t)(((uint32_t)count << 24) | (uint32_t)first); }

struct Builder {
    std::vector<Prim>& prims;
    int task_cutoff;
    int leaf_max;
    float c_node;          // SAH cost of an inner-node step relative to one triangle test

    // builds the subtree over prims[lo, hi) into `out` (indices local to out); returns node index or a leaf code
    int64_t build(std::vector<Node>& out, int lo, int hi, Box& bounds, int depth, int& max_depth, int& n_leaves) {
        bounds.reset();
        for (int i = lo; i < hi; i++) bounds.grow(prims[i].box);
        if (depth > max_depth) max_depth = depth;
        const int n = hi - lo;
        if (n == 1) { n_leaves++; return leaf_code(lo, 1); }
        Box cb; cb.reset();
        for (int i = lo; i < hi; i++) cb.grow_pt(prims[i].c);
        // binned SAH over the three axes
        float best_cost = INFINITY;
        int best_axis = -1, best_bin = -1;
        for (int axis = 0; axis < 3; axis++) {
            float ext = cb.mx[axis] - cb.mn[axis];
            if (!(ext > 0.0f)) continue;
            Box bb[NBINS]; int cnt[NBINS];
            for (int b = 0; b < NBINS; b++) { bb[b].reset(); cnt[b] = 0; }
            float k = (float)NBINS / ext;
            for (int i = lo; i < hi; i++) {
                int b = std::min(NBINS - 1, (int)((prims[i].c[axis] - cb.mn[axis]) * k));
                bb[b].grow(prims[i].box); cnt[b]++;
            }
            float rarea[NBINS]; int rcnt[NBINS];
            Box acc; acc.reset(); int m = 0;
            for (int b = NBINS - 1; b > 0; b--) { acc.grow(bb[b]); m += cnt[b]; rarea[b] = acc.half_area(); rcnt[b] = m; }
            acc.reset(); m = 0;
            for (int b = 0; b < NBINS - 1; b++) {
                acc.grow(bb[b]); m += cnt[b];
                if (m == 0 || rcnt[b + 1] == 0) continue;
                float cost = acc.half_area() * (float)m + rarea[b + 1] * (float)rcnt[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = axis; best_bin = b; }
            }
        }
        if (n <= leaf_max) {
            // leaf when splitting does not pay: n tests here vs one node step + the expected tests below
            float area = bounds.half_area();
            float split = (best_axis >= 0 && area > 0.0f) ? c_node + best_cost / area : INFINITY;
            if ((float)n <= split) { n_leaves++; return leaf_code(lo, n); }
        }
        int mid;
        if (best_axis < 0) {
            mid = (lo + hi) / 2;                       // all centres coincide: split the range in half
        } else {
            float ext = cb.mx[best_axis] - cb.mn[best_axis], k = (float)NBINS / ext, base = cb.mn[best_axis];
            int ax = best_axis, bin = best_bin;
            Prim* m = std::partition(&prims[lo], &prims[lo] + n, [=](const Prim& p) {
                return std::min(NBINS - 1, (int)((p.c[ax] - base) * k)) <= bin;
            });
            mid = (int)(m - &prims[0]);
            if (mid == lo || mid == hi) mid = (lo + hi) / 2;
        }
        const size_t me = out.size();
        out.push_back(Node());
        Node nd;
        if (task_cutoff > 0 && n > task_cutoff) {
            std::vector<Node> lsub, rsub;
            int64_t lroot = 0, rroot = 0;
            int ld = depth, rd = depth, ll = 0, rl = 0;
            Box lb, rb;
            #pragma omp task shared(lsub, lroot, lb, ld, ll)
            { lroot = build(lsub, lo, mid, lb, depth + 1, ld, ll); }
            #pragma omp task shared(rsub, rroot, rb, rd, rl)
            { rroot = build(rsub, mid, hi, rb, depth + 1, rd, rl); }
            #pragma omp taskwait
            max_depth = std::max(max_depth, std::max(ld, rd));
            n_leaves += ll + rl;
            nd.lbox = lb; nd.rbox = rb;
            nd.left = append(out, lsub, lroot);
            nd.right = append(out, rsub, rroot);
        } else {
            nd.left = build(out, lo, mid, nd.lbox, depth + 1, max_depth, n_leaves);
            nd.right = build(out, mid, hi, nd.rbox, depth + 1, max_depth, n_leaves);
        }
        out[me] = nd;
        return (int64_t)me;
    }

    static int64_t append(std::vector<Node>& out, const std::vector<Node>& sub, int64_t sub_root) {
        if (sub_root < 0) return sub_root;            // a leaf: nothing to append
        const int64_t base = (int64_t)out.size();
        for (Node n : sub) {
            if (n.left >= 0) n.left += base;
            if (n.right >= 0) n.right += base;
            out.push_back(n);
        }
        return sub_root + base;
    }
};

int env_int(const char* name, int dflt, int lo, int hi) {
    const char* s = std::getenv(name);
    if (!s || !*s) return dflt;
    int v = std::atoi(s);
    return v < lo ? lo : (v > hi ? hi : v);
}

}  // namespace

// tris_raw:  the reference's Triangle_encoded array (36 floats per triangle; p1,p2,p3 first)
// nodes_raw: the reference's BVHNode_encoded array (12 floats per node; node 0 dummy, root 1)
void build_fast_tree(const float* tris_raw, int nTriangles, const float* nodes_raw, int nNodes, FastTree& out, int threads) {
    std::vector<Prim> prims((size_t)nTriangles);
    float radius = 0.0f;
    for (int i = 0; i < nTriangles; i++) {
        const float* t = tris_raw + 36 * (size_t)i;
        Prim& p = prims[i];
        p.box.reset();
        p.box.grow_pt(t); p.box.grow_pt(t + 3); p.box.grow_pt(t + 6);
        for (int a = 0; a < 3; a++) {
            p.c[a] = 0.5f * (p.box.mn[a] + p.box.mx[a]);
            if (std::isfinite(p.box.mn[a])) radius = std::max(radius, std::fabs(p.box.mn[a]));
            if (std::isfinite(p.box.mx[a])) radius = std::max(radius, std::fabs(p.box.mx[a]));
        }
        p.tri = i;
    }
    // reference leaf of every uploaded triangle
    std::vector<int> leaf_of((size_t)nTriangles, 0);
    for (int i = 1; i < nNodes; i++) {
        const float* nd = nodes_raw + 12 * (size_t)i;
        int n = (int)nd[3], index = (int)nd[4];
        for (int k = 0; k < n; k++) leaf_of[(size_t)index + k] = i;
    }
    out.pad = std::ldexp(std::max(radius, 1e-6f), -17);
    out.wnodes.clear();
    out.depth = 1;
    out.n_leaves = 0;

    // Outlier-large triangles (a ground quad under a million-triangle mesh) would be binned by their centre, land
    // deep inside the mesh's subtree and inflate every box on the way up.  They get their own subtree next to the
    // rest under a new root: "large" = box half-area >= 5 % of the scene's.
    Box scene; scene.reset();
    for (const Prim& p : prims) scene.grow(p.box);
    const float large_area = 0.05f * scene.half_area();
    int n_large = 0;
    if (nTriangles > 64 && large_area > 0.0f) {
        Prim* m = std::partition(prims.data(), prims.data() + prims.size(), [=](const Prim& p) { return p.box.half_area() >= large_area; });
        n_large = (int)(m - prims.data());
        if (n_large > 64 || n_large == nTriangles) n_large = 0;   // not outliers: one ordinary tree
    }

    std::vector<Node> nodes;
    nodes.reserve(prims.size());
    Builder b{prims, 0, env_int("PT_LEAF_MAX", 4, 1, 255), 0.01f * (float)env_int("PT_SAH_CNODE_PCT", 100, 1, 10000)};
    Box bounds;
    int64_t root = 0;
    int max_depth = 1, n_leaves = 0;
    auto build_range = [&](int lo, int hi, Box& bb, int depth0) -> int64_t {
        int64_t r = 0;
#ifdef _OPENMP
        if (threads > 1 && hi - lo > 8192) {
            b.task_cutoff = std::max<int>(4096, (hi - lo) / (8 * threads));
            #pragma omp parallel num_threads(threads)
            #pragma omp single
            { r = b.build(nodes, lo, hi, bb, depth0, max_depth, n_leaves); }
            b.task_cutoff = 0;
            return r;
        }
#endif
        return b.build(nodes, lo, hi, bb, depth0, max_depth, n_leaves);
    };
    (void)threads;
    if (n_large > 0) {
        Node top;
        top.left = build_range(0, n_large, top.lbox, 2);
        top.right = build_range(n_large, nTriangles, top.rbox, 2);
        nodes.push_back(top);
        root = (int64_t)nodes.size() - 1;
        bounds = top.lbox; bounds.grow(top.rbox);
    } else {
        root = build_range(0, nTriangles, bounds, 1);
    }
    out.depth = max_depth;
    out.n_leaves = n_leaves;
    out.order.resize((size_t)nTriangles);
    out.leaf_node.resize((size_t)nTriangles);
    for (int i = 0; i < nTriangles; i++) { out.order[i] = prims[i].tri; out.leaf_node[i] = leaf_of[(size_t)prims[i].tri]; }
    const float pad = out.pad;
    out.wnodes.resize(nodes.size() * 16);
    for (size_t i = 0; i < nodes.size(); i++) {
        const Node& n = nodes[i];
        float* w = &out.wnodes[16 * i];
        w[0] = n.lbox.mn[0] - pad; w[1] = n.lbox.mn[1] - pad; w[2] = n.lbox.mn[2] - pad; w[3] = n.lbox.mx[0] + pad;
        w[4] = n.lbox.mx[1] + pad; w[5] = n.lbox.mx[2] + pad; w[6] = n.rbox.mn[0] - pad; w[7] = n.rbox.mn[1] - pad;
        w[8] = n.rbox.mn[2] - pad; w[9] = n.rbox.mx[0] + pad; w[10] = n.rbox.mx[1] + pad; w[11] = n.rbox.mx[2] + pad;
        uint32_t lref = n.left < 0 ? (uint32_t)(~n.left) : (uint32_t)n.left;
        uint32_t rref = n.right < 0 ? (uint32_t)(~n.right) : (uint32_t)n.right;
        std::memcpy(&w[12], &lref, 4); std::memcpy(&w[13], &rref, 4);
        w[14] = w[15] = 0.0f;
    }
    out.root_ref = root < 0 ? (uint32_t)(~root) : (uint32_t)root;   // inner node index (count bits = 0) or a leaf ref
}

}  // namespace pt
