// pt_build.cpp — host half of pt_upload_scene: the TRAVERSAL topology.
//
// The reference's builders (buildBVHwithSAH, IS/main.cpp:431-552) give the node array that defines WHICH
// triangles a ray may hit: a triangle counts only if the ray passes hitAABB() > 0 on the box of the reference
// LEAF that holds it (pass1.fsh:356-379; a leaf's box being hit implies every ancestor's box is hit, because the
// slab arithmetic is monotone in the box corners and ancestors are exact unions).  The reference topology above
// the leaves decides only the visiting order, i.e. who wins an exact-distance tie.  On large meshes that topology
// degenerates (its left sweep mixes p2.x into the z extent, IS/main.cpp:481-485): the 1M-triangle scene gets a
// tree 6781 levels deep with ~200 inner nodes on the path of an average ray.
//
// So the kernels traverse a second topology built here over the SAME leaves (same boxes, same triangle ranges,
// same per-leaf triangle order): a binned-SAH binary tree, 16 bins x 3 axes, one reference leaf per traversal
// leaf.  Results are identical by construction except for exact ties between triangles of different leaves,
// which the kernel detects and resolves by re-tracing that ray on the reference topology (pt_trace.cuh).
//
// Output: the same 64-byte wide-node records the device builds for the reference topology (child boxes in the
// parent; packed child refs: count << 24 | index).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace pt {

struct FastTree {
    std::vector<float> wnodes;       // 16 floats per inner node
    uint32_t root_ref = 0;
    int depth = 0;
    int n_leaves = 0;
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
    Box box;               // the reference leaf's box, bit for bit
    float c[3];            // centroid
    uint32_t ref;          // packed (count << 24 | first triangle)
};

const int NBINS = 16;

struct Node {              // build-time node; children < 0 => ~prim index
    Box lbox, rbox;
    int left, right;
};

struct Builder {
    std::vector<Prim>& prims;
    int task_cutoff;

    // builds the subtree over prims[lo, hi) into `out` (indices local to out); returns root index, or ~prim
    int build(std::vector<Node>& out, int lo, int hi, Box& bounds, int depth, int& max_depth) {
        bounds.reset();
        for (int i = lo; i < hi; i++) bounds.grow(prims[i].box);
        if (depth > max_depth) max_depth = depth;
        if (hi - lo == 1) return ~lo;
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
            Box acc; acc.reset(); int n = 0;
            for (int b = NBINS - 1; b > 0; b--) { acc.grow(bb[b]); n += cnt[b]; rarea[b] = acc.half_area(); rcnt[b] = n; }
            acc.reset(); n = 0;
            for (int b = 0; b < NBINS - 1; b++) {
                acc.grow(bb[b]); n += cnt[b];
                if (n == 0 || rcnt[b + 1] == 0) continue;
                float cost = acc.half_area() * (float)n + rarea[b + 1] * (float)rcnt[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = axis; best_bin = b; }
            }
        }
        int mid;
        if (best_axis < 0) {
            mid = (lo + hi) / 2;                       // all centroids coincide: split the range in half
        } else {
            float ext = cb.mx[best_axis] - cb.mn[best_axis], k = (float)NBINS / ext, base = cb.mn[best_axis];
            int ax = best_axis, bin = best_bin;
            Prim* m = std::partition(&prims[lo], &prims[lo] + (hi - lo), [=](const Prim& p) {
                return std::min(NBINS - 1, (int)((p.c[ax] - base) * k)) <= bin;
            });
            mid = (int)(m - &prims[0]);
            if (mid == lo || mid == hi) mid = (lo + hi) / 2;
        }
        const size_t me = out.size();
        out.push_back(Node());
        Node nd;
        if (task_cutoff > 0 && hi - lo > task_cutoff) {
            std::vector<Node> lsub, rsub;
            int lroot = 0, rroot = 0, ld = depth, rd = depth;
            Box lb, rb;
            #pragma omp task shared(lsub, lroot, lb, ld)
            { lroot = build(lsub, lo, mid, lb, depth + 1, ld); }
            #pragma omp task shared(rsub, rroot, rb, rd)
            { rroot = build(rsub, mid, hi, rb, depth + 1, rd); }
            #pragma omp taskwait
            max_depth = std::max(max_depth, std::max(ld, rd));
            nd.lbox = lb; nd.rbox = rb;
            nd.left = append(out, lsub, lroot);
            nd.right = append(out, rsub, rroot);
        } else {
            nd.left = build(out, lo, mid, nd.lbox, depth + 1, max_depth);
            nd.right = build(out, mid, hi, nd.rbox, depth + 1, max_depth);
        }
        out[me] = nd;
        return (int)me;
    }

    static int append(std::vector<Node>& out, const std::vector<Node>& sub, int sub_root) {
        if (sub_root < 0) return sub_root;            // a single prim: nothing to append
        const int base = (int)out.size();
        for (Node n : sub) {
            if (n.left >= 0) n.left += base;
            if (n.right >= 0) n.right += base;
            out.push_back(n);
        }
        return sub_root + base;
    }
};

}  // namespace

// nodes_raw: the reference's BVHNode_encoded array (12 floats per node; node 0 dummy, root 1)
void build_fast_tree(const float* nodes_raw, int nNodes, FastTree& out, int threads) {
    std::vector<Prim> prims;
    for (int i = 1; i < nNodes; i++) {
        const float* nd = nodes_raw + 12 * (size_t)i;
        int n = (int)nd[3], index = (int)nd[4];
        if (n <= 0) continue;
        Prim p;
        for (int a = 0; a < 3; a++) { p.box.mn[a] = nd[6 + a]; p.box.mx[a] = nd[9 + a]; p.c[a] = 0.5f * (nd[6 + a] + nd[9 + a]); }
        p.ref = ((uint32_t)n << 24) | (uint32_t)index;
        prims.push_back(p);
    }
    out.n_leaves = (int)prims.size();
    out.wnodes.clear();
    out.depth = 1;
    if (prims.size() == 1) { out.root_ref = prims[0].ref; return; }
    std::vector<Node> nodes;
    nodes.reserve(prims.size());
    Builder b{prims, 0};
    Box bounds;
    int root = 0, max_depth = 1;
#ifdef _OPENMP
    if (threads > 1 && prims.size() > 8192) {
        b.task_cutoff = std::max<int>(4096, (int)prims.size() / (8 * threads));
        #pragma omp parallel num_threads(threads)
        #pragma omp single
        { root = b.build(nodes, 0, (int)prims.size(), bounds, 1, max_depth); }
    } else
#endif
    {
        (void)threads;
        root = b.build(nodes, 0, (int)prims.size(), bounds, 1, max_depth);
    }
    out.depth = max_depth;
    out.wnodes.resize(nodes.size() * 16);
    for (size_t i = 0; i < nodes.size(); i++) {
        const Node& n = nodes[i];
        float* w = &out.wnodes[16 * i];
        w[0] = n.lbox.mn[0]; w[1] = n.lbox.mn[1]; w[2] = n.lbox.mn[2]; w[3] = n.lbox.mx[0];
        w[4] = n.lbox.mx[1]; w[5] = n.lbox.mx[2]; w[6] = n.rbox.mn[0]; w[7] = n.rbox.mn[1];
        w[8] = n.rbox.mn[2]; w[9] = n.rbox.mx[0]; w[10] = n.rbox.mx[1]; w[11] = n.rbox.mx[2];
        uint32_t lref = n.left < 0 ? prims[~n.left].ref : (uint32_t)n.left;
        uint32_t rref = n.right < 0 ? prims[~n.right].ref : (uint32_t)n.right;
        std::memcpy(&w[12], &lref, 4); std::memcpy(&w[13], &rref, 4);
        w[14] = w[15] = 0.0f;
    }
    out.root_ref = (uint32_t)root;                      // inner node index (count bits = 0)
}

}  // namespace pt
