// Host-side check of pt_build.cpp (the production traversal tree built at upload) — TEST CODE, compiled by
// tests/test_fast_tree.py against opengl_ray_tracing_b200/csrc/pt_build.cpp.  Reads the reference-layout buffers
// (Triangle_encoded: 36 floats, BVHNode_encoded: 12 floats) from a file, builds the tree, and verifies what the CUDA
// traversal relies on:
//   1. `order` is a permutation of the uploaded triangle indices, and every leaf is a contiguous range of it;
//   2. every triangle is reachable through exactly one leaf;
//   3. each child box, as stored (padded), contains every vertex of every triangle below that child;
//   4. `leaf_node[i]` is the reference LEAF whose [index, index + n) range holds the triangle (rule (1) of DESIGN.md §4);
//   5. leaves hold 1 .. 255 triangles and indices stay below 2^24 (the packed ref format).
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

namespace pt {
struct FastTree {                      // MUST mirror pt_build.cpp
    std::vector<float> wnodes;
    std::vector<int> order;
    std::vector<int> leaf_node;
    uint32_t root_ref = 0;
    int depth = 0;
    int n_leaves = 0;
    float pad = 0.0f;
};
void build_fast_tree(const float* tris_raw, int nTriangles, const float* nodes_raw, int nNodes, FastTree& out, int threads);
}

static int fails = 0;
#define CHECK(c, ...) do { if (!(c)) { if (fails++ < 10) { std::printf("FAIL: " __VA_ARGS__); std::printf("\n"); } } } while (0)

struct Ctx {
    const float* tris; const pt::FastTree* ft; std::vector<int> seen; long leaves = 0; int max_depth = 0;
};

static void visit(Ctx& c, uint32_t ref, const float* box /* 6 floats or null */, int depth) {
    if (depth > c.max_depth) c.max_depth = depth;
    const uint32_t count = ref >> 24, index = ref & 0x00ffffffu;
    if (count) {                                                   // leaf: [index, index + count) in the new order
        c.leaves++;
        CHECK(index + count <= c.ft->order.size(), "leaf range %u+%u beyond %zu", index, count, c.ft->order.size());
        for (uint32_t i = index; i < index + count && i < c.ft->order.size(); i++) {
            c.seen[i]++;
            const float* t = c.tris + 36 * (size_t)c.ft->order[i];
            if (box)
                for (int v = 0; v < 3; v++)
                    for (int a = 0; a < 3; a++)
                        CHECK(t[3 * v + a] >= box[a] && t[3 * v + a] <= box[3 + a], "vertex outside its leaf's stored box (tri %d)", c.ft->order[i]);
        }
        return;
    }
    CHECK((size_t)index * 16 + 16 <= c.ft->wnodes.size(), "inner node %u out of range", index);
    if ((size_t)index * 16 + 16 > c.ft->wnodes.size()) return;
    const float* w = &c.ft->wnodes[16 * (size_t)index];
    uint32_t l, r;
    std::memcpy(&l, &w[12], 4); std::memcpy(&r, &w[13], 4);
    const float lb[6] = {w[0], w[1], w[2], w[3], w[4], w[5]}, rb[6] = {w[6], w[7], w[8], w[9], w[10], w[11]};
    if (box)                                                       // a child box never sticks out of its parent's by more than the pad
        for (int a = 0; a < 3; a++) {
            CHECK(lb[a] >= box[a] - 2 * c.ft->pad && lb[3 + a] <= box[3 + a] + 2 * c.ft->pad, "left box outside parent");
            CHECK(rb[a] >= box[a] - 2 * c.ft->pad && rb[3 + a] <= box[3 + a] + 2 * c.ft->pad, "right box outside parent");
        }
    visit(c, l, lb, depth + 1);
    visit(c, r, rb, depth + 1);
}

// every vertex below `ref` inside `box` (checked at EVERY level, not only at the leaves)
static void contain(const Ctx& c, uint32_t ref, const float* box) {
    const uint32_t count = ref >> 24, index = ref & 0x00ffffffu;
    if (count) {
        for (uint32_t i = index; i < index + count && i < c.ft->order.size(); i++) {
            const float* t = c.tris + 36 * (size_t)c.ft->order[i];
            for (int v = 0; v < 3; v++)
                for (int a = 0; a < 3; a++)
                    CHECK(t[3 * v + a] >= box[a] && t[3 * v + a] <= box[3 + a], "vertex outside an ancestor's stored box");
        }
        return;
    }
    const float* w = &c.ft->wnodes[16 * (size_t)index];
    uint32_t l, r;
    std::memcpy(&l, &w[12], 4); std::memcpy(&r, &w[13], 4);
    contain(c, l, box); contain(c, r, box);
}

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    FILE* f = std::fopen(argv[1], "rb");
    if (!f) return 2;
    int nT = 0, nN = 0;
    if (std::fread(&nT, 4, 1, f) != 1 || std::fread(&nN, 4, 1, f) != 1) return 2;
    std::vector<float> tris((size_t)nT * 36), nodes((size_t)nN * 12);
    if (std::fread(tris.data(), 4, tris.size(), f) != tris.size() || std::fread(nodes.data(), 4, nodes.size(), f) != nodes.size()) return 2;
    std::fclose(f);
    for (int threads : {1, 4}) {
        pt::FastTree ft;
        pt::build_fast_tree(tris.data(), nT, nodes.data(), nN, ft, threads);
        CHECK((int)ft.order.size() == nT && (int)ft.leaf_node.size() == nT, "order/leaf_node sizes");
        std::vector<int> hit((size_t)nT, 0);
        for (int i : ft.order) { CHECK(i >= 0 && i < nT, "order entry out of range"); if (i >= 0 && i < nT) hit[i]++; }
        for (int i = 0; i < nT; i++) CHECK(hit[i] == 1, "uploaded triangle %d appears %d times in the new order", i, hit[i]);
        Ctx c{tris.data(), &ft, std::vector<int>((size_t)nT, 0)};
        visit(c, ft.root_ref, nullptr, 1);
        for (int i = 0; i < nT; i++) CHECK(c.seen[i] == 1, "position %d is covered by %d leaves", i, c.seen[i]);
        CHECK(c.leaves == ft.n_leaves, "leaf count %ld vs reported %d", c.leaves, ft.n_leaves);
        CHECK(c.max_depth <= ft.depth + 1 && ft.depth < 128, "depth %d vs reported %d", c.max_depth, ft.depth);
        CHECK(ft.pad > 0.0f, "pad");
        // containment at every level (quadratic in depth only: fine at test sizes)
        for (size_t n = 0; n < ft.wnodes.size() / 16; n++) {
            const float* w = &ft.wnodes[16 * n];
            uint32_t l, r;
            std::memcpy(&l, &w[12], 4); std::memcpy(&r, &w[13], 4);
            const float lb[6] = {w[0], w[1], w[2], w[3], w[4], w[5]}, rb[6] = {w[6], w[7], w[8], w[9], w[10], w[11]};
            if (nT <= 20000 || n % 37 == 0) { contain(c, l, lb); contain(c, r, rb); }
        }
        // rule (1): the recorded reference leaf really holds the triangle
        for (int i = 0; i < nT; i++) {
            const int leaf = ft.leaf_node[i], tri = ft.order[i];
            CHECK(leaf >= 1 && leaf < nN, "leaf_node out of range");
            if (leaf < 1 || leaf >= nN) continue;
            const float* nd = &nodes[12 * (size_t)leaf];
            const int n = (int)nd[3], index = (int)nd[4];
            CHECK(n > 0 && tri >= index && tri < index + n, "triangle %d is not in reference leaf %d [%d, %d)", tri, leaf, index, index + n);
        }
        std::printf("threads %d: %d triangles, %zu inner nodes, %ld leaves, depth %d, pad %g\n", threads, nT, ft.wnodes.size() / 16, c.leaves, c.max_depth, ft.pad);
    }
    std::printf(fails ? "FAILED (%d)\n" : "OK\n", fails);
    return fails ? 1 : 0;
}
