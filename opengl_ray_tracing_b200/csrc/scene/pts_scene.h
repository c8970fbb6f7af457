// pts_scene.h — internal types of libptscene (not installed; the public surface is include/ptscene.h)
#pragma once
#include <string>
#include <vector>
#include "../../../include/ptscene.h"
#include "pts_math.h"

namespace pts {

// struct Triangle of the reference (M:44-48) is p1,p2,p3,n1,n2,n3 + Material = 36 packed floats, which is
// also exactly Triangle_encoded (M:50-60): the "encode" loop M:793-812 is a field-by-field copy.
struct Tri {
    V3 p1, p2, p3;
    V3 n1, n2, n3;
    PtsMaterial m;
};
static_assert(sizeof(Tri) == 144, "Tri must be 36 packed floats");

// struct BVHNode, M:62-66
struct Node {
    int left, right;
    int n, index;
    V3 AA, BB;
};

void set_error(const std::string& msg);

// bvh.cpp — returns the node array INCLUDING the dummy node 0, reorders tris
int build_bvh(std::vector<Tri>& tris, std::vector<Node>& nodes, int leaf_n, bool sah, int n_threads);

}  // namespace pts

struct pts_scene {
    std::vector<pts::Tri> tris;
    std::vector<pts::Node> nodes;              // empty until pts_build_bvh
    mutable std::vector<float> nodes_encoded;  // lazily refreshed
    mutable bool nodes_encoded_valid = false;
};
