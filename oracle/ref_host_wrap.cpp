/* oracle/ref_host_wrap.cpp — TEST INFRASTRUCTURE (never linked into the product).
 *
 * C-ABI shim around the reference's OWN host functions.  oracle/build_ref.py copies these line
 * ranges of /root/reference/ImportanceSampling_LowDiscrepancySequence/main.cpp into
 * oracle/_ref/ref_host_extract.inc (git-ignored; no reference source is stored in this repo):
 *     19-21 (INF, using namespace glm), 26-72 (Material/Triangle/BVHNode + encoded structs),
 *     153-167 (cmpx/cmpy/cmpz), 243-259 (getTransformMatrix), 262-373 (readObj),
 *     377-552 (buildBVH, buildBVHwithSAH), 555-652 (calculateHdrCache)
 * and compiles hdrloader.cpp with its LP64 bug fixed (`%ld` into an int -> `%d`, hdrloader.cpp:67).
 * The GL-dependent remainder of main.cpp cannot be built here (no graphics stack), so the camera
 * lines of display() (main.cpp:671-674) and the encode loops (main.cpp:793-821) are re-stated below
 * against the same glm the reference vendors.
 *
 * libref_host.so is what pins opengl_ray_tracing_b200/csrc/scene/* (the product's own bit-exact
 * re-implementation of these functions): see tests/test_scene_vs_ref.py and tests/golden/.
 */
#include <iostream>
#include <iomanip>
#include <string>
#include <fstream>
#include <vector>
#include <sstream>
#include <algorithm>
#include <cstring>
#include <cstdlib>
#include <glm/glm.hpp>
#include <glm/gtc/matrix_transform.hpp>
#include <glm/gtc/type_ptr.hpp>
#include "hdrloader.h"

typedef float GLfloat;
typedef unsigned int GLuint;

namespace refhost {
#include "ref_host_extract.inc"
}

static_assert(sizeof(refhost::Triangle) == 144, "Triangle is 36 packed floats");
static_assert(sizeof(refhost::Triangle_encoded) == 144, "Triangle_encoded is 36 packed floats");
static_assert(sizeof(refhost::BVHNode_encoded) == 48, "BVHNode_encoded is 12 packed floats");

extern "C" {

/* tris: T x 36 floats in Triangle / Triangle_encoded order (identical memory layout), reordered in
   place.  nodes_out: cap x 12 floats (BVHNode_encoded, node 0 = the dummy of main.cpp:780-786).
   Returns nNodes, or -needed if cap is too small. */
int ref_build_bvh(float* tris, int T, int leaf_n, int use_sah, float* nodes_out, int cap)
{
    using namespace refhost;
    std::vector<Triangle> triangles(T);
    memcpy((void*)triangles.data(), tris, (size_t)T * sizeof(Triangle));
    BVHNode testNode;
    testNode.left = 255; testNode.right = 128; testNode.n = 30; testNode.index = 0;   /* index is uninitialised in the reference */
    testNode.AA = glm::vec3(1, 1, 0); testNode.BB = glm::vec3(0, 1, 0);
    std::vector<BVHNode> nodes{ testNode };
    if (use_sah) buildBVHwithSAH(triangles, nodes, 0, (int)triangles.size() - 1, leaf_n);
    else         buildBVH(triangles, nodes, 0, (int)triangles.size() - 1, leaf_n);
    int nNodes = (int)nodes.size();
    memcpy(tris, (void*)triangles.data(), (size_t)T * sizeof(Triangle));
    if (nNodes > cap) return -nNodes;
    for (int i = 0; i < nNodes; i++) {               /* main.cpp:815-821 */
        BVHNode_encoded e;
        e.childs = glm::vec3(nodes[i].left, nodes[i].right, 0);
        e.leafInfo = glm::vec3(nodes[i].n, nodes[i].index, 0);
        e.AA = nodes[i].AA; e.BB = nodes[i].BB;
        memcpy(nodes_out + 12 * (size_t)i, &e, sizeof e);
    }
    return nNodes;
}

void ref_transform_matrix(const float* rot, const float* trans, const float* scale, float* out16)
{
    glm::mat4 m = refhost::getTransformMatrix(glm::vec3(rot[0], rot[1], rot[2]),
                                              glm::vec3(trans[0], trans[1], trans[2]),
                                              glm::vec3(scale[0], scale[1], scale[2]));
    memcpy(out16, glm::value_ptr(m), 64);
}

/* readObj appends to a vector; returns the triangle count and copies up to cap triangles */
int ref_read_obj(const char* path, const float* material18, const float* trans16, int smooth,
                 float* tris_out, int cap)
{
    using namespace refhost;
    Material m;
    memcpy((void*)&m, material18, sizeof(Material));
    glm::mat4 trans;
    memcpy(glm::value_ptr(trans), trans16, 64);
    std::vector<Triangle> triangles;
    readObj(path, triangles, m, trans, smooth != 0);
    int T = (int)triangles.size();
    memcpy(tris_out, (void*)triangles.data(), (size_t)std::min(T, cap) * sizeof(Triangle));
    return T;
}

int ref_hdr_load(const char* path, int* w, int* h, float* out, long cap_floats)
{
    HDRLoaderResult res;
    res.cols = 0; res.width = res.height = 0;
    if (!HDRLoader::load(path, res)) return -1;
    *w = res.width; *h = res.height;
    long n = 3L * res.width * res.height;
    if (out && n <= cap_floats) memcpy(out, res.cols, n * sizeof(float));
    delete[] res.cols;
    return 0;
}

void ref_hdr_cache(const float* hdr, int w, int h, float* out)
{
    float* cache = refhost::calculateHdrCache(const_cast<float*>(hdr), w, h);
    memcpy(out, cache, 3UL * w * h * sizeof(float));
    delete[] cache;
}

/* display(): main.cpp:671-674 */
void ref_camera(float upAngle, float rotateAngle, float r, float* eye3, float* cam16)
{
    using namespace glm;
    vec3 eye = vec3(-sin(radians(rotateAngle)) * cos(radians(upAngle)), sin(radians(upAngle)), cos(radians(rotateAngle)) * cos(radians(upAngle)));
    eye.x *= r; eye.y *= r; eye.z *= r;
    mat4 cameraRotate = lookAt(eye, vec3(0, 0, 0), vec3(0, 1, 0));
    cameraRotate = inverse(cameraRotate);
    memcpy(eye3, value_ptr(eye), 12);
    memcpy(cam16, value_ptr(cameraRotate), 64);
}

} /* extern "C" */
