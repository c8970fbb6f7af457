// pts_scene.cpp — scene container, transforms, camera, OBJ front-end and the encoded arrays.
// Reference: ImportanceSampling_LowDiscrepancySequence/main.cpp ("M:") lines cited per function.
#include "pts_scene.h"

#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>

namespace pts {

static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }

static M4 load_m4(const float* p) {
    M4 m;
    std::memcpy(&m, p, 64);
    return m;
}

// What readObj does once `vertices` and `indices` exist (M:321-372).  bbox[] = maxx,maxy,maxz,minx,miny,minz
// as left by the reference's parsing loop.
static int append_mesh(pts_scene* s, std::vector<V3>& vertices, const std::vector<unsigned>& indices,
                       const float bbox[6], const PtsMaterial& material, const M4& trans, bool smooth) {
    for (unsigned id : indices)
        if (id >= vertices.size()) { set_error("face index out of range"); return PTS_EFORMAT; }
    // M:322-330 unit-normalise by the longest bbox axis
    float lenx = bbox[0] - bbox[3], leny = bbox[1] - bbox[4], lenz = bbox[2] - bbox[5];
    float m2 = (leny < lenz) ? lenz : leny;           // glm::max(a,b) = (a<b)?b:a
    float maxaxis = (lenx < m2) ? m2 : lenx;
    for (auto& v : vertices) { v.x /= maxaxis; v.y /= maxaxis; v.z /= maxaxis; }
    // M:333-337 model matrix
    for (auto& v : vertices) {
        V4 vv = mul(trans, V4{v.x, v.y, v.z, 1.0f});
        v = V3{vv.x, vv.y, vv.z};
    }
    // M:340-349 area-unweighted vertex normals
    std::vector<V3> normals;
    if (smooth) {
        normals.assign(vertices.size(), V3{0, 0, 0});
        for (size_t i = 0; i + 2 < indices.size(); i += 3) {
            V3 p1 = vertices[indices[i]], p2 = vertices[indices[i + 1]], p3 = vertices[indices[i + 2]];
            V3 n = normalize(cross(p2 - p1, p3 - p1));
            normals[indices[i]] = normals[indices[i]] + n;
            normals[indices[i + 1]] = normals[indices[i + 1]] + n;
            normals[indices[i + 2]] = normals[indices[i + 2]] + n;
        }
    }
    // M:352-372
    size_t offset = s->tris.size(), nt = indices.size() / 3;
    s->tris.resize(offset + nt);
    for (size_t k = 0; k < nt; k++) {
        Tri& t = s->tris[offset + k];
        unsigned a = indices[3 * k], b = indices[3 * k + 1], c = indices[3 * k + 2];
        t.p1 = vertices[a]; t.p2 = vertices[b]; t.p3 = vertices[c];
        if (!smooth) {
            V3 n = normalize(cross(t.p2 - t.p1, t.p3 - t.p1));
            t.n1 = n; t.n2 = n; t.n3 = n;
        } else {
            t.n1 = normalize(normals[a]); t.n2 = normalize(normals[b]); t.n3 = normalize(normals[c]);
        }
        t.m = material;
    }
    s->nodes.clear();
    s->nodes_encoded_valid = false;
    return (int)nt;
}

// The running bbox of M:274-299 with the reference's quirk: y and z extents are tracked against maxx/minx.
struct RefBBox {
    float b[6] = {-2147483648.0f, -2147483648.0f, -2147483648.0f, 2147483648.0f, 2147483648.0f, 2147483648.0f};
    static float mx(float a, float c) { return (a < c) ? c : a; }
    static float mn(float a, float c) { return (c < a) ? c : a; }
    void add(float x, float y, float z) {
        b[0] = mx(b[0], x); b[1] = mx(b[0], y); b[2] = mx(b[0], z);
        b[3] = mn(b[3], x); b[4] = mn(b[3], y); b[5] = mn(b[3], z);
    }
};

}  // namespace pts

using namespace pts;

extern "C" {

const char* pts_last_error(void) { return g_err.c_str(); }

void pts_material_default(PtsMaterial* m) {
    std::memset(m, 0, sizeof *m);
    m->baseColor[0] = m->baseColor[1] = m->baseColor[2] = 1.0f;
    m->IOR = 1.0f;
}

void pts_get_transform_matrix(const float r[3], const float t[3], const float sc[3], float out16[16]) {
    M4 unit = identity();
    M4 scl = scale(unit, V3{sc[0], sc[1], sc[2]});
    M4 trn = translate(unit, V3{t[0], t[1], t[2]});
    M4 rot = unit;
    rot = rotate(rot, radians(r[0]), V3{1, 0, 0});
    rot = rotate(rot, radians(r[1]), V3{0, 1, 0});
    rot = rotate(rot, radians(r[2]), V3{0, 0, 1});
    M4 model = mul(mul(trn, rot), scl);              // translate * rotate * scale, M:257
    std::memcpy(out16, &model, 64);
}

void pts_camera(float upAngle, float rotateAngle, float r, float eye3[3], float cam16[16]) {
    V3 eye{-std::sin(radians(rotateAngle)) * std::cos(radians(upAngle)), std::sin(radians(upAngle)),
           std::cos(radians(rotateAngle)) * std::cos(radians(upAngle))};
    eye.x *= r; eye.y *= r; eye.z *= r;
    M4 cam = inverse(lookAt(eye, V3{0, 0, 0}, V3{0, 1, 0}));
    std::memcpy(eye3, &eye, 12);
    std::memcpy(cam16, &cam, 64);
}

pts_scene* pts_scene_create(void) { return new (std::nothrow) pts_scene(); }
void pts_scene_destroy(pts_scene* s) { delete s; }

int pts_read_obj(pts_scene* s, const char* path, const PtsMaterial* m, const float trans16[16], int smooth) {
    if (!s || !path || !m || !trans16) { set_error("pts_read_obj: null argument"); return PTS_EINVAL; }
    std::ifstream fin(path);
    if (!fin.is_open()) { set_error(std::string("obj file ") + path + " open failed"); return PTS_EIO; }
    std::vector<V3> vertices;
    std::vector<unsigned> indices;
    RefBBox bb;
    std::string line;
    while (std::getline(fin, line)) {                 // M:281-319
        std::istringstream sin(line);
        std::string type;
        float x, y, z;
        int v0 = 0, v1 = 0, v2 = 0, vt, vn;
        char slash;
        int slashCnt = 0;
        for (char ch : line) if (ch == '/') slashCnt++;
        sin >> type;
        if (type == "v") {
            sin >> x >> y >> z;
            if (sin.fail()) { set_error("malformed vertex line: " + line); return PTS_EFORMAT; }
            vertices.push_back(V3{x, y, z});
            bb.add(x, y, z);
        }
        if (type == "f") {
            if (slashCnt == 6) {
                sin >> v0 >> slash >> vt >> slash >> vn;
                sin >> v1 >> slash >> vt >> slash >> vn;
                sin >> v2 >> slash >> vt >> slash >> vn;
            } else if (slashCnt == 3) {
                sin >> v0 >> slash >> vt; sin >> v1 >> slash >> vt; sin >> v2 >> slash >> vt;
            } else {
                sin >> v0 >> v1 >> v2;
            }
            // the reference reads garbage on a malformed face (e.g. "1//2"); that is UB, so reject it
            if (sin.fail() || v0 < 1 || v1 < 1 || v2 < 1) { set_error("unsupported face line: " + line); return PTS_EFORMAT; }
            indices.push_back((unsigned)(v0 - 1));
            indices.push_back((unsigned)(v1 - 1));
            indices.push_back((unsigned)(v2 - 1));
        }
    }
    return append_mesh(s, vertices, indices, bb.b, *m, load_m4(trans16), smooth != 0);
}

int pts_add_mesh(pts_scene* s, const float* vx, int nv, const unsigned* idx, int ni, const PtsMaterial* m,
                 const float trans16[16], int smooth) {
    if (!s || !vx || !idx || !m || !trans16 || nv <= 0 || ni < 3) { set_error("pts_add_mesh: bad argument"); return PTS_EINVAL; }
    std::vector<V3> vertices((size_t)nv);
    RefBBox bb;
    for (int i = 0; i < nv; i++) {
        vertices[i] = V3{vx[3 * i], vx[3 * i + 1], vx[3 * i + 2]};
        bb.add(vx[3 * i], vx[3 * i + 1], vx[3 * i + 2]);
    }
    std::vector<unsigned> indices(idx, idx + (ni / 3) * 3);
    return append_mesh(s, vertices, indices, bb.b, *m, load_m4(trans16), smooth != 0);
}

int pts_add_triangles(pts_scene* s, const float* tris, int n) {
    if (!s || !tris || n < 0) { set_error("pts_add_triangles: bad argument"); return PTS_EINVAL; }
    size_t off = s->tris.size();
    s->tris.resize(off + (size_t)n);
    std::memcpy((void*)(s->tris.data() + off), tris, (size_t)n * sizeof(Tri));
    s->nodes.clear();
    s->nodes_encoded_valid = false;
    return n;
}

int pts_triangle_count(const pts_scene* s) { return s ? (int)s->tris.size() : PTS_EINVAL; }
int pts_node_count(const pts_scene* s) { return s ? (int)s->nodes.size() : PTS_EINVAL; }

int pts_build_bvh(pts_scene* s, int leaf_n, int use_sah, int n_threads) {
    if (!s || leaf_n < 1) { set_error("pts_build_bvh: bad argument"); return PTS_EINVAL; }
    if (s->tris.size() >= (1u << 24)) { set_error("triangle index no longer exact in the float encoding (>= 2^24)"); return PTS_ERANGE; }
    s->nodes_encoded_valid = false;
    int n = build_bvh(s->tris, s->nodes, leaf_n, use_sah != 0, n_threads);
    if (n >= (1 << 24)) { set_error("node index no longer exact in the float encoding (>= 2^24)"); return PTS_ERANGE; }
    return n;
}

const float* pts_triangles_encoded(const pts_scene* s) {
    return s && !s->tris.empty() ? reinterpret_cast<const float*>(s->tris.data()) : nullptr;
}

const float* pts_nodes_encoded(const pts_scene* s) {     // M:815-821
    if (!s || s->nodes.empty()) return nullptr;
    if (!s->nodes_encoded_valid) {
        s->nodes_encoded.resize(s->nodes.size() * PTS_NODE_FLOATS);
        float* o = s->nodes_encoded.data();
        for (const Node& nd : s->nodes) {
            o[0] = (float)nd.left; o[1] = (float)nd.right; o[2] = 0.0f;
            o[3] = (float)nd.n; o[4] = (float)nd.index; o[5] = 0.0f;
            o[6] = nd.AA.x; o[7] = nd.AA.y; o[8] = nd.AA.z;
            o[9] = nd.BB.x; o[10] = nd.BB.y; o[11] = nd.BB.z;
            o += PTS_NODE_FLOATS;
        }
        s->nodes_encoded_valid = true;
    }
    return s->nodes_encoded.data();
}

// ---- binary cache: "PTSC" u32 version, u64 nTris, u64 nNodes, tris, nodes -----------------------
int pts_scene_save(const pts_scene* s, const char* path) {
    if (!s || !path) return PTS_EINVAL;
    FILE* f = std::fopen(path, "wb");
    if (!f) { set_error(std::string("cannot write ") + path); return PTS_EIO; }
    unsigned magic = 0x43535450u, ver = 1;
    unsigned long long nt = s->tris.size(), nn = s->nodes.size();
    bool ok = std::fwrite(&magic, 4, 1, f) == 1 && std::fwrite(&ver, 4, 1, f) == 1 &&
              std::fwrite(&nt, 8, 1, f) == 1 && std::fwrite(&nn, 8, 1, f) == 1 &&
              (nt == 0 || std::fwrite(s->tris.data(), sizeof(Tri), nt, f) == nt) &&
              (nn == 0 || std::fwrite(s->nodes.data(), sizeof(Node), nn, f) == nn);
    std::fclose(f);
    if (!ok) { set_error("short write"); return PTS_EIO; }
    return PTS_OK;
}

int pts_scene_load(pts_scene* s, const char* path) {
    if (!s || !path) return PTS_EINVAL;
    FILE* f = std::fopen(path, "rb");
    if (!f) { set_error(std::string("cannot read ") + path); return PTS_EIO; }
    unsigned magic = 0, ver = 0;
    unsigned long long nt = 0, nn = 0;
    bool ok = std::fread(&magic, 4, 1, f) == 1 && std::fread(&ver, 4, 1, f) == 1 &&
              std::fread(&nt, 8, 1, f) == 1 && std::fread(&nn, 8, 1, f) == 1 &&
              magic == 0x43535450u && ver == 1 && nt < (1ull << 24) && nn < (1ull << 24);
    if (ok) {
        s->tris.resize(nt); s->nodes.resize(nn);
        ok = (nt == 0 || std::fread((void*)s->tris.data(), sizeof(Tri), nt, f) == nt) &&
             (nn == 0 || std::fread((void*)s->nodes.data(), sizeof(Node), nn, f) == nn);
    }
    std::fclose(f);
    s->nodes_encoded_valid = false;
    if (!ok) { set_error("not a ptscene cache file or truncated"); s->tris.clear(); s->nodes.clear(); return PTS_EFORMAT; }
    return PTS_OK;
}

void pts_free(void* p) { std::free(p); }

}  // extern "C"
