// pts_meshgen.cpp — seeded procedural meshes standing in for the reference's missing models/*.obj
// (.gitignore:77 of the reference ignores *.obj; bunny/teapot/quad are not in the tree).
//
// Every generated mesh goes through pts_add_mesh, i.e. the readObj pipeline after parsing (bbox
// normalisation, model matrix, face/smooth normals: ImportanceSampling_LowDiscrepancySequence/
// main.cpp:321-372), so a generated scene has exactly the shape of a loaded one.  The generators use
// integer hashing and basic binary32 arithmetic only — same triangles on every machine.
#include "pts_scene.h"

#include <cmath>
#include <cstdio>
#include <map>
#include <utility>

namespace pts {
namespace {

inline unsigned hash3(int x, int y, int z, unsigned seed) {
    unsigned h = seed * 0x9e3779b9u + 0x7f4a7c15u;
    h ^= (unsigned)x * 0x85ebca6bu; h = (h << 13) | (h >> 19); h *= 0xc2b2ae35u;
    h ^= (unsigned)y * 0x27d4eb2fu; h = (h << 13) | (h >> 19); h *= 0xc2b2ae35u;
    h ^= (unsigned)z * 0x165667b1u; h = (h << 13) | (h >> 19); h *= 0xc2b2ae35u;
    h ^= h >> 16; h *= 0x7feb352du; h ^= h >> 15; h *= 0x846ca68bu; h ^= h >> 16;
    return h;
}
inline float lattice(int x, int y, int z, unsigned seed) {
    return (float)(hash3(x, y, z, seed) >> 8) * (1.0f / 16777216.0f) * 2.0f - 1.0f;      // [-1, 1)
}
inline float smooth(float t) { return t * t * (3.0f - 2.0f * t); }

float value_noise(float x, float y, float z, unsigned seed) {
    float fx = std::floor(x), fy = std::floor(y), fz = std::floor(z);
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    float tx = smooth(x - fx), ty = smooth(y - fy), tz = smooth(z - fz);
    float c[2][2][2];
    for (int a = 0; a < 2; a++) for (int b = 0; b < 2; b++) for (int d = 0; d < 2; d++)
        c[a][b][d] = lattice(ix + a, iy + b, iz + d, seed);
    float x00 = c[0][0][0] + (c[1][0][0] - c[0][0][0]) * tx, x10 = c[0][1][0] + (c[1][1][0] - c[0][1][0]) * tx;
    float x01 = c[0][0][1] + (c[1][0][1] - c[0][0][1]) * tx, x11 = c[0][1][1] + (c[1][1][1] - c[0][1][1]) * tx;
    float y0 = x00 + (x10 - x00) * ty, y1 = x01 + (x11 - x01) * ty;
    return y0 + (y1 - y0) * tz;
}

float fbm(V3 p, unsigned seed) {
    float amp = 0.5f, freq = 2.0f, sum = 0.0f;
    for (int o = 0; o < 5; o++) {
        sum += amp * value_noise(p.x * freq + 17.0f, p.y * freq + 31.0f, p.z * freq + 47.0f, seed + (unsigned)o);
        amp *= 0.5f;
        freq *= 2.0f;
    }
    return sum;                                        // roughly [-1, 1]
}

// Class-I geodesic subdivision of the icosahedron at frequency n, vertices welded across faces.
void geodesic_icosphere(int n, float amplitude, unsigned seed, std::vector<float>& verts, std::vector<unsigned>& idx) {
    const float t = 1.6180339887498949f;
    const V3 corner[12] = {{-1, t, 0}, {1, t, 0}, {-1, -t, 0}, {1, -t, 0}, {0, -1, t}, {0, 1, t},
                           {0, -1, -t}, {0, 1, -t}, {t, 0, -1}, {t, 0, 1}, {-t, 0, -1}, {-t, 0, 1}};
    const int face[20][3] = {{0, 11, 5}, {0, 5, 1}, {0, 1, 7}, {0, 7, 10}, {0, 10, 11}, {1, 5, 9}, {5, 11, 4},
                             {11, 10, 2}, {10, 7, 6}, {7, 1, 8}, {3, 9, 4}, {3, 4, 2}, {3, 2, 6}, {3, 6, 8},
                             {3, 8, 9}, {4, 9, 5}, {2, 4, 11}, {6, 2, 10}, {8, 6, 7}, {9, 8, 1}};
    std::map<std::pair<int, int>, int> edge_id;
    auto edge_of = [&](int a, int b) {
        auto key = std::make_pair(a < b ? a : b, a < b ? b : a);
        auto it = edge_id.find(key);
        if (it != edge_id.end()) return it->second;
        int id = (int)edge_id.size();
        edge_id[key] = id;
        return id;
    };
    const long per_edge = n - 1, per_face = (long)(n - 1) * (n - 2) / 2;
    const long nv = 12 + 30 * per_edge + 20 * per_face;
    verts.assign((size_t)nv * 3, 0.0f);
    std::vector<unsigned char> done((size_t)nv, 0);
    std::vector<long> row_start((size_t)n + 1, 0);                    // interior rows i = 1..n-2 hold j = 1..n-1-i
    for (int r = 2; r <= n; r++) row_start[(size_t)r] = row_start[(size_t)r - 1] + (n - 1 - (r - 1));

    auto place = [&](long id, V3 a, V3 b, V3 c, int wa, int wb, int wc) {
        if (done[(size_t)id]) return;
        done[(size_t)id] = 1;
        float fa = (float)wa / (float)n, fb = (float)wb / (float)n, fc = (float)wc / (float)n;
        V3 p = a * fa + b * fb + c * fc;
        p = normalize(p);
        float rad = 1.0f + amplitude * fbm(p, seed);
        verts[3 * (size_t)id] = p.x * rad; verts[3 * (size_t)id + 1] = p.y * rad; verts[3 * (size_t)id + 2] = p.z * rad;
    };
    // id of grid point (i, j) of face f: barycentric weights (n-i-j, i, j) on corners (A, B, C)
    auto vid = [&](int f, int i, int j) -> long {
        const int A = face[f][0], B = face[f][1], C = face[f][2];
        const int k = n - i - j;
        if (k == n) return A;
        if (i == n) return B;
        if (j == n) return C;
        auto on_edge = [&](int u, int v, int wu, int wv) -> long {   // wu + wv = n, weight of u and v
            int e = edge_of(u, v);
            int lo = u < v ? u : v;
            int steps = (lo == u) ? wv : wu;                         // steps away from the lower corner
            long id = 12 + (long)e * per_edge + (steps - 1);
            V3 a = corner[lo], b = corner[lo == u ? v : u];
            place(id, a, b, V3{0, 0, 0}, n - steps, steps, 0);      // canonical orientation: identical bits from both faces
            return id;
        };
        if (j == 0) return on_edge(A, B, k, i);
        if (i == 0) return on_edge(A, C, k, j);
        if (k == 0) return on_edge(B, C, i, j);
        long id = 12 + 30 * per_edge + (long)f * per_face + row_start[(size_t)i] + (j - 1);
        place(id, corner[A], corner[B], corner[C], k, i, j);
        return id;
    };
    for (int c = 0; c < 12; c++) place(c, corner[c], V3{0, 0, 0}, V3{0, 0, 0}, n, 0, 0);
    idx.clear();
    idx.reserve((size_t)20 * n * n * 3);
    for (int f = 0; f < 20; f++)
        for (int i = 0; i < n; i++)
            for (int j = 0; j < n - i; j++) {
                long a = vid(f, i, j), b = vid(f, i + 1, j), c = vid(f, i, j + 1);
                idx.push_back((unsigned)a); idx.push_back((unsigned)b); idx.push_back((unsigned)c);
                if (i + j < n - 1) {
                    long d = vid(f, i + 1, j + 1);
                    idx.push_back((unsigned)b); idx.push_back((unsigned)d); idx.push_back((unsigned)c);
                }
            }
}

}  // namespace
}  // namespace pts

using namespace pts;

extern "C" {

int pts_add_icosphere(pts_scene* s, int frequency, float amplitude, unsigned seed, const PtsMaterial* m,
                      const float trans16[16], int smooth) {
    if (!s || !m || !trans16 || frequency < 1 || frequency > 2000) { set_error("pts_add_icosphere: bad argument"); return PTS_EINVAL; }
    std::vector<float> verts;
    std::vector<unsigned> idx;
    geodesic_icosphere(frequency, amplitude, seed, verts, idx);
    return pts_add_mesh(s, verts.data(), (int)(verts.size() / 3), idx.data(), (int)idx.size(), m, trans16, smooth);
}

// models/quad.obj is absent from the reference tree; every use of it (OpenglRayTracing/main.cpp:658-670,
// IS/main.cpp:771) scales y by 0.01 and x,z by the floor size, i.e. it is a unit square in the xz plane.
// Its bbox-normalised extent is 1 in x (the reference's bbox quirk compares y and z against x).
int pts_add_quad(pts_scene* s, const PtsMaterial* m, const float trans16[16]) {
    static const float v[12] = {-0.5f, 0.0f, -0.5f, 0.5f, 0.0f, -0.5f, 0.5f, 0.0f, 0.5f, -0.5f, 0.0f, 0.5f};
    static const unsigned i[6] = {0, 2, 1, 0, 3, 2};             // +y facing
    return pts_add_mesh(s, v, 4, i, 6, m, trans16, 0);
}

int pts_write_obj_icosphere(const char* path, int frequency, float amplitude, unsigned seed) {
    if (!path || frequency < 1 || frequency > 2000) return PTS_EINVAL;
    std::vector<float> verts;
    std::vector<unsigned> idx;
    geodesic_icosphere(frequency, amplitude, seed, verts, idx);
    FILE* f = std::fopen(path, "w");
    if (!f) { set_error(std::string("cannot write ") + path); return PTS_EIO; }
    for (size_t k = 0; k + 2 < verts.size(); k += 3) std::fprintf(f, "v %.9g %.9g %.9g\n", verts[k], verts[k + 1], verts[k + 2]);
    for (size_t k = 0; k + 2 < idx.size(); k += 3) std::fprintf(f, "f %u %u %u\n", idx[k] + 1, idx[k + 1] + 1, idx[k + 2] + 1);
    std::fclose(f);
    return PTS_OK;
}

}  // extern "C"
