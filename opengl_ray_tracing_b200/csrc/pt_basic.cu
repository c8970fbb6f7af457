// The reference's CPU tracer (BasicRayTracingWithC++/main.cpp, "B:") on the GPU — SURVEY.md §8f.4, BASELINE.json config 1.
//
// B: is a recursive path tracer over a std::vector<Shape*> (3 spheres + 14 triangles in its literal scene), brute-force
// closest hit, Russian roulette with P = 0.8, depth limit 8, a diffuse / mirror / refract choice per hit, and a per-pixel
// SUM over SAMPLE passes kept in doubles.  Its random numbers come from an mt19937 seeded by random_device and shared by
// 50 racing threads, so the reference's OUTPUT is not reproducible; its ARITHMETIC is.  This file restates that arithmetic
// operation for operation (glm 0.9.9.8's float formulas, the double promotions of std::pow / mix / uniform_real) and draws
// its random numbers from a counter-based stream keyed by (seed, sample, row, column), in the order the reference's code
// consumes them when g++ compiles it (the three randf() of `vec3(randf(), randf(), randf())`, B:221, are evaluated last
// argument first).  oracle/_ref/libref_basic_det.so is the reference's own source with the same stream plugged into its
// randf(): tests/test_basic.py requires the two images to be identical bit for bit.
//
// Compiled with -fmad=false: every * and + below is its own IEEE operation, as in the reference's x86-64 build.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/pt.h"

namespace {

struct F3 { float x, y, z; };
__host__ __device__ inline F3 f3(float x, float y, float z) { F3 v; v.x = x; v.y = y; v.z = z; return v; }
__host__ __device__ inline F3 operator+(F3 a, F3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ inline F3 operator-(F3 a, F3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ inline F3 operator-(F3 a) { return f3(-a.x, -a.y, -a.z); }
__host__ __device__ inline F3 operator*(F3 a, F3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
__host__ __device__ inline F3 operator*(F3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
__host__ __device__ inline F3 operator*(float s, F3 a) { return f3(s * a.x, s * a.y, s * a.z); }
__host__ __device__ inline F3 operator/(F3 a, float s) { return f3(a.x / s, a.y / s, a.z / s); }
// glm::dot (tmp = a * b; tmp.x + tmp.y + tmp.z), cross, inversesqrt = 1 / sqrt, normalize = v * inversesqrt(dot(v, v))
__host__ __device__ inline float dot3(F3 a, F3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__host__ __device__ inline F3 cross3(F3 x, F3 y) { return f3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
__host__ __device__ inline F3 normalize3(F3 v) { return v * (1.0f / sqrtf(dot3(v, v))); }
__host__ __device__ inline float length3(F3 v) { return sqrtf(dot3(v, v)); }
__host__ __device__ inline F3 reflect3(F3 I, F3 N) { return I - N * dot3(N, I) * 2.0f; }
__host__ __device__ inline F3 refract3(F3 I, F3 N, float eta) {
    const float d = dot3(N, I);
    const float k = 1.0f - eta * eta * (1.0f - d * d);
    return k >= 0.0f ? (eta * I - (eta * d + sqrtf(k)) * N) : f3(0.0f, 0.0f, 0.0f);
}
// glm::mix(vec3, vec3, double a): evaluated in double, vec3(dvec3(x) * (1 - a) + dvec3(y) * a)
__host__ __device__ inline F3 mix3(F3 x, F3 y, double a) {
    const double w = 1.0 - a;
    return f3((float)((double)x.x * w + (double)y.x * a), (float)((double)x.y * w + (double)y.y * a), (float)((double)x.z * w + (double)y.z * a));
}

struct Shape {                 // device copy of a PtBasicShape, triangle normal precomputed as the constructor does (B:82-86)
    int kind, emissive;
    F3 p1, p2, p3, normal, color;
    double R, specularRate, roughness, refractRate, refractAngle, refractRoughness;
};
#define PT_BASIC_MAX_SHAPES 64
__constant__ Shape c_shapes[PT_BASIC_MAX_SHAPES];

struct Hit { bool isHit; double distance; F3 P, normal; int shape; };

__device__ inline uint32_t wang(uint32_t s) {
    s = (s ^ 61u) ^ (s >> 16);
    s *= 9u;
    s = s ^ (s >> 4);
    s *= 0x27d4eb2du;
    s = s ^ (s >> 15);
    return s;
}
struct Rng {
    uint32_t s;
    __device__ void seed(uint32_t seed, int k, int i, int j) {
        s = wang(seed ^ ((uint32_t)k * 0x9e3779b1u));
        s = wang(s ^ ((uint32_t)i * 0x85ebca77u));
        s = wang(s ^ ((uint32_t)j * 0xc2b2ae3du));
    }
    __device__ double next() { s = wang(s); return (double)s * (1.0 / 4294967296.0); }   // randf(), B:211-214
};

// Triangle::intersect, B:91-123
__device__ inline bool hit_triangle(const Shape& sh, F3 S, F3 d, Hit& res) {
    F3 N = sh.normal;
    if (dot3(N, d) > 0.0f) N = -N;
    if (fabsf(dot3(N, d)) < 0.00001f) return false;
    const float t = (dot3(N, sh.p1) - dot3(S, N)) / dot3(d, N);
    if (t < 0.0005f) return false;
    const F3 P = S + d * t;
    const F3 c1 = cross3(sh.p2 - sh.p1, P - sh.p1), c2 = cross3(sh.p3 - sh.p2, P - sh.p2), c3 = cross3(sh.p1 - sh.p3, P - sh.p3);
    const F3 n = sh.normal;
    if (dot3(c1, n) < 0 || dot3(c2, n) < 0 || dot3(c3, n) < 0) return false;
    res.isHit = true; res.distance = t; res.P = P; res.normal = N;
    return true;
}
// Sphere::intersect, B:137-166 — std::pow(float, int) and the sqrt around it are double
__device__ inline bool hit_sphere(const Shape& sh, F3 S, F3 d, Hit& res) {
    const F3 O = sh.p1;
    const float OS = length3(O - S);
    const float SH = dot3(O - S, d);
    const float OH = (float)sqrt((double)OS * (double)OS - (double)SH * (double)SH);
    if (OH > sh.R) return false;
    const float PH = (float)sqrt(sh.R * sh.R - (double)OH * (double)OH);
    const float t1 = fabsf(SH) - PH;
    const float t2 = fabsf(SH) + PH;
    const float t = (t1 < 0) ? t2 : t1;
    const F3 P = S + t * d;
    if (fabsf(t1) < 0.0005f || fabsf(t2) < 0.0005f) return false;
    res.isHit = true; res.distance = t; res.P = P; res.normal = normalize3(P - O);
    return true;
}
// shoot, B:192-205
__device__ inline Hit shoot(int n_shapes, F3 S, F3 d) {
    Hit res; res.isHit = false; res.distance = (double)1145141919.810f; res.shape = -1;
    for (int s = 0; s < n_shapes; s++) {
        Hit r; r.isHit = false;
        const bool h = c_shapes[s].kind == 1 ? hit_sphere(c_shapes[s], S, d, r) : hit_triangle(c_shapes[s], S, d, r);
        if (h && r.distance < res.distance) { res = r; res.shape = s; }
    }
    return res;
}
// randomVec3 (B:216-233) + randomDirection (B:236-249); g++ evaluates vec3(randf(), randf(), randf()) last argument first
__device__ inline F3 random_direction(Rng& rng, F3 n) {
    F3 d;
    do {
        const double c = rng.next(), b = rng.next(), a = rng.next();
        d = 2.0f * f3((float)a, (float)b, (float)c) - f3(1, 1, 1);
    } while ((double)dot3(d, d) > 1.0);
    return normalize3(normalize3(d) + n);
}

struct Level { float cosine; int diffuse; F3 src; };

// what a hit does next (B:266-293 and the same block in main, B:392-417): returns the next ray direction, sets `diffuse`
__device__ inline F3 scatter(Rng& rng, const Shape& m, F3 in_dir, F3 normal, bool rr_first, bool& absorbed, int& diffuse) {
    absorbed = false;
    if (rr_first) {                                            // pathTracing only: Russian roulette before the bounce (B:260-263)
        const double r = rng.next();
        const float P = 0.8f;
        if (r > P) { absorbed = true; return f3(0, 0, 0); }
    }
    const F3 rnd = random_direction(rng, normal);
    const double r = rng.next();
    diffuse = 0;
    if (r < m.specularRate) {
        const F3 ref = normalize3(reflect3(in_dir, normal));
        return mix3(ref, rnd, m.roughness);
    } else if (m.specularRate <= r && r <= m.refractRate) {
        const F3 ref = normalize3(refract3(in_dir, normal, (float)m.refractAngle));
        return mix3(ref, -rnd, m.refractRoughness);
    }
    diffuse = 1;
    return rnd;
}

// pathTracing(shapes, ray, 0), B:251-296, with the recursion unrolled: walk down recording (cosine, branch, colour), then
// apply `* cosine [* srcColor] / P` from the deepest level up — the order the returns execute in
__device__ inline F3 path_tracing(int n_shapes, Rng& rng, F3 S, F3 d) {
    Level lv[9];
    int depth = 0;
    F3 value = f3(0, 0, 0);
    for (;;) {
        if (depth > 8) break;
        const Hit res = shoot(n_shapes, S, d);
        if (!res.isHit) break;
        const Shape& m = c_shapes[res.shape];
        if (m.emissive) { value = m.color; break; }
        bool absorbed; int diffuse;
        const F3 next = scatter(rng, m, d, res.normal, true, absorbed, diffuse);
        if (absorbed) break;
        lv[depth].cosine = fabsf(dot3(-d, res.normal));
        lv[depth].diffuse = diffuse;
        lv[depth].src = m.color;
        S = res.P; d = next;
        depth++;
    }
    const float P = 0.8f;
    for (int k = depth - 1; k >= 0; k--) {
        value = value * lv[k].cosine;
        if (lv[k].diffuse) value = value * lv[k].src;
        value = value / P;
    }
    return value;
}

// one thread = one (sample k, pixel) path of main's loop nest, B:361-426; colours[k_local][pixel]
__global__ void __launch_bounds__(128) k_basic_paths(int n_shapes, int width, int height, int samples_total, int k0, int nk, uint32_t seed, float* __restrict__ colours) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t npx = (size_t)width * height;
    if (idx >= npx * (size_t)nk) return;
    const int kl = (int)(idx / npx);
    const size_t pix = idx - (size_t)kl * npx;
    const int i = (int)(pix / (size_t)width), j = (int)(pix - (size_t)i * width);
    Rng rng;
    rng.seed(seed, k0 + kl, i, j);
    double x = 2.0 * double(j) / double(width) - 1.0;
    double y = 2.0 * double(height - i) / double(height) - 1.0;
    x += (rng.next() - 0.5f) / double(width);
    y += (rng.next() - 0.5f) / double(height);
    const F3 coord = f3((float)x, (float)y, (float)1.1);                       // SCREEN_Z, B:27
    const F3 dir = normalize3(coord - f3(0, 0, 4.0f));                         // EYE, B:28
    const Hit res = shoot(n_shapes, coord, dir);
    F3 color = f3(0, 0, 0);
    if (res.isHit) {
        const Shape& m = c_shapes[res.shape];
        if (m.emissive) {
            color = m.color;
        } else {
            bool absorbed; int diffuse;
            const F3 next = scatter(rng, m, dir, res.normal, false, absorbed, diffuse);
            color = path_tracing(n_shapes, rng, res.P, next);
            if (diffuse) color = color * m.color;
            const double brightness = (2.0f * 3.1415926f) * (1.0f / double(samples_total));   // B:20
            color = color * (float)brightness;                                 // vec3 *= double casts the scalar to float
        }
    }
    float* o = colours + 3 * idx;
    o[0] = color.x; o[1] = color.y; o[2] = color.z;
}

// image[pixel] += colour of sample k, k in order: the reference's `*p += color.x` in doubles (B:421-423)
__global__ void k_basic_accumulate(const float* __restrict__ colours, size_t npx, int nk, double* __restrict__ image) {
    const size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= npx * 3) return;
    double s = image[c];
    for (int k = 0; k < nk; k++) s += (double)colours[(size_t)k * npx * 3 + c];
    image[c] = s;
}

thread_local std::string g_basic_error;

}  // namespace

extern "C" {

const char* pt_basic_last_error(void) { return g_basic_error.c_str(); }

// the scene literal of main(), B:304-353
int pt_basic_default_scene(PtBasicShape* out, int capacity) {
    std::vector<PtBasicShape> v;
    auto base = [](int kind, const float col[3]) {
        PtBasicShape s;
        std::memset(&s, 0, sizeof s);
        s.kind = kind;
        s.color[0] = col[0]; s.color[1] = col[1]; s.color[2] = col[2];
        s.roughness = 1.0f; s.refractAngle = 1.0f;            // Material's member initialisers, B:50-58
        return s;
    };
    const float RED[3] = {1, 0.5f, 0.5f}, GREEN[3] = {0.5f, 1, 0.5f}, BLUE[3] = {0.5f, 0.5f, 1}, YELLOW[3] = {1.0f, 1.0f, 0.1f},
                CYAN[3] = {0.1f, 1.0f, 1.0f}, WHITE[3] = {1, 1, 1};
    (void)RED; (void)GREEN;
    auto sphere = [&](double ox, double oy, double oz, double r, const float col[3]) {
        PtBasicShape s = base(1, col);
        s.p1[0] = (float)ox; s.p1[1] = (float)oy; s.p1[2] = (float)oz; s.radius = r;
        return s;
    };
    auto tri = [&](double ax, double ay, double az, double bx, double by, double bz, double cx, double cy, double cz, const float col[3]) {
        PtBasicShape s = base(0, col);
        s.p1[0] = (float)ax; s.p1[1] = (float)ay; s.p1[2] = (float)az;
        s.p2[0] = (float)bx; s.p2[1] = (float)by; s.p2[2] = (float)bz;
        s.p3[0] = (float)cx; s.p3[1] = (float)cy; s.p3[2] = (float)cz;
        return s;
    };
    PtBasicShape s1 = sphere(-0.65, -0.7, 0.0, 0.3, GREEN), s2 = sphere(0.0, -0.3, 0.0, 0.4, WHITE), s3 = sphere(0.65, 0.1, 0.0, 0.3, BLUE);
    s1.specularRate = 0.3; s1.roughness = 0.1;
    s2.specularRate = 0.3; s2.refractRate = 0.95; s2.refractAngle = 0.1;
    s3.specularRate = 0.3;
    v.push_back(s1); v.push_back(s2); v.push_back(s3);
    v.push_back(tri(-0.15, 0.4, -0.6, -0.15, -0.95, -0.6, 0.15, 0.4, -0.6, YELLOW));
    v.push_back(tri(0.15, 0.4, -0.6, -0.15, -0.95, -0.6, 0.15, -0.95, -0.6, YELLOW));
    PtBasicShape l1 = tri(0.4, 0.99, 0.4, -0.4, 0.99, -0.4, -0.4, 0.99, 0.4, WHITE), l2 = tri(0.4, 0.99, 0.4, 0.4, 0.99, -0.4, -0.4, 0.99, -0.4, WHITE);
    l1.isEmissive = 1; l2.isEmissive = 1;
    v.push_back(l1); v.push_back(l2);
    v.push_back(tri(1, -1, 1, -1, -1, -1, -1, -1, 1, WHITE));   v.push_back(tri(1, -1, 1, 1, -1, -1, -1, -1, -1, WHITE));    // bottom
    v.push_back(tri(1, 1, 1, -1, 1, 1, -1, 1, -1, WHITE));      v.push_back(tri(1, 1, 1, -1, 1, -1, 1, 1, -1, WHITE));       // top
    v.push_back(tri(1, -1, -1, -1, 1, -1, -1, -1, -1, CYAN));   v.push_back(tri(1, -1, -1, 1, 1, -1, -1, 1, -1, CYAN));      // back
    v.push_back(tri(-1, -1, -1, -1, 1, 1, -1, -1, 1, BLUE));    v.push_back(tri(-1, -1, -1, -1, 1, -1, -1, 1, 1, BLUE));     // left
    v.push_back(tri(1, 1, 1, 1, -1, -1, 1, -1, 1, RED));        v.push_back(tri(1, -1, -1, 1, 1, 1, 1, 1, -1, RED));         // right
    if (out) for (int k = 0; k < (int)v.size() && k < capacity; k++) out[k] = v[k];
    return (int)v.size();
}

int pt_basic_render(int device, const PtBasicShape* shapes, int n_shapes, int width, int height, int samples, unsigned seed,
                    double* image, double* device_ms) {
    auto fail = [](int code, const std::string& m) { g_basic_error = m; return code; };
    if (!shapes || !image || n_shapes <= 0 || n_shapes > PT_BASIC_MAX_SHAPES || width <= 0 || height <= 0 || samples <= 0)
        return fail(PT_EINVAL, "pt_basic_render: bad argument (1..64 shapes, positive sizes)");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || device < 0 || device >= ndev) return fail(PT_ECUDA, "pt_basic_render: no such CUDA device (this library has no CPU fallback)");
    cudaSetDevice(device);
    std::vector<Shape> dev(n_shapes);
    for (int k = 0; k < n_shapes; k++) {
        const PtBasicShape& s = shapes[k];
        Shape& d = dev[k];
        d.kind = s.kind; d.emissive = s.isEmissive;
        d.p1 = f3(s.p1[0], s.p1[1], s.p1[2]); d.p2 = f3(s.p2[0], s.p2[1], s.p2[2]); d.p3 = f3(s.p3[0], s.p3[1], s.p3[2]);
        d.color = f3(s.color[0], s.color[1], s.color[2]);
        d.normal = f3(0, 0, 0);
        if (s.kind == 0) {                                     // Triangle's constructor, B:82-86, in separate float operations
            const F3 a = d.p2 - d.p1, b = d.p3 - d.p1;
            volatile float cx = a.y * b.z, cx2 = b.y * a.z, cy = a.z * b.x, cy2 = b.z * a.x, cz = a.x * b.y, cz2 = b.x * a.y;
            const F3 c = f3(cx - cx2, cy - cy2, cz - cz2);
            volatile float xx = c.x * c.x, yy = c.y * c.y, zz = c.z * c.z;
            volatile float s2 = xx + yy;
            volatile float s3 = s2 + zz;
            const float inv = 1.0f / std::sqrt((float)s3);
            d.normal = f3(c.x * inv, c.y * inv, c.z * inv);
        }
        d.R = s.radius; d.specularRate = s.specularRate; d.roughness = s.roughness; d.refractRate = s.refractRate;
        d.refractAngle = s.refractAngle; d.refractRoughness = s.refractRoughness;
    }
    const size_t npx = (size_t)width * height;
    const int chunk = (int)std::max<size_t>(1, std::min<size_t>((size_t)samples, ((size_t)96 << 20) / (npx * 12)));   // <= 96 MiB of colours in flight
    float* colours = nullptr;
    double* d_image = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    e = cudaMemcpyToSymbol(c_shapes, dev.data(), sizeof(Shape) * n_shapes);
    if (e == cudaSuccess) e = cudaMalloc(&colours, npx * 3 * sizeof(float) * chunk);
    if (e == cudaSuccess) e = cudaMalloc(&d_image, npx * 3 * sizeof(double));
    if (e == cudaSuccess) e = cudaMemset(d_image, 0, npx * 3 * sizeof(double));
    if (e == cudaSuccess) e = cudaEventCreate(&ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&ev1);
    if (e == cudaSuccess) {
        cudaEventRecord(ev0, 0);
        for (int k0 = 0; k0 < samples; k0 += chunk) {
            const int nk = std::min(chunk, samples - k0);
            const size_t n = npx * (size_t)nk;
            k_basic_paths<<<(unsigned)((n + 127) / 128), 128>>>(n_shapes, width, height, samples, k0, nk, seed, colours);
            k_basic_accumulate<<<(unsigned)((npx * 3 + 255) / 256), 256>>>(colours, npx, nk, d_image);
        }
        cudaEventRecord(ev1, 0);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaMemcpy(image, d_image, npx * 3 * sizeof(double), cudaMemcpyDeviceToHost);
        if (e == cudaSuccess && device_ms) { float ms = 0; cudaEventElapsedTime(&ms, ev0, ev1); *device_ms = ms; }
    }
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    cudaFree(colours); cudaFree(d_image);
    if (e != cudaSuccess) return fail(PT_ECUDA, std::string("pt_basic_render: ") + cudaGetErrorString(e));
    return PT_OK;
}

// imshow's conversion, B:169-190: (unsigned char)clamp(pow(S, 1.0f / 2.2f) * 255, 0, 255) per channel, rows as stored (top first)
int pt_basic_imshow(const double* image, int width, int height, unsigned char* rgb8) {
    if (!image || !rgb8 || width <= 0 || height <= 0) { g_basic_error = "pt_basic_imshow: bad argument"; return PT_EINVAL; }
    const size_t n = (size_t)width * height * 3;
    for (size_t k = 0; k < n; k++) {
        double v = std::pow(image[k], 1.0f / 2.2f) * 255;
        v = v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v);           // glm::clamp = min(max(x, lo), hi)
        rgb8[k] = (unsigned char)v;
    }
    return PT_OK;
}

}  // extern "C"
