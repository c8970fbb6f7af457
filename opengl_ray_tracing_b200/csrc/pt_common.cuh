// pt_common.cuh — device-side types shared by the upload, traversal and shading code of libpt.so.
//
// Arithmetic discipline (DESIGN.md "Arithmetic modes"):
//   * "strict" helpers (s_add, s_mul, …) map to the CUDA _rn intrinsics, which are IEEE-754 binary32,
//     never fused into FMAs and unaffected by fast-math flags.  Ray generation and every ray/box and
//     ray/triangle test use them in ALL modes: that is what makes primary-hit triangle ids bit-exact.
//   * plain operators are used in shading.  The parity translation unit is compiled with -fmad=false,
//     so there they are exactly the oracle's operations; the fast translation unit lets nvcc fuse them.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pt {

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 mk3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(V3 a, V3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ V3 operator-(V3 a) { return mk3(-a.x, -a.y, -a.z); }
// the oracle's definitions of the GLSL built-ins (oracle/glsl_compat.h): same association order
__device__ __forceinline__ float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
    return mk3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
__device__ __forceinline__ V3 normalize(V3 v) { return v * (1.0f / sqrtf(dot(v, v))); }
__device__ __forceinline__ float mixf(float x, float y, float a) { return x * (1.0f - a) + y * a; }
__device__ __forceinline__ V3 mix3(V3 x, V3 y, float a) { return x * (1.0f - a) + y * a; }
__device__ __forceinline__ float sqr(float x) { return x * x; }

// ---- strict (never fused, IEEE) arithmetic ------------------------------------------------------
__device__ __forceinline__ float s_add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float s_sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float s_mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float s_div(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float s_sqrt(float a) { return __fsqrt_rn(a); }
__device__ __forceinline__ float s_rcp(float a) { return __frcp_rn(a); }      // == s_div(1.0f, a): both correctly rounded
__device__ __forceinline__ V3 s_sub3(V3 a, V3 b) { return mk3(s_sub(a.x, b.x), s_sub(a.y, b.y), s_sub(a.z, b.z)); }
__device__ __forceinline__ float s_dot(V3 a, V3 b) {
    return s_add(s_add(s_mul(a.x, b.x), s_mul(a.y, b.y)), s_mul(a.z, b.z));
}
__device__ __forceinline__ V3 s_cross(V3 a, V3 b) {
    return mk3(s_sub(s_mul(a.y, b.z), s_mul(b.y, a.z)), s_sub(s_mul(a.z, b.x), s_mul(b.z, a.x)),
               s_sub(s_mul(a.x, b.y), s_mul(b.x, a.y)));
}
__device__ __forceinline__ V3 s_normalize(V3 v) {
    float k = s_rcp(s_sqrt(s_dot(v, v)));
    return mk3(s_mul(v.x, k), s_mul(v.y, k), s_mul(v.z, k));
}
// S + d*t
__device__ __forceinline__ V3 s_point(V3 S, V3 d, float t) {
    return mk3(s_add(S.x, s_mul(d.x, t)), s_add(S.y, s_mul(d.y, t)), s_add(S.z, s_mul(d.z, t)));
}

// ---- read-only loads: ld.global.nc (the scene is immutable while a render kernel runs) ----------
__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }
__device__ __forceinline__ float ldg1(const float* p) { return __ldg(p); }
// 32 bytes in ONE load (sm_100: LDG.E.256): a 64-byte traversal record is two L1 tag look-ups instead of four.  With the
// lanes of a warp at 32 different records the L1's tag stage, not bandwidth, is what a node step waits for
// (profiles/r02_wavefront.md: l1tex throughput 91 % of peak in the traversal-only kernel).  p must be 32-byte aligned.
#ifndef PT_LD256
#define PT_LD256 1
#endif
__device__ __forceinline__ void ldg8(const float4* p, float4& a, float4& b) {
#if PT_LD256
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
#else
    a = __ldg(p); b = __ldg(p + 1);
#endif
}

// ---- packed child reference of the wide BVH ----------------------------------------------------
// bits 0..23 = index (inner: wide-node index; leaf: first triangle), bits 24..31 = triangle count
// (0 => inner node).  Lossless: the reference's own float encoding caps every index below 2^24.
#define PT_REF_NONE 0xffffffffu
__host__ __device__ __forceinline__ uint32_t ref_pack(uint32_t index, uint32_t count) { return (count << 24) | index; }
__host__ __device__ __forceinline__ uint32_t ref_index(uint32_t r) { return r & 0x00ffffffu; }
__host__ __device__ __forceinline__ uint32_t ref_count(uint32_t r) { return r >> 24; }

// Everything a render kernel reads.  Passed by value in kernel parameter space.
struct DevScene {
    // reference layout, exactly as uploaded (PT_MODE_PARITY_REF traverses these)
    const float* tris_raw;      // nTriangles x 36 floats  (Triangle_encoded)
    const float* nodes_raw;     // nNodes x 12 floats      (BVHNode_encoded)
    // traversal layout built on the device at upload
    const float4* wnodes;       // nInner x 4 float4 (64 B): child boxes + packed child refs, REFERENCE topology
    const float4* wnodes_fast;  // same record format, binned-SAH topology over the triangles (pt_build.cpp), padded boxes
    const float4* trigeo_fast;  // nTriangles x 4 float4 (64 B), in the order of wnodes_fast's leaves: trigeo's three + (uploaded triangle
                                // index, reference leaf node holding it, -, -) as int bits: one 64-byte record per triangle test (wf_trace)
    const float4* trigeo_fast48; // the same triangles as 48-byte records (3 float4) + trimeta below: what the persistent / megakernel read
    const int2* trimeta;        // per re-ordered triangle: (uploaded triangle index, reference leaf node holding it); candidates only
    const float4* trigeo;       // nTriangles x 3 float4 (48 B): (p1,N.x) (p2,N.y) (p3,N.z), N = unit face normal
    const float4* trishade;     // nTriangles x 3 float4 (48 B): (n1, matId) (n2, -) (n3, -)
    const float4* materials;    // nMaterials x 8 float4 (128 B): the 16 parameters, then 16 values derived from them (pt_shade.cuh MaterialD)
    const float4* hdr4;         // hdrH x hdrW float4 (rgb, 0)
    const float4* cache4;       // hdrH x hdrW float4 (x/w, y/h, pdf, 0) or null
    const float4* light_tab;    // hdrH x hdrW x 2 float4: what a light sample through each cache texel evaluates to (pt_shade.cuh) or null
    // PT_ENV_CDF_SEARCH (cache4 and light_tab are null then): calculateHdrCache's intermediates, searched per sample
    const float* env_cdf_x;     // hdrW: marginal CDF over the columns
    const float* env_cdf_y;     // hdrW x hdrH: conditional CDF of every column, contiguous per column
    const float* env_pdf;       // hdrH x hdrW: pdf per texel (the table's B channel)
    int nTriangles, nNodes, nInner, nMaterials, nFastInner;
    int hdrW, hdrH;
    uint32_t root_ref, root_fast;
    int stack_depth;            // depth of the tree = worst-case traversal stack
    unsigned long long* ray_histogram;   // [33] log2 histogram of steps per ray (PT_MODE_FAST_COUNT)
    int* error_flag;            // set to PT_ERR_* by a kernel that had to give up (checked at the next sync)
};
#define PT_ERR_STACK 1          // traversal stack overflow: more deferred subtrees than PT_STACK (pt_trace.cuh: 128; trace_ref keeps the reference's stack[256], P:340)

// PT_DEBUG_BOUNDS build (tools/build_variant.py dbg -DPT_DEBUG_BOUNDS=1): every index the kernels form from scene or queue data
// is range-checked before use; a violation raises PT_ERR_BOUNDS (reported as PT_ERANGE at the next synchronising call) and
// the access is redirected to element 0.  compute-sanitizer is closed on the GPU pool this was developed on, so the GPU test
// suite is run once against this build instead (profiles/r02_bounds_check.md).
#define PT_ERR_BOUNDS 2
#if defined(PT_DEBUG_BOUNDS) && PT_DEBUG_BOUNDS
#define PT_IDX(sc, i, n) pt::checked_index((sc).error_flag, (size_t)(i), (size_t)(n))
__device__ __forceinline__ size_t checked_index(int* flag, size_t i, size_t n) {
    if (i < n) return i;
    *flag = PT_ERR_BOUNDS;
    return 0;
}
#else
#define PT_IDX(sc, i, n) (i)
#endif

// Path state of the wavefront pipeline (pt_wavefront.cuh), struct-of-arrays in HBM.
//   per PATH (compact ids handed out by wf_head to the paths that survive their primary ray; `pcap` of them)
//   per IMAGE SLOT = (frame within the wave) * npix + py * width + px (`icap` of them): the finished colour of every sample
#define PT_WF_MAX_BOUNCE 8
struct WaveBuffers {
    float4* ray_o;              // [pcap] xyz = origin of the rays in flight (the hit point they leave), w = sampling pdf of the bounce ray
    float4* ray_d;              //        xyz = bounce-ray direction, w = N.L of the sample that produced it
    float4* sh_d;               //        xyz = shadow-ray direction
    float4* hist;               //        xyz = path throughput before the bounce ray ("history", P:765), w = wang-hash state
    float4* fr;                 //        xyz = f_r of the sample that produced the bounce ray, w = PT_WF_FLAG_* bits
    float4* lo;                 //        xyz = Lo so far, w = material id of the primary hit (Le0)
    float4* cs;                 //        xyz = the pending light term of this bounce (zeroed when its shadow ray is occluded)
    int2* hit;                  //        bounce ray's hit: x = triangle | inside << 31, y = bits of t
    uint32_t* pix;              //        image slot of the path
    float4* img;                // [icap] colour of the sample (rgb, 1) once its path has ended
    uint32_t* q_ray;            // [2 pcap] rays of the current bounce: path | PT_WF_SHADOW_BIT
    int2* res;                  // [2 pcap] what wf_trace found for q_ray[i]: x = triangle | inside << 31 (-1: miss), y = bits of t
    uint32_t* q_hit;            // [pcap] paths whose bounce ray hit
    uint32_t* q_miss[2];        // [pcap] paths whose bounce ray escaped or never flew; by bounce parity (shade(b) reads b-1 while it appends to b)
    uint32_t* counters;         // queue tails and work-claim heads, one per 256 bytes
    size_t pcap, icap;
    uint32_t npix;              // width * height
    float fill;                 // share of pcap the host sizes the next wave for (0.75; PT_WF_FILL_PCT)
};

struct RenderParams {
    DevScene sc;
    float eye[3];
    float cam[16];              // column-major cameraRotate
    int width, height;
    int x0, y0, x1, y1;         // region
    int band_world, band_rank;  // multi-GPU tile split, see TileMap
    int band_diag;              // 0: whole 8-row bands, b % world == rank; 1: 8x4 tiles, (tx + strip) % world == rank
    unsigned first_frame, n_frames, frame_stride;
    int accumulate_sum;         // 0: running mean (P:869), 1: += colour
    int max_bounce;
    float4* accum;              // W*H float4, row 0 = bottom
    int* primary_ids;           // W*H
    unsigned long long* ray_counter;
    unsigned long long* work_counters;   // [4]: inner nodes fetched, triangles tested, leaf-box candidates, - (PT_MODE_FAST_COUNT)
    WaveBuffers wf;             // PT_PIPELINE_WAVEFRONT only
    uint32_t sched_stride;      // PT_PIPELINE_PERSISTENT: tile rows are visited in this many interleaved passes (0: the kernel's rule)
    const uint32_t* tile_tab;   // PT_PIPELINE_PERSISTENT: [owned tiles] the k-th tile of a frame in visiting order, tx | strip << 16
                                // (written per launch by tile_order_kernel), or null: every warp derives it per item
    const float* sobol_tab;     // variant 2: [n_frames][8] Sobol coordinates of this launch's frames (written by sobol_table_kernel) or null
    uint32_t* sched;            // PT_PIPELINE_PERSISTENT: [0] work head (the next frame a pixel accepts travels in the pixel's own .w)
    unsigned* cost;             // optional W*H cost map (SM cycles spent on the pixel by the last launch) or null
    int eye_far;                // the eye is so far outside the scene that float rounding of ray/triangle points (|o| 2^-23) can exceed the
                                // production tree's box padding (scene radius 2^-17): primary rays then take the reference topology
};

// Multi-GPU tile split: which 8x4 pixel warp tiles (tx = column, strip = 4-row strip, both counted from the region's
// corner) a launch owns, and how the owned ones are numbered 0 .. rows * per_row - 1.
//   bands    (pt_set_band_interleave): every tile of the 8-row band b = strip / 2 iff b % world == rank
//   diagonal (pt_set_tile_interleave): tile (tx, strip) iff (tx + strip) % world == rank — every rank gets an even sample
//            of the whole image (measured at 8 GPUs: slowest/mean rank 1.06 with bands)
// A numbered tile can lie right of the region when the column count does not divide; its pixels fail the x1 test.
struct TileMap {
    int per_row, rows, world, rank, diag;
    __host__ __device__ uint32_t count() const { return (uint32_t)per_row * (uint32_t)rows; }
    __host__ __device__ void locate(uint32_t t, int& tx, int& strip) const {
        const int row = (int)(t / (uint32_t)per_row), j = (int)(t - (uint32_t)row * (uint32_t)per_row);
        if (diag) { strip = row; tx = ((rank - strip) % world + world) % world + j * world; }
        else { strip = ((row >> 1) * world + rank) * 2 + (row & 1); tx = j; }
    }
};
__host__ __device__ inline TileMap make_tile_map(int x0, int y0, int x1, int y1, int world, int rank, int diag) {
    const int ntx = (x1 - x0 + 7) >> 3, strips = (y1 - y0 + 3) >> 2, bands = (y1 - y0 + 7) >> 3;
    TileMap m;
    m.world = world; m.rank = rank; m.diag = diag && world > 1;
    if (m.diag) { m.per_row = (ntx + world - 1) / world; m.rows = strips; }
    else { m.per_row = ntx; m.rows = 2 * (bands > rank ? (bands - rank + world - 1) / world : 0); }
    return m;
}

struct Material {               // P:41-56 without the two fields no shader reads (IOR, transmission)
    V3 emissive, baseColor;
    float subsurface, metallic, specular, specularTint, roughness, anisotropic;
    float sheen, sheenTint, clearcoat, clearcoatGloss;
};

struct Hit {                    // P:63-71
    bool isHit, isInside;
    float distance;
    V3 hitPoint, normal, viewDir;
    int tri;
};

// NEAREST + CLAMP_TO_EDGE texel index (M:190-193): floor(u*n) clamped to [0, n-1]; NaN -> 0
__device__ __forceinline__ int nearest_texel(float u, int n) {
    float f = floorf(s_mul(u, (float)n));
    f = fminf(fmaxf(f, 0.0f), (float)(n - 1));
    return (int)f;
}

}  // namespace pt
