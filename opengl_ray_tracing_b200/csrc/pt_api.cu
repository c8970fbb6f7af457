// pt_api.cu — the C ABI of libpt.so (include/pt.h) and the upload-time repack kernels.
//
// pt_upload_scene keeps a verbatim device copy of the reference-layout buffers (what the reference hands to
// glBufferData / glTexImage2D, IS/main.cpp:825-853) and derives the traversal layout from it ON THE DEVICE:
//   wide nodes   64 B / inner node : both child boxes + packed child refs, inner nodes compacted in the
//                                    reference's (DFS pre-order) numbering so a left child is the next record
//   trigeo       48 B / triangle   : p1,p2,p3 + the unit face normal, computed with the exact operations of
//                                    hitTriangle (P:263) so the precomputation changes no bit
//   trishade     48 B / triangle   : vertex normals + material id (fetched once per ray, for the winning hit)
//   materials   128 B / material   : de-duplicated (the reference stores all 72 B per triangle); second half = what the
//                                    Disney functions derive from the material alone (pt_shade.cuh MaterialD)
//   hdr4/cache4  16 B / texel      : float4 instead of 12-byte texels that straddle sectors
//   light_tab    32 B / texel      : variant 2 — what a light sample through each hdrCache texel evaluates to (pt_shade.cuh)
#include <cub/device/device_scan.cuh>
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <chrono>
#include <unordered_map>
#include <vector>

#include "pt_internal.h"

namespace pt {
cudaError_t launch_parity(const RenderParams& p, int variant, int which, cudaStream_t stream);
struct FastTree {
    std::vector<float> wnodes; std::vector<int> order, leaf_node;
    uint32_t root_ref = 0; int depth = 0; int n_leaves = 0; float pad = 0.0f;
};
void build_fast_tree(const float* tris_raw, int nTriangles, const float* nodes_raw, int nNodes, FastTree& out, int threads);
cudaError_t hdr_cache_device(const float* d_rgb, int W, int H, float4* d_cache4, float* d_cache3, cudaStream_t st);   // pt_hdrcache.cu
cudaError_t hdr_cdf_device(const float* d_rgb, int W, int H, float** pdf_out, float** cdf_x_out, float** cdf_yT_out, cudaStream_t st);
cudaError_t launch_persistent_fast_cdf(const RenderParams& p, int grid, cudaStream_t stream);
cudaError_t launch_persistent_parity_cdf(const RenderParams& p, int grid, cudaStream_t stream);
cudaError_t launch_fast(const RenderParams& p, int variant, bool count, cudaStream_t stream);
cudaError_t build_light_table_parity(const DevScene& sc, float4* tab, cudaStream_t stream);   // pt_kernels_parity.cu
cudaError_t derive_materials_parity(float4* materials, int n, cudaStream_t stream);
struct WaveStats { unsigned long long waves = 0, retries = 0, frames = 0; };
cudaError_t launch_wavefront_fast(const RenderParams& p, int variant, bool count, int num_sms, cudaStream_t stream, WaveStats* ws);
cudaError_t launch_persistent_fast(const RenderParams& p, int variant, bool count, int grid, cudaStream_t stream);
cudaError_t launch_persistent_parity(const RenderParams& p, int variant, int which, int grid, cudaStream_t stream);
cudaError_t launch_wavefront_parity(const RenderParams& p, int variant, int which, int num_sms, cudaStream_t stream, WaveStats* ws);
}

static thread_local std::string g_err;
int pt_fail(int code, const std::string& msg) { g_err = msg; return code; }

// ------------------------------------------------------------------------------------------------
// upload kernels
// ------------------------------------------------------------------------------------------------
namespace {

using namespace pt;

__global__ void k_flag_inner(const float* __restrict__ nodes, int nNodes, int* __restrict__ flag) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nNodes) return;
    // node 0 is the reference's dummy (IS/main.cpp:780-786); inner <=> n == 0
    flag[i] = (i >= 1 && (int)nodes[12 * (size_t)i + 3] == 0) ? 1 : 0;
}

__device__ __forceinline__ uint32_t child_ref(const float* __restrict__ nodes, const int* __restrict__ widx, int c) {
    const float* nd = nodes + 12 * (size_t)c;
    int n = (int)nd[3], index = (int)nd[4];
    return n > 0 ? ref_pack((uint32_t)index, (uint32_t)n) : ref_pack((uint32_t)widx[c], 0u);
}

__global__ void k_build_wide(const float* __restrict__ nodes, int nNodes, const int* __restrict__ flag,
                             const int* __restrict__ widx, float4* __restrict__ wnodes) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nNodes || !flag[i]) return;
    const float* nd = nodes + 12 * (size_t)i;
    int l = (int)nd[0], r = (int)nd[1];
    const float* L = nodes + 12 * (size_t)l;
    const float* R = nodes + 12 * (size_t)r;
    float4* w = wnodes + 4 * (size_t)widx[i];
    w[0] = make_float4(L[6], L[7], L[8], L[9]);
    w[1] = make_float4(L[10], L[11], R[6], R[7]);
    w[2] = make_float4(R[8], R[9], R[10], R[11]);
    w[3] = make_float4(__uint_as_float(child_ref(nodes, widx, l)), __uint_as_float(child_ref(nodes, widx, r)), 0.f, 0.f);
}

__global__ void k_build_tris(const float* __restrict__ tris, int nTriangles, const int* __restrict__ mat_id,
                             float4* __restrict__ geo, float4* __restrict__ shade) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nTriangles) return;
    const float* t = tris + 36 * (size_t)i;
    V3 p1 = mk3(t[0], t[1], t[2]), p2 = mk3(t[3], t[4], t[5]), p3 = mk3(t[6], t[7], t[8]);
    V3 N = s_normalize(s_cross(s_sub3(p2, p1), s_sub3(p3, p1)));          // hitTriangle, P:263
    geo[3 * (size_t)i] = make_float4(p1.x, p1.y, p1.z, N.x);
    geo[3 * (size_t)i + 1] = make_float4(p2.x, p2.y, p2.z, N.y);
    geo[3 * (size_t)i + 2] = make_float4(p3.x, p3.y, p3.z, N.z);
    shade[3 * (size_t)i] = make_float4(t[9], t[10], t[11], __int_as_float(mat_id[i]));
    shade[3 * (size_t)i + 1] = make_float4(t[12], t[13], t[14], 0.f);
    shade[3 * (size_t)i + 2] = make_float4(t[15], t[16], t[17], 0.f);
}

// trigeo in the traversal tree's order + (uploaded index, reference leaf node) per re-ordered triangle: 64-byte records
__global__ void k_gather_tris(const float4* __restrict__ geo, const int* __restrict__ order, const int* __restrict__ leaf_node,
                              int nTriangles, float4* __restrict__ geo_fast) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nTriangles) return;
    const int o = order[i];
    geo_fast[4 * (size_t)i] = geo[3 * (size_t)o];
    geo_fast[4 * (size_t)i + 1] = geo[3 * (size_t)o + 1];
    geo_fast[4 * (size_t)i + 2] = geo[3 * (size_t)o + 2];
    geo_fast[4 * (size_t)i + 3] = make_float4(__int_as_float(o), __int_as_float(leaf_node[i]), 0.f, 0.f);
}
__global__ void k_gather_tris48(const float4* __restrict__ geo, const int* __restrict__ order, const int* __restrict__ leaf_node,
                                int nTriangles, float4* __restrict__ geo_fast, int2* __restrict__ meta) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nTriangles) return;
    const int o = order[i];
    geo_fast[3 * (size_t)i] = geo[3 * (size_t)o];
    geo_fast[3 * (size_t)i + 1] = geo[3 * (size_t)o + 1];
    geo_fast[3 * (size_t)i + 2] = geo[3 * (size_t)o + 2];
    meta[i] = make_int2(o, leaf_node[i]);
}

__global__ void k_rgb_to_float4(const float* __restrict__ src, size_t n, float4* __restrict__ dst) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    dst[i] = make_float4(src[3 * i], src[3 * i + 1], src[3 * i + 2], 0.f);
}

// pass3.fsh:14-24 — toneMapping(c, 1.5) = c * 1.0 / (1.0 + (0.3 r + 0.6 g + 0.1 b) / 1.5), then the window's 8-bit
// UNORM conversion (round(clamp(v, 0, 1) * 255)).  Strict arithmetic: a float32 host restatement gives the same bytes.
// Output rows run top to bottom (image order); the framebuffer's row 0 is the bottom row.
__global__ void k_display(const float4* __restrict__ accum, int width, int height, uchar4* __restrict__ out) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= width || y >= height) return;
    const float4 c = accum[(size_t)(height - 1 - y) * width + x];
    const float lum = s_add(s_add(s_mul(0.3f, c.x), s_mul(0.6f, c.y)), s_mul(0.1f, c.z));
    const float k = s_add(1.0f, s_div(lum, 1.5f));
    float v[3] = {s_div(s_mul(c.x, 1.0f), k), s_div(s_mul(c.y, 1.0f), k), s_div(s_mul(c.z, 1.0f), k)};
    unsigned char q[3];
    for (int i = 0; i < 3; i++) {
        float t = fminf(fmaxf(v[i], 0.0f), 1.0f);            // NaN -> 0
        q[i] = (unsigned char)(int)s_add(s_mul(t, 255.0f), 0.5f);
    }
    out[(size_t)y * width + x] = make_uchar4(q[0], q[1], q[2], 255);
}

__global__ void k_scale(float4* __restrict__ a, size_t n, float s) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 v = a[i];
    a[i] = make_float4(v.x * s, v.y * s, v.z * s, v.w);
}

struct MatKey {
    float f[18];
    bool operator==(const MatKey& o) const { return std::memcmp(f, o.f, sizeof f) == 0; }
};
struct MatHash {
    size_t operator()(const MatKey& k) const {
        uint64_t h = 1469598103934665603ull;
        const unsigned char* p = reinterpret_cast<const unsigned char*>(k.f);
        for (size_t i = 0; i < sizeof k.f; i++) { h ^= p[i]; h *= 1099511628211ull; }
        return (size_t)h;
    }
};

cudaError_t create_events(pt_ctx* c) {
    for (int i = 0; i < pt_ctx::NEV; i++) {
        cudaError_t e = cudaEventCreate(&c->ev0[i]);
        if (e == cudaSuccess) e = cudaEventCreate(&c->ev1[i]);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

void free_scene(pt_ctx* c) {
    cudaFree(c->trigeo_fast); cudaFree(c->trigeo_fast48); cudaFree(c->trimeta);
    c->trigeo_fast = c->trigeo_fast48 = nullptr; c->trimeta = nullptr;
    cudaFree(c->tris_raw); cudaFree(c->nodes_raw); cudaFree(c->wnodes); cudaFree(c->wnodes_fast); cudaFree(c->trigeo); cudaFree(c->trishade);
    cudaFree(c->materials); cudaFree(c->hdr4); cudaFree(c->cache4); cudaFree(c->light_tab);
    cudaFree(c->env_pdf); cudaFree(c->env_cdf_x); cudaFree(c->env_cdf_y);
    c->env_pdf = c->env_cdf_x = c->env_cdf_y = nullptr;
    c->tris_raw = c->nodes_raw = nullptr;
    c->wnodes = c->wnodes_fast = c->trigeo = c->trishade = c->materials = c->hdr4 = c->cache4 = c->light_tab = nullptr;
    c->has_scene = false;
}

// Host-side validation of the reference-layout tree: every index in range, children after their parent (the builders emit
// pre-order), both children present on inner nodes, every node reachable, and the leaves' ranges a partition of the
// triangles.  `nested` reports whether every child box lies inside its parent's (the root's box aside: the reference never
// tests it, P:342-343) — the premise of the production traversal's "a hit on the leaf's box implies one on all its
// ancestors'" (pt_trace.cuh trace_fast).  Returns the tree depth or < 0.
int validate_tree(const float* nodes, int nNodes, int nTriangles, std::string& why, bool& nested) {
    nested = true;
    if (nNodes < 2) { why = "need the dummy node 0 and a root node 1"; return -1; }
    std::vector<int> depth((size_t)nNodes, 0);
    std::vector<unsigned char> owned((size_t)nTriangles, 0);
    depth[1] = 1;
    int max_depth = 1;
    long long covered = 0;
    for (int i = 1; i < nNodes; i++) {
        const float* nd = nodes + 12 * (size_t)i;
        int l = (int)nd[0], r = (int)nd[1], n = (int)nd[3], index = (int)nd[4];
        if ((float)l != nd[0] || (float)r != nd[1] || (float)n != nd[3] || (float)index != nd[4]) { why = "non-integer index in node " + std::to_string(i); return -1; }
        if (n < 0) { why = "negative triangle count in node " + std::to_string(i); return -1; }
        if (depth[i] == 0) { why = "node " + std::to_string(i) + " is unreachable"; return -1; }
        if (n > 0) {
            if (index < 0 || (long long)index + n > nTriangles) { why = "leaf range out of bounds in node " + std::to_string(i); return -1; }
            if (n > 255) { why = "leaf with more than 255 triangles (node " + std::to_string(i) + ")"; return -1; }
            for (int k = 0; k < n; k++) {
                if (owned[(size_t)index + k]) { why = "triangle " + std::to_string(index + k) + " belongs to two leaves (node " + std::to_string(i) + ")"; return -1; }
                owned[(size_t)index + k] = 1;
            }
            covered += n;
        } else {
            if (l <= i || r <= i || l >= nNodes || r >= nNodes || l == r) { why = "child index out of order/bounds in node " + std::to_string(i); return -1; }
            if (depth[l] || depth[r]) { why = "node " + std::to_string(depth[l] ? l : r) + " has two parents"; return -1; }
            depth[l] = depth[r] = depth[i] + 1;
            if (depth[i] + 1 > max_depth) max_depth = depth[i] + 1;
            if (i != 1) {
                for (int c : {l, r}) {
                    const float* cd = nodes + 12 * (size_t)c;
                    for (int a = 0; a < 3; a++)
                        if (!(cd[6 + a] >= nd[6 + a]) || !(cd[9 + a] <= nd[9 + a])) nested = false;
                }
            }
        }
    }
    if (covered != nTriangles) { why = std::to_string(nTriangles - covered) + " triangles belong to no leaf"; return -1; }
    return max_depth;
}

// One device allocation carved into the wavefront's struct-of-arrays (DESIGN.md "HBM layout").  PT_WF_PATHS_M million path
// slots (default 64; 156 B each incl. queue space) + image slots for the finished colours (16 B each): PT_WF_IMAGE_GB (default
// 16) worth of whole frames, at least one frame and at least as many frames as pcap / pixels.
int alloc_wavefront(pt_ctx* c) {
    if (c->wf_block) return PT_OK;
    const size_t np = (size_t)c->cfg.width * c->cfg.height;
    size_t pcap = (size_t)64 << 20, img_bytes = (size_t)16 << 30;
    if (const char* e = std::getenv("PT_WF_PATHS_M")) { const long v = std::atol(e); if (v > 0) pcap = (size_t)v << 20; }
    if (const char* e = std::getenv("PT_WF_IMAGE_GB")) { const long v = std::atol(e); if (v > 0) img_bytes = (size_t)v << 30; }
    if (pcap < np) pcap = np;                                     // one frame with every pixel alive always fits
    if (pcap >= ((size_t)1 << 31)) pcap = ((size_t)1 << 31) - 1;  // bit 31 of a queue entry marks a shadow ray
    if (pcap < np) return fail(PT_EINVAL, "wavefront pipeline: the framebuffer has more than 2^31 pixels");
    size_t frames = std::max<size_t>(std::max<size_t>(1, std::min<size_t>(4096, img_bytes / (np * 16))), pcap / np);
    if (frames * np >= ((size_t)1 << 32)) frames = (((size_t)1 << 32) - 1) / np;   // image slots are 32-bit
    const size_t icap = frames * np;
    const size_t ctr_bytes = (size_t)(6 * (PT_WF_MAX_BOUNCE + 1) + 4) * 64 * sizeof(uint32_t);
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t o_ro = take(pcap * 16), o_rd = take(pcap * 16), o_sh = take(pcap * 16), o_hist = take(pcap * 16), o_fr = take(pcap * 16);
    const size_t o_lo = take(pcap * 16), o_cs = take(pcap * 16), o_hit = take(pcap * 8), o_pix = take(pcap * 4);
    const size_t o_img = take(icap * 16);
    const size_t o_qr = take(pcap * 8), o_res = take(pcap * 16), o_qh = take(pcap * 4), o_qm0 = take(pcap * 4), o_qm1 = take(pcap * 4);
    const size_t o_ctr = take(ctr_bytes);
    cudaError_t e = cudaMalloc(&c->wf_block, off);
    if (e != cudaSuccess) { c->wf_block = nullptr; return fail(PT_ECUDA, std::string("wavefront path state: ") + cudaGetErrorString(e)); }
    char* b = (char*)c->wf_block;
    pt::WaveBuffers& w = c->wf;
    w.ray_o = (float4*)(b + o_ro); w.ray_d = (float4*)(b + o_rd); w.sh_d = (float4*)(b + o_sh); w.hist = (float4*)(b + o_hist);
    w.fr = (float4*)(b + o_fr); w.lo = (float4*)(b + o_lo); w.cs = (float4*)(b + o_cs); w.hit = (int2*)(b + o_hit);
    w.pix = (uint32_t*)(b + o_pix); w.img = (float4*)(b + o_img);
    w.q_ray = (uint32_t*)(b + o_qr); w.res = (int2*)(b + o_res); w.q_hit = (uint32_t*)(b + o_qh);
    w.q_miss[0] = (uint32_t*)(b + o_qm0); w.q_miss[1] = (uint32_t*)(b + o_qm1);
    w.counters = (uint32_t*)(b + o_ctr);
    w.pcap = pcap; w.icap = icap;
    w.npix = (uint32_t)np;
    w.fill = 0.75f;
    if (const char* e = std::getenv("PT_WF_FILL_PCT")) { const long v = std::atol(e); if (v > 0) w.fill = 0.01f * (float)v; }
    return PT_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------
extern "C" {

const char* pt_last_error(void) { return g_err.c_str(); }

const char* pt_build_info(void) {
    static char buf[160];
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    std::snprintf(buf, sizeof buf, "libpt sm_100a cuda-runtime %d devices=%d%s", CUDART_VERSION, e == cudaSuccess ? n : 0,
                  e == cudaSuccess ? "" : " (no CUDA device: this library has no CPU fallback)");
    return buf;
}

int pt_create(const PtConfig* cfg, pt_ctx** out) {
    if (!cfg || !out) return fail(PT_EINVAL, "pt_create: null argument");
    if (cfg->variant < 0 || cfg->variant > 2) return fail(PT_EINVAL, "pt_create: variant must be 0, 1 or 2");
    if (cfg->mode < 0 || cfg->mode > 4) return fail(PT_EINVAL, "pt_create: mode must be one of PT_MODE_*");
    if (cfg->width <= 0 || cfg->height <= 0 || cfg->width > 16384 || cfg->height > 16384)
        return fail(PT_EINVAL, "pt_create: width/height out of range");
    int mb = cfg->max_bounce > 0 ? cfg->max_bounce : (cfg->variant == PT_VARIANT_DISNEY_UNIFORM ? 5 : 2);
    if (cfg->variant == PT_VARIANT_DISNEY_IS_MIS && mb > 4)
        return fail(PT_EINVAL, "pt_create: variant 2 supports max_bounce <= 4 (8 Sobol dimensions, pass1.fsh:92)");
    if (cfg->pipeline < PT_PIPELINE_MEGAKERNEL || cfg->pipeline > PT_PIPELINE_PERSISTENT)
        return fail(PT_EINVAL, "pt_create: pipeline must be one of PT_PIPELINE_*");
    if (cfg->pipeline == PT_PIPELINE_WAVEFRONT && cfg->mode == PT_MODE_PARITY_REF)
        return fail(PT_EINVAL, "pt_create: PT_MODE_PARITY_REF (the reference's own layout and traversal) runs as a megakernel only");
    if (cfg->pipeline == PT_PIPELINE_WAVEFRONT && mb > PT_WF_MAX_BOUNCE)
        return fail(PT_EINVAL, "pt_create: the wavefront pipeline supports max_bounce <= 8");
    if (cfg->env_sampler != PT_ENV_TABLE && cfg->env_sampler != PT_ENV_CDF_SEARCH) return fail(PT_EINVAL, "pt_create: env_sampler must be one of PT_ENV_*");
    if (cfg->env_sampler == PT_ENV_CDF_SEARCH &&
        (cfg->variant != PT_VARIANT_DISNEY_IS_MIS || cfg->pipeline != PT_PIPELINE_PERSISTENT || (cfg->mode != PT_MODE_FAST && cfg->mode != PT_MODE_PARITY_FAST)))
        return fail(PT_EINVAL, "pt_create: PT_ENV_CDF_SEARCH needs variant 2, the persistent pipeline and PT_MODE_FAST or PT_MODE_PARITY_FAST");
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (cfg->device < 0 || cfg->device >= ndev) return fail(PT_EINVAL, "pt_create: no such CUDA device");
    CK(cudaSetDevice(cfg->device));
    pt_ctx* c = new (std::nothrow) pt_ctx();
    if (!c) return fail(PT_ENOMEM, "pt_create: out of host memory");
    c->cfg = *cfg;
    c->cfg.max_bounce = mb;
    c->x1 = cfg->width; c->y1 = cfg->height;
    std::memset(&c->sc, 0, sizeof c->sc);
    std::memset(&c->wf, 0, sizeof c->wf);
    cudaDeviceGetAttribute(&c->num_sms, cudaDevAttrMultiProcessorCount, cfg->device);
    size_t npx = (size_t)cfg->width * cfg->height;
    cudaError_t e;
    if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = create_events(c)) != cudaSuccess ||
        (e = cudaMalloc(&c->accum, npx * sizeof(float4))) != cudaSuccess ||
        (e = cudaMalloc(&c->primary_ids, npx * sizeof(int))) != cudaSuccess ||
        (e = cudaMalloc(&c->ray_counter, 48 * sizeof(unsigned long long))) != cudaSuccess ||
        (e = cudaMalloc(&c->error_flag, sizeof(int))) != cudaSuccess ||
        (e = cudaMemsetAsync(c->error_flag, 0, sizeof(int), c->stream)) != cudaSuccess ||
        (e = cudaMemsetAsync(c->accum, 0, npx * sizeof(float4), c->stream)) != cudaSuccess ||
        (e = cudaMemsetAsync(c->primary_ids, 0xff, npx * sizeof(int), c->stream)) != cudaSuccess ||
        (e = cudaMemsetAsync(c->ray_counter, 0, 48 * sizeof(unsigned long long), c->stream)) != cudaSuccess) {
        pt_destroy(c);
        return fail(PT_ECUDA, std::string("pt_create: ") + cudaGetErrorString(e));
    }
    *out = c;
    return PT_OK;
}

int pt_destroy(pt_ctx* c) {
    if (!c) return PT_OK;
    cudaSetDevice(c->cfg.device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    pt_comm_release(c);
    free_scene(c);
    cudaFree(c->wf_block); cudaFree(c->sched); cudaFree(c->tile_tab); cudaFree(c->sobol_tab); cudaFree(c->display); cudaFree(c->cost); cudaFree(c->accum); cudaFree(c->primary_ids); cudaFree(c->ray_counter); cudaFree(c->error_flag);
    for (int i = 0; i < pt_ctx::NEV; i++) {
        if (c->ev0[i]) cudaEventDestroy(c->ev0[i]);
        if (c->ev1[i]) cudaEventDestroy(c->ev1[i]);
    }
    if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
    if (c->ev_image) cudaEventDestroy(c->ev_image);
    if (c->ev_copied) cudaEventDestroy(c->ev_copied);
    cudaFree(c->snapshot);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return PT_OK;
}

int pt_upload_scene(pt_ctx* c, const float* tris, int nTriangles, const float* nodes, int nNodes, const float* hdr,
                    const float* hdrCache, int hdrW, int hdrH) {
    if (!c || !tris || !nodes || !hdr) return fail(PT_EINVAL, "pt_upload_scene: null argument");
    if (nTriangles <= 0 || nNodes < 2 || hdrW <= 0 || hdrH <= 0) return fail(PT_EINVAL, "pt_upload_scene: empty scene or HDR map");
    if (nTriangles >= (1 << 24) || nNodes >= (1 << 24))
        return fail(PT_ERANGE, "pt_upload_scene: index not exact in the reference's float encoding (>= 2^24)");
    std::string why;
    bool nested = true;
    int depth = validate_tree(nodes, nNodes, nTriangles, why, nested);
    if (depth < 0) return fail(PT_EFORMAT, "pt_upload_scene: " + why);
    if (!nested && c->cfg.mode != PT_MODE_PARITY_REF && c->cfg.mode != PT_MODE_PARITY_WIDE)
        return fail(PT_EFORMAT, "pt_upload_scene: a child box is not inside its parent's; the production traversal (PT_MODE_FAST, "
                                "PT_MODE_PARITY_FAST) needs nested boxes, as buildBVH / buildBVHwithSAH produce them — "
                                "PT_MODE_PARITY_REF and PT_MODE_PARITY_WIDE follow the array as it is");
    // depth > 256 is accepted: the reference's stack[256] (pass1.fsh:340) only overflows if a RAY defers more than
    // 256 subtrees at once; the kernels detect an overflow of their own stacks at run time (PT_MODE_PARITY_REF: 256
    // entries like the reference; the wide/production traversals: PT_STACK = 128 deferred subtrees on trees of depth
    // < 64) and report PT_ERANGE at the next synchronising call.

    // material de-duplication (host hash over the 18 material floats of each triangle)
    std::vector<int> mat_id((size_t)nTriangles);
    std::vector<float> mat_table;
    {
        std::unordered_map<MatKey, int, MatHash> seen;
        MatKey k, last;
        int last_id = -1;
        for (int i = 0; i < nTriangles; i++) {
            std::memcpy(k.f, tris + 36 * (size_t)i + 18, sizeof k.f);
            if (last_id >= 0 && k == last) { mat_id[i] = last_id; continue; }
            auto it = seen.find(k);
            int id;
            if (it == seen.end()) {
                id = (int)seen.size();
                seen.emplace(k, id);
                const float* f = k.f;                 // emissive, baseColor, 12 params (IS/main.cpp:806-811)
                const float rec[16] = {f[0], f[1], f[2], f[6], f[3], f[4], f[5], f[7],
                                       f[8], f[9], f[10], f[11], f[12], f[13], f[14], f[15]};
                mat_table.insert(mat_table.end(), rec, rec + 16);
                mat_table.insert(mat_table.end(), 16, 0.0f);     // derived half, filled on the device (material_derive_kernel)
            } else id = it->second;
            mat_id[i] = id; last = k; last_id = id;
        }
    }
    const int nMaterials = (int)(mat_table.size() / 32);

    CK(cudaSetDevice(c->cfg.device));
    CK(cudaStreamSynchronize(c->stream));
    free_scene(c);
    cudaStream_t st = c->stream;
    const size_t ntex = (size_t)hdrW * hdrH;
    int *d_flag = nullptr, *d_widx = nullptr, *d_matid = nullptr, *d_order = nullptr, *d_leaf = nullptr;
    float* d_tmp = nullptr;
    void* d_scan = nullptr;
    size_t scan_bytes = 0;
    int nInner = 0, nFastInner = 0, fast_depth = 0, cache_launches = 0;
    uint32_t root_fast = 0;
    int rc = PT_OK;
#define CKU(call)                                                                                  \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) { rc = fail(PT_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); goto done; } \
    } while (0)
    CKU(cudaMalloc(&c->tris_raw, (size_t)nTriangles * 36 * sizeof(float)));
    CKU(cudaMalloc(&c->nodes_raw, (size_t)nNodes * 12 * sizeof(float)));
    CKU(cudaMemcpyAsync(c->tris_raw, tris, (size_t)nTriangles * 36 * sizeof(float), cudaMemcpyHostToDevice, st));
    CKU(cudaMemcpyAsync(c->nodes_raw, nodes, (size_t)nNodes * 12 * sizeof(float), cudaMemcpyHostToDevice, st));
    // wide nodes: flag inner nodes, exclusive scan -> compact index, build
    CKU(cudaMalloc(&d_flag, (size_t)nNodes * sizeof(int)));
    CKU(cudaMalloc(&d_widx, (size_t)nNodes * sizeof(int)));
    k_flag_inner<<<(nNodes + 255) / 256, 256, 0, st>>>(c->nodes_raw, nNodes, d_flag);
    CKU(cudaGetLastError());
    CKU(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, d_flag, d_widx, nNodes, st));
    CKU(cudaMalloc(&d_scan, scan_bytes ? scan_bytes : 1));
    CKU(cub::DeviceScan::ExclusiveSum(d_scan, scan_bytes, d_flag, d_widx, nNodes, st));
    {
        int last_flag = 0, last_idx = 0;
        CKU(cudaMemcpyAsync(&last_flag, d_flag + (nNodes - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
        CKU(cudaMemcpyAsync(&last_idx, d_widx + (nNodes - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
        CKU(cudaStreamSynchronize(st));
        nInner = last_idx + last_flag;
    }
    CKU(cudaMalloc(&c->wnodes, (size_t)(nInner > 0 ? nInner : 1) * 4 * sizeof(float4)));
    if (nInner > 0) {
        k_build_wide<<<(nNodes + 255) / 256, 256, 0, st>>>(c->nodes_raw, nNodes, d_flag, d_widx, c->wnodes);
        CKU(cudaGetLastError());
    }
    // triangles
    CKU(cudaMalloc(&d_matid, (size_t)nTriangles * sizeof(int)));
    CKU(cudaMemcpyAsync(d_matid, mat_id.data(), (size_t)nTriangles * sizeof(int), cudaMemcpyHostToDevice, st));
    CKU(cudaMalloc(&c->trigeo, (size_t)nTriangles * 3 * sizeof(float4)));
    CKU(cudaMalloc(&c->trishade, (size_t)nTriangles * 3 * sizeof(float4)));
    k_build_tris<<<(nTriangles + 255) / 256, 256, 0, st>>>(c->tris_raw, nTriangles, d_matid, c->trigeo, c->trishade);
    CKU(cudaGetLastError());
    // traversal topology: binned SAH over the triangles (host build, pt_build.cpp), then gather trigeo into its order
    {
        pt::FastTree ft;
        int hw = (int)std::thread::hardware_concurrency();
        auto t0 = std::chrono::steady_clock::now();
        pt::build_fast_tree(tris, nTriangles, nodes, nNodes, ft, hw > 0 ? (hw > 32 ? 32 : hw) : 1);
        c->build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        root_fast = ft.root_ref; fast_depth = ft.depth; nFastInner = (int)(ft.wnodes.size() / 16);
        c->fast_pad = ft.pad;
        if (nFastInner >= (1 << 24)) { rc = fail(PT_ERANGE, "pt_upload_scene: traversal tree has >= 2^24 inner nodes"); goto done; }
        CKU(cudaMalloc(&c->wnodes_fast, (ft.wnodes.empty() ? 16 : ft.wnodes.size()) * sizeof(float)));
        if (!ft.wnodes.empty())
            CKU(cudaMemcpyAsync(c->wnodes_fast, ft.wnodes.data(), ft.wnodes.size() * sizeof(float), cudaMemcpyHostToDevice, st));
        CKU(cudaMalloc(&d_order, (size_t)nTriangles * sizeof(int)));
        CKU(cudaMalloc(&d_leaf, (size_t)nTriangles * sizeof(int)));
        CKU(cudaMemcpyAsync(d_order, ft.order.data(), (size_t)nTriangles * sizeof(int), cudaMemcpyHostToDevice, st));
        CKU(cudaMemcpyAsync(d_leaf, ft.leaf_node.data(), (size_t)nTriangles * sizeof(int), cudaMemcpyHostToDevice, st));
        CKU(cudaMalloc(&c->trigeo_fast, (size_t)nTriangles * 4 * sizeof(float4)));
        k_gather_tris<<<(nTriangles + 255) / 256, 256, 0, st>>>(c->trigeo, d_order, d_leaf, nTriangles, c->trigeo_fast);
        CKU(cudaGetLastError());
        CKU(cudaMalloc(&c->trigeo_fast48, (size_t)nTriangles * 3 * sizeof(float4)));
        CKU(cudaMalloc(&c->trimeta, (size_t)nTriangles * sizeof(int2)));
        k_gather_tris48<<<(nTriangles + 255) / 256, 256, 0, st>>>(c->trigeo, d_order, d_leaf, nTriangles, c->trigeo_fast48, c->trimeta);
        CKU(cudaGetLastError());
        CKU(cudaStreamSynchronize(st));               // ft goes out of scope
    }
    CKU(cudaMalloc(&c->materials, (size_t)nMaterials * 8 * sizeof(float4)));
    CKU(cudaMemcpyAsync(c->materials, mat_table.data(), (size_t)nMaterials * 32 * sizeof(float), cudaMemcpyHostToDevice, st));
    // second half of each record: what the Disney functions derive from the material alone.  Evaluated in the EXACT arithmetic
    // (the parity translation unit) for every mode: a^2 - 1 and friends feed the cancellations of the lobe samplers
    // (pt_shade.cuh SampleBRDF), where an ulp decides which environment texel a mirror-like surface reflects.
    CKU(pt::derive_materials_parity(c->materials, nMaterials, st));
    // environment
    CKU(cudaMalloc(&d_tmp, ntex * 3 * sizeof(float)));
    CKU(cudaMalloc(&c->hdr4, ntex * sizeof(float4)));
    CKU(cudaMemcpyAsync(d_tmp, hdr, ntex * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
    k_rgb_to_float4<<<(unsigned)((ntex + 255) / 256), 256, 0, st>>>(d_tmp, ntex, c->hdr4);
    CKU(cudaGetLastError());
    if (c->cfg.env_sampler == PT_ENV_CDF_SEARCH) {
        // no table at all: the marginal / conditional CDFs calculateHdrCache inverts, searched per sample (pt_shade.cuh SampleHdr)
        CKU(pt::hdr_cdf_device(d_tmp, hdrW, hdrH, &c->env_pdf, &c->env_cdf_x, &c->env_cdf_y, st));
        cache_launches = 6;
    } else if (hdrCache) {
        CKU(cudaMalloc(&c->cache4, ntex * sizeof(float4)));
        CKU(cudaStreamSynchronize(st));
        CKU(cudaMemcpyAsync(d_tmp, hdrCache, ntex * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
        k_rgb_to_float4<<<(unsigned)((ntex + 255) / 256), 256, 0, st>>>(d_tmp, ntex, c->cache4);
        CKU(cudaGetLastError());
    } else if (c->cfg.variant == PT_VARIANT_DISNEY_IS_MIS) {
        // no table supplied: calculateHdrCache (M:555-652) runs here on the device, same arithmetic, same table
        CKU(cudaMalloc(&c->cache4, ntex * sizeof(float4)));
        CKU(pt::hdr_cache_device(d_tmp, hdrW, hdrH, c->cache4, nullptr, st));
        cache_launches = 6;
    }
    CKU(cudaStreamSynchronize(st));
    {
        pt::DevScene& s = c->sc;
        s.tris_raw = c->tris_raw; s.nodes_raw = c->nodes_raw; s.wnodes = c->wnodes; s.wnodes_fast = c->wnodes_fast; s.trigeo = c->trigeo;
        s.trigeo_fast = c->trigeo_fast; s.trigeo_fast48 = c->trigeo_fast48; s.trimeta = c->trimeta;
        s.root_fast = root_fast; c->fast_depth = fast_depth; c->n_fast_inner = nFastInner;
        s.trishade = c->trishade; s.materials = c->materials; s.hdr4 = c->hdr4; s.cache4 = c->cache4;
        s.env_pdf = c->env_pdf; s.env_cdf_x = c->env_cdf_x; s.env_cdf_y = c->env_cdf_y;
        s.nTriangles = nTriangles; s.nNodes = nNodes; s.nInner = nInner; s.nMaterials = nMaterials; s.nFastInner = nFastInner;
        s.hdrW = hdrW; s.hdrH = hdrH; s.stack_depth = depth; s.error_flag = c->error_flag; s.ray_histogram = c->ray_counter + 8;
        // root = node 1: a leaf when the whole scene fits one leaf, else wide node 0
        int rn = (int)nodes[12 + 3], rindex = (int)nodes[12 + 4];
        s.root_ref = rn > 0 ? pt::ref_pack((uint32_t)rindex, (uint32_t)rn) : pt::ref_pack(0u, 0u);
        s.light_tab = nullptr;
        if (c->cfg.variant == PT_VARIANT_DISNEY_IS_MIS && c->cache4 && c->cfg.mode != PT_MODE_PARITY_REF) {
            // light table: what a light sample through each hdrCache texel evaluates to, in this context's arithmetic (pt_shade.cuh)
            CKU(cudaMalloc(&c->light_tab, ntex * 2 * sizeof(float4)));
            // ... in the exact arithmetic for every mode (upload-time work: the production kernel only reads the table)
            CKU(pt::build_light_table_parity(s, c->light_tab, st));
            CKU(cudaStreamSynchronize(st));
            s.light_tab = c->light_tab;
            c->launches += 1;
        }
        c->has_scene = true;
        c->launches += 6 + (hdrCache ? 1 : cache_launches);
        // Optional L2 persistence window over the production tree's nodes (north star item 2); PT_L2_WINDOW_MB=0 (default)
        // leaves L2 to its own replacement: measured no gain, DESIGN.md §7.
        if (const char* e = std::getenv("PT_L2_WINDOW_MB")) {
            const size_t want = (size_t)std::atoi(e) << 20;
            if (want > 0 && nFastInner > 0) {
                cudaDeviceProp prop;
                if (cudaGetDeviceProperties(&prop, c->cfg.device) == cudaSuccess && prop.accessPolicyMaxWindowSize > 0) {
                    size_t bytes = (size_t)nFastInner * 64;
                    if (bytes > want) bytes = want;
                    if (bytes > (size_t)prop.accessPolicyMaxWindowSize) bytes = (size_t)prop.accessPolicyMaxWindowSize;
                    cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, bytes);
                    cudaStreamAttrValue attr;
                    std::memset(&attr, 0, sizeof attr);
                    attr.accessPolicyWindow.base_ptr = (void*)c->wnodes_fast;
                    attr.accessPolicyWindow.num_bytes = bytes;
                    attr.accessPolicyWindow.hitRatio = 1.0f;
                    attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                    attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                    cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &attr);
                    cudaGetLastError();
                }
            }
        }
    }
done:
    cudaFree(d_order); cudaFree(d_leaf);
    cudaFree(d_flag); cudaFree(d_widx); cudaFree(d_matid); cudaFree(d_tmp); cudaFree(d_scan);
    if (rc != PT_OK) free_scene(c);
    return rc;
#undef CKU
}

int pt_set_camera(pt_ctx* c, const float eye[3], const float cam[16]) {
    if (!c || !eye || !cam) return fail(PT_EINVAL, "pt_set_camera: null argument");
    std::memcpy(c->eye, eye, 12);
    std::memcpy(c->cam, cam, 64);
    return PT_OK;
}

int pt_set_region(pt_ctx* c, int x0, int y0, int x1, int y1) {
    if (!c) return fail(PT_EINVAL, "pt_set_region: null context");
    if (x0 < 0 || y0 < 0 || x1 > c->cfg.width || y1 > c->cfg.height || x0 >= x1 || y0 >= y1)
        return fail(PT_EINVAL, "pt_set_region: rectangle outside the framebuffer or empty");
    c->x0 = x0; c->y0 = y0; c->x1 = x1; c->y1 = y1;
    return PT_OK;
}

int pt_set_band_interleave(pt_ctx* c, int world, int rank) {
    if (!c) return fail(PT_EINVAL, "pt_set_band_interleave: null context");
    if (world < 1 || rank < 0 || rank >= world) return fail(PT_EINVAL, "pt_set_band_interleave: need 0 <= rank < world");
    c->band_world = world; c->band_rank = rank; c->band_diag = 0;
    return PT_OK;
}

int pt_set_tile_interleave(pt_ctx* c, int world, int rank) {
    if (!c) return fail(PT_EINVAL, "pt_set_tile_interleave: null context");
    if (world < 1 || rank < 0 || rank >= world) return fail(PT_EINVAL, "pt_set_tile_interleave: need 0 <= rank < world");
    c->band_world = world; c->band_rank = rank; c->band_diag = 1;
    return PT_OK;
}

// Folds finished launch timings into render_ms.  wait = false: only pairs that have completed (never blocks);
// wait = true (synchronising calls: readback, synchronize, get_stats): all of them, and the kernels' error flag is read.
static int collect_timing(pt_ctx* c, bool wait = true) {
    while (c->ev_tail != c->ev_head) {
        const int i = c->ev_tail % pt_ctx::NEV;
        if (!wait) {
            const cudaError_t q = cudaEventQuery(c->ev1[i]);
            if (q == cudaErrorNotReady) break;
            if (q != cudaSuccess) return fail(PT_ECUDA, std::string("cudaEventQuery: ") + cudaGetErrorString(q));
        } else {
            CK(cudaEventSynchronize(c->ev1[i]));
        }
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, c->ev0[i], c->ev1[i]));
        c->render_ms += ms;
        c->ev_tail++;
    }
    if (!wait) return PT_OK;
    int flag = 0;
    CK(cudaMemcpy(&flag, c->error_flag, sizeof flag, cudaMemcpyDeviceToHost));
    if (flag == PT_ERR_BOUNDS) {
        CK(cudaMemset(c->error_flag, 0, sizeof flag));
        return fail(PT_ERANGE, "render kernel: an index left its array (PT_DEBUG_BOUNDS build) - image is invalid");
    }
    if (flag) {
        CK(cudaMemset(c->error_flag, 0, sizeof flag));
        return fail(PT_ERANGE, "render kernel: traversal stack overflow (a ray deferred more subtrees than the traversal stack holds: "
                               "256 in PT_MODE_PARITY_REF, where the reference's own stack[256], pass1.fsh:340, would have been overrun; "
                               "128 in the other modes) - image is invalid");
    }
    return PT_OK;
}

int pt_render_spp(pt_ctx* c, unsigned first_frame, unsigned n_frames) {
    return pt_render_frames(c, first_frame, n_frames, 1u, PT_ACC_MEAN);
}

int pt_render_frames(pt_ctx* c, unsigned first_frame, unsigned n_frames, unsigned stride, int accumulate) {
    if (!c) return fail(PT_EINVAL, "pt_render_spp: null context");
    if (stride == 0 || (accumulate != PT_ACC_MEAN && accumulate != PT_ACC_SUM)) return fail(PT_EINVAL, "pt_render_frames: bad stride/accumulate");
    if (accumulate == PT_ACC_MEAN && stride != 1) return fail(PT_EINVAL, "pt_render_frames: the running mean needs stride 1");
    if (!c->has_scene) return fail(PT_ENOSCENE, "pt_render_spp: no scene uploaded");
    if (n_frames == 0) return PT_OK;
    CK(cudaSetDevice(c->cfg.device));
    int rc = collect_timing(c, false);
    if (rc != PT_OK) return rc;
    if (c->ev_head - c->ev_tail >= pt_ctx::NEV) {             // the ring is full: wait for its oldest pair only
        const int i = c->ev_tail % pt_ctx::NEV;
        float ms = 0.f;
        CK(cudaEventSynchronize(c->ev1[i]));
        CK(cudaEventElapsedTime(&ms, c->ev0[i], c->ev1[i]));
        c->render_ms += ms;
        c->ev_tail++;
    }
    pt::RenderParams p;
    p.sc = c->sc;
    std::memcpy(p.eye, c->eye, 12);
    std::memcpy(p.cam, c->cam, 64);
    p.width = c->cfg.width; p.height = c->cfg.height;
    p.x0 = c->x0; p.y0 = c->y0; p.x1 = c->x1; p.y1 = c->y1;
    p.band_world = c->band_world; p.band_rank = c->band_rank; p.band_diag = c->band_diag;
    p.first_frame = first_frame; p.n_frames = n_frames; p.frame_stride = stride; p.accumulate_sum = accumulate;
    p.max_bounce = c->cfg.max_bounce;
    {
        // padding = radius 2^-17 covers the rounding of ray points while |eye| <~ 16 radii (error ~ 4 |o| 2^-24); beyond 8 radii
        // primary rays are traced on the reference topology (pt_shade.cuh trace<>)
        const float radius = std::ldexp(c->fast_pad, 17);
        const float far = std::max(std::fabs(c->eye[0]), std::max(std::fabs(c->eye[1]), std::fabs(c->eye[2])));
        p.eye_far = (!(far <= 8.0f * radius)) ? 1 : 0;
    }
    p.accum = c->accum; p.primary_ids = c->primary_ids; p.ray_counter = c->ray_counter; p.work_counters = c->ray_counter + 1; p.cost = c->cost;
    const bool fastmode = c->cfg.mode == PT_MODE_FAST || c->cfg.mode == PT_MODE_FAST_COUNT;
    const bool wavefront = c->cfg.pipeline == PT_PIPELINE_WAVEFRONT;
    if (wavefront) {
        rc = alloc_wavefront(c);
        if (rc != PT_OK) return rc;
        p.wf = c->wf;
    } else {
        std::memset(&p.wf, 0, sizeof p.wf);
    }
    const bool persistent = c->cfg.pipeline == PT_PIPELINE_PERSISTENT;
    if (persistent && !c->sched) CK(cudaMalloc(&c->sched, 64 * sizeof(uint32_t)));   // the persistent kernel's work head
    p.sched = c->sched;
    p.tile_tab = nullptr;
    if (persistent) {
        const pt::TileMap tm0 = pt::make_tile_map(c->x0, c->y0, c->x1, c->y1, c->band_world, c->band_rank, c->band_diag);
        const size_t nt = tm0.count();
        if (c->tile_cap < nt) {
            CK(cudaStreamSynchronize(c->stream));            // an earlier launch may still read the old table
            cudaFree(c->tile_tab); c->tile_tab = nullptr; c->tile_cap = 0;
            CK(cudaMalloc(&c->tile_tab, std::max<size_t>(nt, 1024) * sizeof(uint32_t)));
            c->tile_cap = std::max<size_t>(nt, 1024);
        }
        p.tile_tab = c->tile_tab;
    }
    {
        // visiting order of the tile rows (render_persistent): 0 = the kernel's own rule, PT_SCHED_STRIDE=1 = bottom to top
        const char* e = std::getenv("PT_SCHED_STRIDE");
        p.sched_stride = e && *e ? (uint32_t)std::max(0, std::atoi(e)) : 0u;
    }
    p.sobol_tab = nullptr;
    if (c->cfg.variant == PT_VARIANT_DISNEY_IS_MIS) {
        if (c->sobol_cap < n_frames) {
            CK(cudaStreamSynchronize(c->stream));            // an earlier launch may still read the old table
            cudaFree(c->sobol_tab); c->sobol_tab = nullptr; c->sobol_cap = 0;
            const size_t cap = std::max<size_t>(n_frames, 256);
            CK(cudaMalloc(&c->sobol_tab, cap * 8 * sizeof(float)));
            c->sobol_cap = cap;
        }
        p.sobol_tab = c->sobol_tab;
    }
    const int evi = c->ev_head % pt_ctx::NEV;
    CK(cudaEventRecord(c->ev0[evi], c->stream));
    cudaError_t e;
    pt::WaveStats wstats;
    if (wavefront) {
        const int sms = c->num_sms > 0 ? c->num_sms : 148;
        // (synchronises once per wave: the host sizes each wave from the survivor count of the one before, pt_wavefront.cuh)
        e = fastmode ? pt::launch_wavefront_fast(p, c->cfg.variant, c->cfg.mode == PT_MODE_FAST_COUNT, sms, c->stream, &wstats)
                     : pt::launch_wavefront_parity(p, c->cfg.variant, c->cfg.mode, sms, c->stream, &wstats);
    } else if (persistent) {
        const int grid = c->num_sms > 0 ? c->num_sms : 148;   // the launcher multiplies by the kernel's CTAs per SM
        if (c->cfg.env_sampler == PT_ENV_CDF_SEARCH)
            e = fastmode ? pt::launch_persistent_fast_cdf(p, grid, c->stream) : pt::launch_persistent_parity_cdf(p, grid, c->stream);
        else
            e = fastmode ? pt::launch_persistent_fast(p, c->cfg.variant, c->cfg.mode == PT_MODE_FAST_COUNT, grid, c->stream)
                         : pt::launch_persistent_parity(p, c->cfg.variant, c->cfg.mode, grid, c->stream);
    } else {
        e = fastmode ? pt::launch_fast(p, c->cfg.variant, c->cfg.mode == PT_MODE_FAST_COUNT, c->stream)
                     : pt::launch_parity(p, c->cfg.variant, c->cfg.mode, c->stream);
    }
    if (e != cudaSuccess) return fail(PT_ECUDA, std::string("render kernel launch: ") + cudaGetErrorString(e));
    CK(cudaEventRecord(c->ev1[evi], c->stream));
    c->ev_head++;
    if (wavefront) {
        // per wave: head + max_bounce x (trace + classify + shade) + accumulate; a repeated wave ran its head twice
        c->launches += wstats.waves * (unsigned long long)(2 + 3 * c->cfg.max_bounce) + wstats.retries + (p.sobol_tab ? 1 : 0);
        c->wf_waves += wstats.waves; c->wf_retries += wstats.retries;
    } else {
        c->launches += (p.sobol_tab ? 2 : 1) + (p.tile_tab ? 1 : 0);   // the render kernel (+ the Sobol tabulation, + the tile-order table)
    }
    {
        const pt::TileMap tm = pt::make_tile_map(c->x0, c->y0, c->x1, c->y1, c->band_world, c->band_rank, c->band_diag);
        unsigned long long px = 0;
        for (uint32_t t = 0; t < tm.count(); t++) {
            int tx, strip;
            tm.locate(t, tx, strip);
            const int xa = c->x0 + tx * 8, ya = c->y0 + strip * 4;
            const int w = std::min(c->x1, xa + 8) - xa, h = std::min(c->y1, ya + 4) - ya;
            if (w > 0 && h > 0) px += (unsigned long long)w * h;
        }
        c->samples += px * n_frames;
    }
    return PT_OK;
}

int pt_clear(pt_ctx* c) {
    if (!c) return fail(PT_EINVAL, "pt_clear: null context");
    CK(cudaSetDevice(c->cfg.device));
    size_t npx = (size_t)c->cfg.width * c->cfg.height;
    CK(cudaMemsetAsync(c->accum, 0, npx * sizeof(float4), c->stream));
    CK(cudaMemsetAsync(c->primary_ids, 0xff, npx * sizeof(int), c->stream));
    return PT_OK;
}

int pt_readback(pt_ctx* c, float* rgba, size_t bytes) {
    if (!c || !rgba) return fail(PT_EINVAL, "pt_readback: null argument");
    size_t need = (size_t)c->cfg.width * c->cfg.height * sizeof(float4);
    if (bytes < need) return fail(PT_EINVAL, "pt_readback: buffer too small");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemcpyAsync(rgba, c->accum, need, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return collect_timing(c);
}

// calculateHdrCache (M:555-652) on a device, host buffers in and out: the stand-alone form of what pt_upload_scene does when
// it is handed no table.
int pt_calculate_hdr_cache(int device, const float* hdr, int width, int height, float* cache) {
    if (!hdr || !cache || width <= 0 || height <= 0) return fail(PT_EINVAL, "pt_calculate_hdr_cache: null buffer or empty image");
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(PT_EINVAL, "pt_calculate_hdr_cache: no such CUDA device");
    CK(cudaSetDevice(device));
    const size_t bytes = (size_t)width * height * 3 * sizeof(float);
    float *d_rgb = nullptr, *d_out = nullptr;
    cudaError_t e = cudaMalloc(&d_rgb, bytes);
    if (e == cudaSuccess) e = cudaMalloc(&d_out, bytes);
    if (e == cudaSuccess) e = cudaMemcpy(d_rgb, hdr, bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = pt::hdr_cache_device(d_rgb, width, height, nullptr, d_out, 0);
    if (e == cudaSuccess) e = cudaMemcpy(cache, d_out, bytes, cudaMemcpyDeviceToHost);
    cudaFree(d_rgb); cudaFree(d_out);
    if (e != cudaSuccess) return fail(PT_ECUDA, std::string("pt_calculate_hdr_cache: ") + cudaGetErrorString(e));
    return PT_OK;
}

int pt_readback_display(pt_ctx* c, unsigned char* rgba8, size_t bytes) {
    if (!c || !rgba8) return fail(PT_EINVAL, "pt_readback_display: null argument");
    const size_t npx = (size_t)c->cfg.width * c->cfg.height;
    if (bytes < npx * 4) return fail(PT_EINVAL, "pt_readback_display: buffer too small");
    CK(cudaSetDevice(c->cfg.device));
    if (!c->display) CK(cudaMalloc(&c->display, npx * sizeof(uchar4)));
    if (c->copy_pending) CK(cudaStreamWaitEvent(c->stream, c->ev_copied, 0));     // an asynchronous copy may still read the staging image
    dim3 block(32, 8), grid((c->cfg.width + 31) / 32, (c->cfg.height + 7) / 8);
    k_display<<<grid, block, 0, c->stream>>>(c->accum, c->cfg.width, c->cfg.height, c->display);
    CK(cudaGetLastError());
    c->launches += 1;
    CK(cudaMemcpyAsync(rgba8, c->display, npx * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return collect_timing(c);
}

// pt_readback_display without the wait: the display pass runs now, on the context's stream, into the staging image (which IS
// the snapshot: 8.3 MB at 1080p), and the copy to pinned host memory rides on the copy stream under the next pt_render_spp.
int pt_readback_display_async(pt_ctx* c, unsigned char* rgba8_pinned, size_t bytes) {
    if (!c || !rgba8_pinned) return fail(PT_EINVAL, "pt_readback_display_async: null argument");
    const size_t npx = (size_t)c->cfg.width * c->cfg.height;
    if (bytes < npx * 4) return fail(PT_EINVAL, "pt_readback_display_async: buffer too small");
    CK(cudaSetDevice(c->cfg.device));
    int rc = pt_copy_setup(c);
    if (rc != PT_OK) return rc;
    if (!c->display) CK(cudaMalloc(&c->display, npx * sizeof(uchar4)));
    if (c->copy_pending) CK(cudaStreamWaitEvent(c->stream, c->ev_copied, 0));     // the previous copy still reads the staging image
    dim3 block(32, 8), grid((c->cfg.width + 31) / 32, (c->cfg.height + 7) / 8);
    k_display<<<grid, block, 0, c->stream>>>(c->accum, c->cfg.width, c->cfg.height, c->display);
    CK(cudaGetLastError());
    c->launches += 1;
    return pt_copy_enqueue(c, rgba8_pinned, c->display, npx * 4, true);
}

int pt_readback_primary_ids(pt_ctx* c, int* ids, size_t bytes) {
    if (!c || !ids) return fail(PT_EINVAL, "pt_readback_primary_ids: null argument");
    size_t need = (size_t)c->cfg.width * c->cfg.height * sizeof(int);
    if (bytes < need) return fail(PT_EINVAL, "pt_readback_primary_ids: buffer too small");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemcpyAsync(ids, c->primary_ids, need, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return collect_timing(c);
}

int pt_enable_cost_map(pt_ctx* c, int on) {
    if (!c) return fail(PT_EINVAL, "pt_enable_cost_map: null context");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaStreamSynchronize(c->stream));
    if (on && !c->cost) {
        size_t npx = (size_t)c->cfg.width * c->cfg.height;
        CK(cudaMalloc(&c->cost, npx * sizeof(unsigned)));
        CK(cudaMemset(c->cost, 0, npx * sizeof(unsigned)));
    } else if (!on && c->cost) {
        cudaFree(c->cost);
        c->cost = nullptr;
    }
    return PT_OK;
}

int pt_readback_cost_map(pt_ctx* c, unsigned* cycles, size_t bytes) {
    if (!c || !cycles) return fail(PT_EINVAL, "pt_readback_cost_map: null argument");
    if (!c->cost) return fail(PT_EINVAL, "pt_readback_cost_map: pt_enable_cost_map(ctx, 1) first");
    size_t need = (size_t)c->cfg.width * c->cfg.height * sizeof(unsigned);
    if (bytes < need) return fail(PT_EINVAL, "pt_readback_cost_map: buffer too small");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemcpyAsync(cycles, c->cost, need, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return collect_timing(c);
}

int pt_framebuffer_device_ptr(pt_ctx* c, void** dptr, size_t* bytes) {
    if (!c || !dptr) return fail(PT_EINVAL, "pt_framebuffer_device_ptr: null argument");
    *dptr = c->accum;
    if (bytes) *bytes = (size_t)c->cfg.width * c->cfg.height * sizeof(float4);
    return PT_OK;
}

int pt_stream(pt_ctx* c, void** s) {
    if (!c || !s) return fail(PT_EINVAL, "pt_stream: null argument");
    *s = (void*)c->stream;
    return PT_OK;
}

int pt_scale_framebuffer(pt_ctx* c, float scale) {
    if (!c) return fail(PT_EINVAL, "pt_scale_framebuffer: null context");
    CK(cudaSetDevice(c->cfg.device));
    size_t npx = (size_t)c->cfg.width * c->cfg.height;
    k_scale<<<(unsigned)((npx + 255) / 256), 256, 0, c->stream>>>(c->accum, npx, scale);
    CK(cudaGetLastError());
    c->launches += 1;
    return PT_OK;
}

int pt_synchronize(pt_ctx* c) {
    if (!c) return fail(PT_EINVAL, "pt_synchronize: null context");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaStreamSynchronize(c->stream));
    if (c->copy_pending) {                                   // asynchronous host copies land here
        CK(cudaStreamSynchronize(c->copy_stream));
        c->copy_pending = false;
    }
    if (c->comm) {
        const int rc = pt_comm_check(c);
        if (rc != PT_OK) return rc;
    }
    return collect_timing(c);
}

// ---- asynchronous host copies -------------------------------------------------------------------------------------------
int pt_copy_setup(pt_ctx* c) {
    if (c->copy_stream) return PT_OK;
    // highest priority: a copy, or the small kernels of a pipelined combine (pt_comm.cu), go ahead of the render stream's
    // queued blocks whenever an SM has room
    int prio_lo = 0, prio_hi = 0;
    CK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    CK(cudaStreamCreateWithPriority(&c->copy_stream, cudaStreamNonBlocking, prio_hi));
    CK(cudaEventCreateWithFlags(&c->ev_image, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&c->ev_copied, cudaEventDisableTiming));
    return PT_OK;
}
// `first`: records "image ready" on the context's stream and makes the copy stream wait for it; every copy ends by recording
// "copied", which the NEXT producer of the device image waits for (pt_readback_async's snapshot, pt_combine's barrier A)
int pt_copy_enqueue(pt_ctx* c, void* host, const void* dev, size_t bytes, bool first) {
    if (first) {
        CK(cudaEventRecord(c->ev_image, c->stream));
        CK(cudaStreamWaitEvent(c->copy_stream, c->ev_image, 0));
    }
    CK(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, c->copy_stream));
    CK(cudaEventRecord(c->ev_copied, c->copy_stream));
    c->copy_pending = true;
    return PT_OK;
}

int pt_readback_async(pt_ctx* c, float* rgba_pinned, size_t bytes) {
    if (!c || !rgba_pinned) return fail(PT_EINVAL, "pt_readback_async: null argument");
    const size_t need = (size_t)c->cfg.width * c->cfg.height * sizeof(float4);
    if (bytes < need) return fail(PT_EINVAL, "pt_readback_async: buffer too small");
    CK(cudaSetDevice(c->cfg.device));
    int rc = pt_copy_setup(c);
    if (rc != PT_OK) return rc;
    if (!c->snapshot) CK(cudaMalloc(&c->snapshot, need));
    if (c->copy_pending) CK(cudaStreamWaitEvent(c->stream, c->ev_copied, 0));     // the previous copy still reads the snapshot
    CK(cudaMemcpyAsync(c->snapshot, c->accum, need, cudaMemcpyDeviceToDevice, c->stream));   // 66 MB of HBM traffic at 1080p: ~15 us
    return pt_copy_enqueue(c, rgba_pinned, c->snapshot, need, true);
}

int pt_get_stats(pt_ctx* c, PtStats* out) {
    if (!c || !out) return fail(PT_EINVAL, "pt_get_stats: null argument");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaStreamSynchronize(c->stream));
    int rc = collect_timing(c);
    if (rc != PT_OK) return rc;
    unsigned long long cnt[7] = {0, 0, 0, 0, 0, 0, 0};
    CK(cudaMemcpy(cnt, c->ray_counter, sizeof cnt, cudaMemcpyDeviceToHost));
    out->rays = cnt[0]; out->inner_visits = cnt[1]; out->tri_tests = cnt[2]; out->leaf_candidates = cnt[3]; out->tie_retraces = cnt[4]; out->shaded_hits = cnt[5]; out->env_texels = cnt[6]; out->samples = c->samples; out->render_ms = c->render_ms; out->launches = c->launches;
    out->max_stack_depth = c->has_scene ? c->sc.stack_depth : 0;
    out->n_inner_nodes = c->has_scene ? c->sc.nInner : 0;
    out->n_materials = c->has_scene ? c->sc.nMaterials : 0;
    out->fast_tree_depth = c->fast_depth; out->n_fast_inner = c->n_fast_inner; out->upload_build_ms = c->build_ms;
    return PT_OK;
}

int pt_get_ray_histogram(pt_ctx* c, unsigned long long* buckets, int n) {
    if (!c || !buckets || n < 33) return fail(PT_EINVAL, "pt_get_ray_histogram: need room for 33 buckets");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaMemcpy(buckets, c->ray_counter + 8, 33 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    return PT_OK;
}

int pt_reset_stats(pt_ctx* c) {
    if (!c) return fail(PT_EINVAL, "pt_reset_stats: null context");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaStreamSynchronize(c->stream));
    int rc = collect_timing(c);
    if (rc != PT_OK) return rc;
    CK(cudaMemset(c->ray_counter, 0, 48 * sizeof(unsigned long long)));
    c->samples = 0; c->launches = 0; c->render_ms = 0.0;
    return PT_OK;
}

}  // extern "C"
