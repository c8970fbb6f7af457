// calculateHdrCache (ImportanceSampling_LowDiscrepancySequence/main.cpp:555-652, "M:") on the GPU — SURVEY.md §8f.3.
//
// The table is part of the spec bit for bit: its fp32 running sums drift (the pdf channel sums to 1.0001 … 1.0077 on the
// shipped maps) and the shader inverts exactly that table.  So the sums that are sequential in the reference stay sequential
// here, in the reference's order, with explicit round-to-nearest operations; what is independent runs in parallel:
//   luminance (the one expression the reference evaluates in double, M:567)  one thread per texel
//   lumSum over all texels, row-major (M:568)                                one thread, operands staged through shared memory
//   pdf /= lumSum (M:573-576)                                                one thread per texel
//   column marginals and conditional CDFs (M:579-609)                        one thread per column, rows in order
//   cdf_x prefix over the columns (M:586-588)                                one thread
//   the two lower_bound searches per table entry (M:621-636)                 one thread per entry, libstdc++'s halving loop
// Checked bit for bit against libptscene's host restatement (itself checked against the reference's own code) in tests.
#include <cuda_runtime.h>

#include <cstdint>
#include <mutex>
#include <string>

#include "pt_internal.h"

namespace pt {
namespace {

__global__ void k_luminance(const float* __restrict__ rgb, size_t n, float* __restrict__ pdf) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double r = (double)rgb[3 * i], g = (double)rgb[3 * i + 1], b = (double)rgb[3 * i + 2];
    pdf[i] = __double2float_rn(__dadd_rn(__dadd_rn(__dmul_rn(0.2, r), __dmul_rn(0.7, g)), __dmul_rn(0.1, b)));
}

// lumSum += lum in texel order: the additions are one dependent chain, the loads are not
#define PT_SEQ_CHUNK 4096
__global__ void __launch_bounds__(1024) k_sequential_sum(const float* __restrict__ v, size_t n, float* __restrict__ out) {
    __shared__ float stage[PT_SEQ_CHUNK];
    float s = 0.0f;
    for (size_t base = 0; base < n; base += PT_SEQ_CHUNK) {
        for (int k = threadIdx.x; k < PT_SEQ_CHUNK; k += blockDim.x) stage[k] = base + k < n ? v[base + k] : 0.0f;
        __syncthreads();
        if (threadIdx.x == 0) {
            const int m = (int)(n - base < (size_t)PT_SEQ_CHUNK ? n - base : (size_t)PT_SEQ_CHUNK);
            for (int k = 0; k < m; k++) s = __fadd_rn(s, stage[k]);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = s;
}

__global__ void k_normalise(float* __restrict__ pdf, size_t n, const float* __restrict__ lum_sum) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) pdf[i] = __fdiv_rn(pdf[i], *lum_sum);
}

// column j: pdf_x[j] = sum_i pdf[i][j]; cdf_y[j][i] = running sum of pdf[i][j] / pdf_x[j].  cdf_y is kept row-major
// ([i][j]) so that both loops are coalesced across the columns.
__global__ void k_columns(const float* __restrict__ pdf, int W, int H, float* __restrict__ pdf_x, float* __restrict__ cdf_y) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= W) return;
    float s = 0.0f;
    for (int i = 0; i < H; i++) s = __fadd_rn(s, pdf[(size_t)i * W + j]);
    pdf_x[j] = s;
    float run = 0.0f;
    for (int i = 0; i < H; i++) {
        const float cond = __fdiv_rn(pdf[(size_t)i * W + j], s);
        run = i == 0 ? cond : __fadd_rn(run, cond);
        cdf_y[(size_t)i * W + j] = run;
    }
}

__global__ void k_prefix(const float* __restrict__ pdf_x, int W, float* __restrict__ cdf_x) {
    if (blockIdx.x || threadIdx.x) return;
    float run = pdf_x[0];
    cdf_x[0] = run;
    for (int j = 1; j < W; j++) { run = __fadd_rn(pdf_x[j], run); cdf_x[j] = run; }
}

// std::lower_bound as libstdc++ writes it (first position whose element is not < val), over a strided sequence
__device__ __forceinline__ int lower_bound_strided(const float* __restrict__ a, size_t stride, int n, float val) {
    int first = 0, len = n;
    while (len > 0) {
        const int half = len >> 1, middle = first + half;
        if (a[(size_t)middle * stride] < val) { first = middle + 1; len = len - half - 1; }
        else len = half;
    }
    return first;
}

// entry (i, j): xi_1 = i / H picks the column x through cdf_x, xi_2 = j / W the row y through column x's cdf (the
// reference's transposed roles of i and j, M:621-636); out = (x / W, y / H, pdf[i][j])
template <bool FLOAT4>
__global__ void k_lookup(const float* __restrict__ pdf, const float* __restrict__ cdf_x, const float* __restrict__ cdf_y, int W, int H,
                         float* __restrict__ out) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)W * H) return;
    const int i = (int)(idx / (size_t)W), j = (int)(idx - (size_t)i * W);
    const float xi_1 = __fdiv_rn((float)i, (float)H), xi_2 = __fdiv_rn((float)j, (float)W);
    int x = lower_bound_strided(cdf_x, 1, W, xi_1);
    if (x >= W) x = W - 1;                             // the reference would index past the end here; never hit on real maps
    const int y = lower_bound_strided(cdf_y + x, (size_t)W, H, xi_2);
    const float ox = __fdiv_rn((float)x, (float)W), oy = __fdiv_rn((float)y, (float)H), op = pdf[idx];
    if (FLOAT4) reinterpret_cast<float4*>(out)[idx] = make_float4(ox, oy, op, 0.0f);
    else { out[3 * idx] = ox; out[3 * idx + 1] = oy; out[3 * idx + 2] = op; }
}

// cdf_y [i][j] (coalesced for k_columns) -> [j][i]: the column a binary search walks is contiguous
__global__ void k_transpose(const float* __restrict__ in, int W, int H, float* __restrict__ out) {
    __shared__ float tile[32][33];
    const int j0 = blockIdx.x * 32, i0 = blockIdx.y * 32;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int i = i0 + r, j = j0 + threadIdx.x;
        if (i < H && j < W) tile[r][threadIdx.x] = in[(size_t)i * W + j];
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int j = j0 + r, i = i0 + threadIdx.x;
        if (i < H && j < W) out[(size_t)j * H + i] = tile[threadIdx.x][r];
    }
}

}  // namespace

// The CDF-search environment sampler's tables (PtConfig.env_sampler = PT_ENV_CDF_SEARCH): the intermediates of
// calculateHdrCache, M:555-609, by the same kernels as the table build below — pdf [H][W], the marginal CDF over the
// columns cdf_x [W], the conditional CDFs column by column cdf_yT [W][H].  The caller owns the three allocations.
cudaError_t hdr_cdf_device(const float* d_rgb, int W, int H, float** pdf_out, float** cdf_x_out, float** cdf_yT_out, cudaStream_t st) {
    const size_t n = (size_t)W * H;
    float *pdf = nullptr, *cdf_y = nullptr, *cdf_yT = nullptr, *small = nullptr, *cdf_x = nullptr;
    cudaError_t e = cudaMalloc(&pdf, n * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&cdf_y, n * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&cdf_yT, n * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&small, ((size_t)W + 1) * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&cdf_x, (size_t)W * sizeof(float));
    if (e == cudaSuccess) {
        float *pdf_x = small, *lum_sum = small + W;
        const unsigned blocks = (unsigned)((n + 255) / 256);
        k_luminance<<<blocks, 256, 0, st>>>(d_rgb, n, pdf);
        k_sequential_sum<<<1, 1024, 0, st>>>(pdf, n, lum_sum);
        k_normalise<<<blocks, 256, 0, st>>>(pdf, n, lum_sum);
        k_columns<<<(W + 127) / 128, 128, 0, st>>>(pdf, W, H, pdf_x, cdf_y);
        k_prefix<<<1, 32, 0, st>>>(pdf_x, W, cdf_x);
        k_transpose<<<dim3((W + 31) / 32, (H + 31) / 32), dim3(32, 8), 0, st>>>(cdf_y, W, H, cdf_yT);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    }
    cudaFree(cdf_y); cudaFree(small);
    if (e != cudaSuccess) { cudaFree(pdf); cudaFree(cdf_yT); cudaFree(cdf_x); return e; }
    *pdf_out = pdf; *cdf_x_out = cdf_x; *cdf_yT_out = cdf_yT;
    return cudaSuccess;
}

// d_rgb: W*H*3 floats on the device (file row order).  Exactly one of d_cache4 (W*H float4, the renderer's layout) and
// d_cache3 (W*H*3 floats, the reference's layout) receives the table.  Temporaries are allocated and freed here.
cudaError_t hdr_cache_device(const float* d_rgb, int W, int H, float4* d_cache4, float* d_cache3, cudaStream_t st) {
    const size_t n = (size_t)W * H;
    float *pdf = nullptr, *cdf_y = nullptr, *small = nullptr;
    cudaError_t e = cudaMalloc(&pdf, n * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&cdf_y, n * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&small, (2 * (size_t)W + 1) * sizeof(float));
    if (e == cudaSuccess) {
        float *pdf_x = small, *cdf_x = small + W, *lum_sum = small + 2 * (size_t)W;
        const unsigned blocks = (unsigned)((n + 255) / 256);
        k_luminance<<<blocks, 256, 0, st>>>(d_rgb, n, pdf);
        k_sequential_sum<<<1, 1024, 0, st>>>(pdf, n, lum_sum);
        k_normalise<<<blocks, 256, 0, st>>>(pdf, n, lum_sum);
        k_columns<<<(W + 127) / 128, 128, 0, st>>>(pdf, W, H, pdf_x, cdf_y);
        k_prefix<<<1, 32, 0, st>>>(pdf_x, W, cdf_x);
        if (d_cache4) k_lookup<true><<<blocks, 256, 0, st>>>(pdf, cdf_x, cdf_y, W, H, reinterpret_cast<float*>(d_cache4));
        else k_lookup<false><<<blocks, 256, 0, st>>>(pdf, cdf_x, cdf_y, W, H, d_cache3);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    }
    cudaFree(pdf); cudaFree(cdf_y); cudaFree(small);
    return e;
}

}  // namespace pt

// ---- pt_calculate_hdr_cache_noalloc: the same table without a single allocation ------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The temporaries hdr_cache_device()
// allocates per call (pdf, conditional CDFs, marginal pdf / CDF, lumSum) live in ONE file-scope device array of a fixed,
// modest size; every GPU that loads libpt.so holds those 2.1 MB until the process exits.
//   * In host code `g_noalloc_scratch` is only a host-side stand-in for the symbol, NOT a device address: its address is
//     obtained once per device, with that device current, through cudaGetSymbolAddress, and that pointer is what the
//     kernels receive.
//   * Every call on a device shares the array, which keeps the previous call's contents.  Nothing in it has to start at
//     zero: each kernel below writes every element a later kernel reads (pdf, lumSum, pdf_x, cdf_y, cdf_x are all fully
//     overwritten for the W x H of the call), so no clearing pass is needed.  Calls that overlap on the GPU (two streams,
//     two threads) race on the array: serialise them.
#define PT_NOALLOC_MAX_TEXELS (1 << 18)       // W * H <= 262,144 (512 x 512, 1024 x 256, ...)
#define PT_NOALLOC_MAX_WIDTH 4096
#define PT_NOALLOC_FLOATS (2 * PT_NOALLOC_MAX_TEXELS + 2 * PT_NOALLOC_MAX_WIDTH + 64)
static __device__ __align__(256) float g_noalloc_scratch[PT_NOALLOC_FLOATS];

namespace {
std::mutex g_noalloc_mutex;
float* g_noalloc_ptr[64] = {};                // per device: the DEVICE address of g_noalloc_scratch
}  // namespace

extern "C" int pt_calculate_hdr_cache_noalloc(int device, const float* d_hdr, int width, int height, float* d_cache, void* cuda_stream) {
    if (!d_hdr || !d_cache || width <= 0 || height <= 0) return fail(PT_EINVAL, "pt_calculate_hdr_cache_noalloc: null buffer or empty image");
    if (width > PT_NOALLOC_MAX_WIDTH || (long long)width * height > PT_NOALLOC_MAX_TEXELS)
        return fail(PT_ERANGE, "pt_calculate_hdr_cache_noalloc: " + std::to_string(width) + " x " + std::to_string(height) +
                               " exceeds the static scratch (width <= " + std::to_string(PT_NOALLOC_MAX_WIDTH) + ", width * height <= " +
                               std::to_string(PT_NOALLOC_MAX_TEXELS) + "); use pt_calculate_hdr_cache, which allocates");
    if (device < 0 || device >= 64) return fail(PT_EINVAL, "pt_calculate_hdr_cache_noalloc: device ordinal out of range");
    CK(cudaSetDevice(device));
    float* base = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_noalloc_mutex);
        if (!g_noalloc_ptr[device]) {                       // set-up, once per device, with that device current
            void* p = nullptr;
            CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
            g_noalloc_ptr[device] = static_cast<float*>(p);
        }
        base = g_noalloc_ptr[device];
    }
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const int W = width, H = height;
    const size_t n = (size_t)W * H;
    float *pdf = base, *cdf_y = base + PT_NOALLOC_MAX_TEXELS, *pdf_x = base + 2 * PT_NOALLOC_MAX_TEXELS,
          *cdf_x = pdf_x + PT_NOALLOC_MAX_WIDTH, *lum_sum = cdf_x + PT_NOALLOC_MAX_WIDTH;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    pt::k_luminance<<<blocks, 256, 0, st>>>(d_hdr, n, pdf);
    pt::k_sequential_sum<<<1, 1024, 0, st>>>(pdf, n, lum_sum);
    pt::k_normalise<<<blocks, 256, 0, st>>>(pdf, n, lum_sum);
    pt::k_columns<<<(W + 127) / 128, 128, 0, st>>>(pdf, W, H, pdf_x, cdf_y);
    pt::k_prefix<<<1, 32, 0, st>>>(pdf_x, W, cdf_x);
    pt::k_lookup<false><<<blocks, 256, 0, st>>>(pdf, cdf_x, cdf_y, W, H, d_cache);
    CK(cudaGetLastError());
    return PT_OK;
}

