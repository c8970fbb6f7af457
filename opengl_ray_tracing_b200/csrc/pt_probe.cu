// pt_probe.cu — measurement probes of libpt.so (no reference equivalent; SURVEY.md §8d: "L2 peak must be measured on the
// box with a read microbenchmark over a <= 64 MB ld.global.nc window and recorded").
//
// pt_measure_l2_read_gbs: every resident thread streams 16-byte ld.global.nc loads over a window that fits L2 (126 MB on
// B200) but not the SMs' L1s: the slice a CTA reads ROTATES from pass to pass, so no SM sees an address twice within
// L1's reach.  The figure is the roofline denominator for the traversal's node/triangle fetches, which ncu reports as
// lts__t_sectors_srcunit_tex_op_read x 32 B / time.
#include <cuda_runtime.h>

#include <string>

#include "../../include/pt.h"

namespace {

__global__ void __launch_bounds__(256) k_l2_read(const float4* __restrict__ buf, size_t n_vec, int passes, float4* __restrict__ sink) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const size_t per_cta = (n_vec + gridDim.x - 1) / gridDim.x;          // one contiguous slice per CTA and pass
    for (int p = 0; p < passes; p++) {
        const size_t cta = ((size_t)blockIdx.x + (size_t)p * 61u) % gridDim.x;   // rotate the slices: 61 is prime to any grid used here
        const size_t lo = cta * per_cta, hi = lo + per_cta < n_vec ? lo + per_cta : n_vec;
        size_t i = lo + threadIdx.x;
        for (; i + 3 * 256 < hi; i += 4 * 256) {                          // four independent 16-byte loads in flight per thread
            const float4 a = __ldg(buf + i), b = __ldg(buf + i + 256), c = __ldg(buf + i + 512), d = __ldg(buf + i + 768);
            acc.x += a.x + b.x; acc.y += a.y + b.y; acc.z += c.z + d.z; acc.w += c.w + d.w;
        }
        for (; i < hi; i += 256) { const float4 a = __ldg(buf + i); acc.x += a.x; }
    }
    if (acc.x + acc.y + acc.z + acc.w == 123456.789f) *sink = acc;        // never true for the zero-filled window: keeps the loads alive
}

}  // namespace

extern "C" int pt_measure_l2_read_gbs(int device, int window_mb, int passes, double* gbs, double* hbm_gbs) {
    if (!gbs || window_mb <= 0 || passes <= 0) return PT_EINVAL;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return PT_EINVAL;
    if (cudaSetDevice(device) != cudaSuccess) return PT_ECUDA;
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const size_t bytes = (size_t)window_mb << 20, n_vec = bytes / sizeof(float4);
    float4 *buf = nullptr, *sink = nullptr, *big = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaError_t e = cudaMalloc(&buf, bytes);
    if (e == cudaSuccess) e = cudaMalloc(&sink, sizeof(float4));
    if (e == cudaSuccess) e = cudaMemset(buf, 0, bytes);
    if (e == cudaSuccess) e = cudaEventCreate(&e0);
    if (e == cudaSuccess) e = cudaEventCreate(&e1);
    float best = 0.f, best_hbm = 0.f;
    const int grid = (sms > 0 ? sms : 148) * 8;
    if (e == cudaSuccess) {
        k_l2_read<<<grid, 256>>>(buf, n_vec, 2, sink);                   // warm-up: the window becomes L2-resident
        for (int rep = 0; rep < 5 && e == cudaSuccess; rep++) {
            cudaEventRecord(e0);
            k_l2_read<<<grid, 256>>>(buf, n_vec, passes, sink);
            cudaEventRecord(e1);
            e = cudaEventSynchronize(e1);
            float ms = 0.f;
            if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
            const float g = (float)((double)bytes * passes / (ms * 1e-3) / 1e9);
            if (g > best) best = g;
        }
    }
    if (e == cudaSuccess && hbm_gbs) {
        // the same kernel over a window far larger than L2 (2 GiB, one pass): the HBM read figure of THIS access pattern
        const size_t big_bytes = (size_t)2048 << 20;
        e = cudaMalloc(&big, big_bytes);
        if (e == cudaSuccess) e = cudaMemset(big, 0, big_bytes);
        for (int rep = 0; rep < 3 && e == cudaSuccess; rep++) {
            cudaEventRecord(e0);
            k_l2_read<<<grid, 256>>>(big, big_bytes / sizeof(float4), 1, sink);
            cudaEventRecord(e1);
            e = cudaEventSynchronize(e1);
            float ms = 0.f;
            if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
            const float g = (float)((double)big_bytes / (ms * 1e-3) / 1e9);
            if (g > best_hbm) best_hbm = g;
        }
        *hbm_gbs = best_hbm;
    }
    cudaFree(buf); cudaFree(sink); cudaFree(big);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (e != cudaSuccess) return PT_ECUDA;
    *gbs = best;
    return PT_OK;
}
