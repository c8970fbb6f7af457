// pt_wavefront.cuh — the wavefront pipeline: the same per-sample arithmetic as shade_sample() (pt_shade.cuh), cut
// at every hitBVH call into small kernels that hand paths to each other through queues in HBM.
//
// Why: the megakernel is 5-7 thousand SASS instructions (85-107 KB) against a 32 KB L1.5 instruction cache; with
// warps of one SM spread over the whole program the dominant stall was "no instruction" (5.2 cycles per issue,
// profiles/r01_megakernel_v2.md).  Here every kernel is one phase, all resident warps run the same few hundred
// instructions, terminated paths drop out of the queues (compaction) instead of idling lanes, and every ray is its
// own work item (no pixel-sized stragglers).
//
// One frame (= one sample of every pixel of the region), bounce b = 0 .. max_bounce-1:
//   generate   ray generation P:846-850 + primary hitBVH; miss -> sampleHdr and accumulate; hit -> Q_shade[0]
//   shade(b)   for Q_shade[b]: hit record -> normal, material; closes the previous bounce (emissive term, history);
//              b < max_bounce: environment light sample -> shadow ray + its MIS-weighted term (variant 2), BRDF sample
//              -> next ray, f_r, pdf.  Feeds Q_shadow[b] and Q_extend[b].
//   shadow(b)  for Q_shadow[b]: any-hit traversal; occluded -> the light term is zeroed
//   extend(b)  for Q_extend[b]: closest-hit traversal; miss -> environment term (MIS-weighted in variant 2), path
//              ends; hit -> Q_shade[b+1]
//   accumulate for Q_shade[0]: colour = Le0 + sum of the path's terms IN THE REFERENCE'S ORDER, then the running mean
//
// Exactness: every radiance contribution of a path is stored as its own term and summed at the end in the order
// the reference adds them (light term then BRDF-ray term, bounce by bounce), skipped contributions are +0 (x + 0 is
// exact), so under the parity arithmetic the pipeline is bit-identical to the megakernel and the oracle.
//
// Path state lives in HBM, indexed by pixel (path id = py * width + px), struct-of-arrays; queues hold 4-byte ids.
#pragma once
#include "pt_shade.cuh"

namespace pt {

#define PT_WF_CHUNK 64            // queue entries a warp claims per atomic

__device__ __forceinline__ int wf_light_slot(int b) { return 1 + 2 * b; }
template <int VARIANT> __device__ __forceinline__ int wf_bounce_slot(int b) { return VARIANT == 2 ? 2 + 2 * b : 1 + b; }
template <int VARIANT> __host__ __device__ __forceinline__ int wf_num_slots(int max_bounce) {
    return VARIANT == 2 ? 1 + 2 * max_bounce : 1 + max_bounce;
}

// queue counters: 32-byte apart is not enough (same L2 atomic unit) — one counter per 256 bytes
#define PT_WF_CTR_STRIDE 64       // in uint32
enum { WF_Q_SHADE = 0, WF_Q_SHADOW = PT_WF_MAX_BOUNCE + 1, WF_Q_EXTEND = 2 * PT_WF_MAX_BOUNCE + 1, WF_Q_GEN = 3 * PT_WF_MAX_BOUNCE + 1,
       WF_NUM_Q = 3 * PT_WF_MAX_BOUNCE + 2 };
__device__ __forceinline__ uint32_t* wf_tail(const WaveBuffers& wf, int q) { return wf.counters + (size_t)q * PT_WF_CTR_STRIDE; }
__device__ __forceinline__ uint32_t* wf_head(const WaveBuffers& wf, int q) { return wf.counters + (size_t)(WF_NUM_Q + q) * PT_WF_CTR_STRIDE; }

// warp-aggregated append: all 32 lanes call this; returns nothing, writes id into queue for lanes with pred
__device__ __forceinline__ void wf_push(uint32_t* queue, uint32_t* tail, bool pred, uint32_t id) {
    const uint32_t m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return;
    const int lane = threadIdx.x & 31;
    uint32_t base = 0;
    if (lane == (__ffs(m) - 1)) base = atomicAdd(tail, (uint32_t)__popc(m));
    base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
    if (pred) queue[base + __popc(m & ((1u << lane) - 1u))] = id;
}

// a warp claims the next PT_WF_CHUNK entries of a work list of `count` entries; returns the first, or >= count
__device__ __forceinline__ uint32_t wf_claim(uint32_t* head) {
    uint32_t base = 0;
    if ((threadIdx.x & 31) == 0) base = atomicAdd(head, (uint32_t)PT_WF_CHUNK);
    return __shfl_sync(0xffffffffu, base, 0);
}

__device__ __forceinline__ void wf_flush_rays(uint32_t nrays, unsigned long long* ray_counter) {
    for (int o = 16; o > 0; o >>= 1) nrays += __shfl_xor_sync(0xffffffffu, nrays, o);
    if ((threadIdx.x & 31) == 0 && nrays) atomicAdd(ray_counter, (unsigned long long)nrays);
}

__device__ __forceinline__ void wf_accumulate_pixel(const RenderParams& p, size_t idx, uint32_t frame, V3 c) {
    float4 acc = p.accum[idx];
    if (p.accumulate_sum) {
        acc.x = s_add(acc.x, c.x); acc.y = s_add(acc.y, c.y); acc.z = s_add(acc.z, c.z);
    } else {                                          // mix(lastColor, color, 1/(frameCounter+1)), P:868-869
        float a = s_div(1.0f, (float)(frame + 1u));
        float b = s_sub(1.0f, a);
        acc.x = s_add(s_mul(acc.x, b), s_mul(c.x, a));
        acc.y = s_add(s_mul(acc.y, b), s_mul(c.y, a));
        acc.z = s_add(s_mul(acc.z, b), s_mul(c.z, a));
    }
    acc.w = 1.0f;
    p.accum[idx] = acc;
}

__device__ __forceinline__ float4 f4(V3 v, float w) { return make_float4(v.x, v.y, v.z, w); }
__device__ __forceinline__ V3 xyz(float4 v) { return mk3(v.x, v.y, v.z); }

// ---- generate + primary ray ----------------------------------------------------------------------------
// work item = one 8x4 pixel tile per warp iteration (same pixel <-> lane mapping as the megakernel)
template <class Cfg, int VARIANT>
__global__ void __launch_bounds__(256) wf_generate(const __grid_constant__ RenderParams p, uint32_t frame, int write_ids) {
    const WaveBuffers& wf = p.wf;
    const int lane = threadIdx.x & 31;
    const TileMap tm = make_tile_map(p.x0, p.y0, p.x1, p.y1, p.band_world, p.band_rank, p.band_diag);
    const uint32_t ntiles = tm.count();                // owned 8x4 warp tiles
    uint32_t nrays = 0;
    Counter<Cfg::COUNT> cnt;
    for (;;) {
        const uint32_t base = wf_claim(wf_head(wf, WF_Q_GEN));       // PT_WF_CHUNK pixels = 2 tiles per claim
        if (base >= ntiles * 32u) break;
#pragma unroll 1
        for (uint32_t k = 0; k < PT_WF_CHUNK; k += 32) {
            const uint32_t tile = (base + k) >> 5;
            bool valid = tile < ntiles;
            int tx = 0, strip = 0;
            if (valid) tm.locate(tile, tx, strip);                   // owned tiles as TileMap numbers them
            const int px = p.x0 + tx * 8 + (lane & 7);
            const int py = p.y0 + strip * 4 + (lane >> 3);
            valid = valid && px < p.x1 && py < p.y1;
            bool alive = false;
            const size_t idx = (size_t)py * p.width + px;
            if (valid) {
                uint32_t seed = ((uint32_t)px * 1973u + (uint32_t)py * 9277u + frame * 26699u) | 1u;     // P:73-76
                V3 o, d;
                generate_ray(p, px, py, seed, o, d);
                const TraceResult tr = trace<Cfg>(p.sc, o, d, false, cnt);
                nrays++;
                if (write_ids) p.primary_ids[idx] = tr.tri;
                if (tr.tri < 0) {
                    wf_accumulate_pixel(p, idx, frame, sampleHdr<typename Cfg::M>(p.sc, d));             // P:857-859
                } else {
                    alive = true;
                    wf.ray_o[idx] = f4(o, 1.0f);
                    wf.ray_d[idx] = f4(d, 0.0f);
                    wf.hit[idx] = make_int2(tr.tri | (tr.inside ? (int)0x80000000 : 0), __float_as_int(tr.t));
                    wf.seed[idx] = seed;
                }
            }
            wf_push(wf.q_shade[0], wf_tail(wf, WF_Q_SHADE), alive, (uint32_t)idx);
        }
    }
    wf_flush_rays(nrays, p.ray_counter);
    flush_counts(cnt, p.work_counters);
}

// ---- shade ------------------------------------------------------------------------------------------------
template <class Cfg, int VARIANT>
__global__ void __launch_bounds__(256) wf_shade(const __grid_constant__ RenderParams p, uint32_t frame, int b) {
    typedef typename Cfg::M M;
    const WaveBuffers& wf = p.wf;
    const DevScene& sc = p.sc;
    const int lane = threadIdx.x & 31;
    const uint32_t count = *wf_tail(wf, WF_Q_SHADE + b);
    const uint32_t* q = wf.q_shade[b];
    const size_t np = wf.npaths;
    for (;;) {
        const uint32_t base = wf_claim(wf_head(wf, WF_Q_SHADE + b));
        if (base >= count) break;
#pragma unroll 1
        for (uint32_t k = 0; k < PT_WF_CHUNK; k += 32) {
            const uint32_t i = base + k + lane;
            const bool valid = i < count;
            bool want_shadow = false, want_extend = false;
            uint32_t idx = 0;
            if (valid) {
                idx = q[i];
                const float4 ro = wf.ray_o[idx], rd = wf.ray_d[idx];
                const int2 hr = wf.hit[idx];
                TraceResult tr;
                tr.tri = hr.x & 0x7fffffff; tr.inside = hr.x < 0; tr.t = __int_as_float(hr.y);
                const V3 o = xyz(ro), d = xyz(rd);
                Hit hit;
                int mat_id;
                finalize_hit<Cfg::RAW>(sc, o, d, tr, hit, mat_id);
                Material mat;
                load_material<Cfg::RAW>(sc, mat_id, mat);
                V3 history = mk3(1, 1, 1);
                if (b == 0) {
                    wf.terms[idx] = f4(mat.emissive, 0.0f);                                               // Le0, P:862
                } else {
                    const V3 f_r = xyz(wf.fr[idx]);
                    const float cosw = rd.w, pdf = ro.w;
                    history = xyz(wf.hist[idx]);
                    wf.terms[(size_t)wf_bounce_slot<VARIANT>(b - 1) * np + idx] = f4(history * mat.emissive * f_r * cosw / pdf, 0.0f);   // P:833-834
                    history = history * (f_r * cosw / pdf);                                               // P:837
                }
                if (b >= p.max_bounce) {
                    wf.nb[idx] = b;
                } else {
                    uint32_t seed = wf.seed[idx];
                    const V3 V = -hit.viewDir, N = hit.normal;
                    V3 L, f_r = mk3(0, 0, 0);
                    float cosw, pdf = 1.0f;
                    bool ok = true;
                    if (VARIANT == 2) {
                        // light sampling through the hdrCache + MIS, P:770-793.  The term is computed BEFORE the shadow
                        // ray is traced (nothing in it depends on the trace); the shadow kernel zeroes it if occluded.
                        const float e1 = rnd(seed);
                        const float e2 = rnd(seed);
                        const V3 Ld = SampleHdr<M>(sc, e1, e2);
                        V3 Cs = mk3(0, 0, 0);
                        if (dot(N, Ld) > 0.0f) {
                            V3 color; float pdf_light;
                            hdr_color_pdf<M>(sc, Ld, color, pdf_light);
                            const V3 fl = BRDF_Evaluate_iso<M>(V, N, Ld, mat);
                            const float pdf_b = BRDF_Pdf<M>(V, N, Ld, mat);
                            const float mis_weight = misMixWeight(pdf_light, pdf_b);
                            Cs = (history * mis_weight) * color * fl * dot(N, Ld) / pdf_light;
                            wf.sh_d[idx] = f4(Ld, 0.0f);
                            want_shadow = true;
                        }
                        wf.terms[(size_t)wf_light_slot(b) * np + idx] = f4(Cs, 0.0f);
                        // BRDF sampling driven by Sobol + Cranley-Patterson, P:795-816
                        const int px = (int)(idx % (uint32_t)p.width), py = (int)(idx / (uint32_t)p.width);
                        float u = sobol((uint32_t)b * 2u, (frame + 1u) ^ ((frame + 1u) >> 1));
                        float v = sobol((uint32_t)b * 2u + 1u, (frame + 1u) ^ ((frame + 1u) >> 1));
                        cranley_patterson(px, py, u, v);
                        const float xi_3 = rnd(seed);
                        L = SampleBRDF<M>(u, v, xi_3, V, N, mat);
                        cosw = dot(N, L);
                        ok = cosw > 0.0f;
                        if (ok) {
                            f_r = BRDF_Evaluate_iso<M>(V, N, L, mat);
                            pdf = BRDF_Pdf<M>(V, N, L, mat);
                            ok = pdf > 0.0f;
                        }
                    } else {
                        // uniform hemisphere, O:334-346 / D:448-456
                        const float z = rnd(seed);
                        const float r = fmaxf(0.0f, sqrtf(1.0f - z * z));
                        const float phi = 2.0f * PT_PI * rnd(seed);
                        L = toNormalHemisphere(mk3(r * M::cos(phi), r * M::sin(phi), z), N);
                        pdf = 1.0f / (2.0f * PT_PI);
                        cosw = fmaxf(0.0f, dot(L, N));
                        if (VARIANT == 1) {
                            V3 tangent, bitangent;
                            getTangent(N, tangent, bitangent);
                            f_r = BRDF_Evaluate_aniso<M>(V, N, L, tangent, bitangent, mat);
                        } else {
                            f_r = mat.baseColor / PT_PI;
                        }
                    }
                    wf.ray_o[idx] = f4(hit.hitPoint, pdf);          // origin of both the shadow ray and the next ray
                    wf.seed[idx] = seed;
                    if (ok) {
                        wf.ray_d[idx] = f4(L, cosw);
                        wf.hist[idx] = f4(history, 0.0f);
                        wf.fr[idx] = f4(f_r, 0.0f);
                        want_extend = true;
                    } else {
                        wf.terms[(size_t)wf_bounce_slot<VARIANT>(b) * np + idx] = make_float4(0.f, 0.f, 0.f, 0.f);
                        wf.nb[idx] = b + 1;
                    }
                }
            }
            if (VARIANT == 2) wf_push(wf.q_shadow, wf_tail(wf, WF_Q_SHADOW + b), want_shadow, idx);
            wf_push(wf.q_extend, wf_tail(wf, WF_Q_EXTEND + b), want_extend, idx);
        }
    }
}

// ---- shadow rays --------------------------------------------------------------------------------------------
template <class Cfg, int VARIANT>
__global__ void __launch_bounds__(256) wf_shadow(const __grid_constant__ RenderParams p, int b) {
    const WaveBuffers& wf = p.wf;
    const int lane = threadIdx.x & 31;
    const uint32_t count = *wf_tail(wf, WF_Q_SHADOW + b);
    const uint32_t* q = wf.q_shadow;
    uint32_t nrays = 0;
    Counter<Cfg::COUNT> cnt;
    for (;;) {
        const uint32_t base = wf_claim(wf_head(wf, WF_Q_SHADOW + b));
        if (base >= count) break;
#pragma unroll 1
        for (uint32_t k = 0; k < PT_WF_CHUNK; k += 32) {
            const uint32_t i = base + k + lane;
            if (i < count) {
                const uint32_t idx = q[i];
                const TraceResult sh = trace<Cfg>(p.sc, xyz(wf.ray_o[idx]), xyz(wf.sh_d[idx]), true, cnt);
                nrays++;
                if (sh.tri >= 0) wf.terms[(size_t)wf_light_slot(b) * wf.npaths + idx] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
    }
    wf_flush_rays(nrays, p.ray_counter);
    flush_counts(cnt, p.work_counters);
}

// ---- extension rays -------------------------------------------------------------------------------------------
template <class Cfg, int VARIANT>
__global__ void __launch_bounds__(256) wf_extend(const __grid_constant__ RenderParams p, int b) {
    typedef typename Cfg::M M;
    const WaveBuffers& wf = p.wf;
    const int lane = threadIdx.x & 31;
    const uint32_t count = *wf_tail(wf, WF_Q_EXTEND + b);
    const uint32_t* q = wf.q_extend;
    uint32_t nrays = 0;
    Counter<Cfg::COUNT> cnt;
    for (;;) {
        const uint32_t base = wf_claim(wf_head(wf, WF_Q_EXTEND + b));
        if (base >= count) break;
#pragma unroll 1
        for (uint32_t k = 0; k < PT_WF_CHUNK; k += 32) {
            const uint32_t i = base + k + lane;
            bool alive = false;
            uint32_t idx = 0;
            if (i < count) {
                idx = q[i];
                const float4 ro = wf.ray_o[idx], rd = wf.ray_d[idx];
                const V3 d = xyz(rd);
                const TraceResult tr = trace<Cfg>(p.sc, xyz(ro), d, false, cnt);
                nrays++;
                if (tr.tri < 0) {
                    const V3 history = xyz(wf.hist[idx]), f_r = xyz(wf.fr[idx]);
                    const float cosw = rd.w, pdf = ro.w;
                    V3 term;
                    if (VARIANT == 2) {                                                                   // P:819-830
                        V3 color; float pdf_light;
                        hdr_color_pdf<M>(p.sc, d, color, pdf_light);
                        const float mis_weight = misMixWeight(pdf, pdf_light);
                        term = (history * mis_weight) * color * f_r * cosw / pdf;
                    } else {                                                                              // O:349-353, D:465-469
                        const V3 sky = sampleHdr<M>(p.sc, d);
                        term = history * sky * f_r * cosw / pdf;
                    }
                    wf.terms[(size_t)wf_bounce_slot<VARIANT>(b) * wf.npaths + idx] = f4(term, 0.0f);
                    wf.nb[idx] = b + 1;
                } else {
                    alive = true;
                    wf.hit[idx] = make_int2(tr.tri | (tr.inside ? (int)0x80000000 : 0), __float_as_int(tr.t));
                }
            }
            wf_push(wf.q_shade[b + 1], wf_tail(wf, WF_Q_SHADE + b + 1), alive, idx);
        }
    }
    wf_flush_rays(nrays, p.ray_counter);
    flush_counts(cnt, p.work_counters);
}

// ---- accumulate -----------------------------------------------------------------------------------------------
template <class Cfg, int VARIANT>
__global__ void __launch_bounds__(256) wf_accumulate(const __grid_constant__ RenderParams p, uint32_t frame) {
    const WaveBuffers& wf = p.wf;
    const uint32_t count = *wf_tail(wf, WF_Q_SHADE);
    const uint32_t* q = wf.q_shade[0];
    const size_t np = wf.npaths;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const uint32_t idx = q[i];
        const int n = wf.nb[idx];
        V3 Lo = mk3(0, 0, 0);
        for (int b = 0; b < n; b++) {
            if (VARIANT == 2) Lo = Lo + xyz(wf.terms[(size_t)wf_light_slot(b) * np + idx]);
            Lo = Lo + xyz(wf.terms[(size_t)wf_bounce_slot<VARIANT>(b) * np + idx]);
        }
        wf_accumulate_pixel(p, idx, frame, xyz(wf.terms[idx]) + Lo);
    }
}

// ---- host side: one frame after another on the context's stream --------------------------------------------------
template <class Cfg, int VARIANT>
cudaError_t launch_wavefront_variant(const RenderParams& p, int grid, cudaStream_t stream) {
    const WaveBuffers& wf = p.wf;
    const int mb = p.max_bounce;
    for (uint32_t f = 0; f < p.n_frames; f++) {
        const uint32_t frame = p.first_frame + f * p.frame_stride;
        cudaError_t e = cudaMemsetAsync(wf.counters, 0, (size_t)2 * WF_NUM_Q * PT_WF_CTR_STRIDE * sizeof(uint32_t), stream);
        if (e != cudaSuccess) return e;
        wf_generate<Cfg, VARIANT><<<grid, 256, 0, stream>>>(p, frame, f + 1 == p.n_frames ? 1 : 0);
        for (int b = 0; b <= mb; b++) {
            wf_shade<Cfg, VARIANT><<<grid, 256, 0, stream>>>(p, frame, b);
            if (b == mb) break;
            if (VARIANT == 2) wf_shadow<Cfg, VARIANT><<<grid, 256, 0, stream>>>(p, b);
            wf_extend<Cfg, VARIANT><<<grid, 256, 0, stream>>>(p, b);
        }
        wf_accumulate<Cfg, VARIANT><<<grid, 256, 0, stream>>>(p, frame);
    }
    return cudaGetLastError();
}

template <class Cfg>
cudaError_t launch_wavefront(const RenderParams& p, int variant, int grid, cudaStream_t stream) {
    if (variant == 0) return launch_wavefront_variant<Cfg, 0>(p, grid, stream);
    if (variant == 1) return launch_wavefront_variant<Cfg, 1>(p, grid, stream);
    return launch_wavefront_variant<Cfg, 2>(p, grid, stream);
}

// launches per frame: memset is not a kernel; generate + (mb+1) shade + mb extend (+ mb shadow) + accumulate
__host__ inline int wavefront_launches_per_frame(int variant, int max_bounce) {
    return 1 + (max_bounce + 1) + max_bounce * (variant == 2 ? 2 : 1) + 1;
}

}  // namespace pt
