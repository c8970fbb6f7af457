// pt_wavefront.cuh — the wavefront pipeline: the same per-sample arithmetic as shade_sample() (pt_shade.cuh), cut at
// every secondary hitBVH call into kernels that hand rays and paths to each other through queues in HBM.
//
// Round-2 design (the round-1 form processed ONE frame per wave with one ray per lane and no refill, and lost 9x to the
// persistent kernel; profiles/r01_wavefront.md):
//   * a WAVE is F frames of the whole region at once (F = path capacity / pixels: 32 frames at 1080p with the default
//     64 M path slots, ~9 GB of the 180 GB), so every launch carries tens of millions of rays and its tail is noise;
//   * secondary rays — shadow and bounce rays of one bounce together — go through ONE queue into a TRAVERSAL-ONLY
//     persistent kernel (wf_trace) that holds nothing but ray state per lane: a lane whose ray is finished takes the next
//     ray of the warp's chunk of the queue INSIDE the while-while loop (__ballot_sync / __popc ranks, one atomicAdd per
//     chunk of 256 rays), so the node loop runs with full warps instead of the 7-10 of 32 lanes the megakernel averages
//     on mesh rays (north star items 1 and 4);
//   * primary rays are coherent (8x4 pixel tiles) and need no queue: wf_head generates, traces and shades them like the
//     persistent kernel does, then parks the path;
//   * shading runs over COMPACTED queues (wf_shade: paths whose bounce ray hit; then paths whose bounce ray escaped), so
//     all 32 lanes of a warp run the same ~1.5 k instructions of Disney BRDF code.
//
// One wave, bounce b = 0 .. max_bounce-1 (launches: 2 + 2 max_bounce):
//   wf_head        ray generation P:846-850 + primary hitBVH; miss -> sampleHdr, path done; hit -> bounce 0 is shaded:
//                  light sample -> shadow ray + its MIS-weighted term Cs (pending), BRDF sample -> bounce ray, f_r, pdf
//   wf_trace(b)    every ray of bounce b: shadow rays (any hit; occluded -> Cs := 0) and bounce rays (closest hit ->
//                  hit record; the path goes to Q_hit or Q_miss)
//   wf_shade(b+1)  Q_hit: Lo += Cs, emissive term, history (P:833-837), then bounce b+1 as above;
//                  Q_miss: Lo += Cs, environment term (MIS-weighted in variant 2, P:819-830), path done
//   wf_accumulate  per pixel, the wave's F colours folded into the framebuffer IN FRAME ORDER (P:868-869)
//
// Exactness: a path's radiance terms are added to its Lo in the reference's order — the light term of bounce b is
// computed when bounce b is shaded, parked in `cs`, zeroed by an occluded shadow ray (x + 0 is exact; Lo is never -0)
// and added by whichever stage touches the path next, BEFORE that stage's own term — and the running mean sees the
// frames in order, so under the parity arithmetic the pipeline is bit-identical to the megakernel and to the oracle.
//
// Path state lives in HBM, struct-of-arrays, indexed by slot = (frame within the wave) * W*H + py*W + px.
#pragma once
#include "pt_shade.cuh"

namespace pt {

#ifndef PT_WF_CHUNK
#define PT_WF_CHUNK 256           // queue entries a warp claims per atomic
#endif
#ifndef PT_WF_REFILL_MIN
#define PT_WF_REFILL_MIN 4        // wf_trace: idle lanes of a warp before it stops to fetch new rays (all idle: always)
#endif
#ifndef PT_WF_TRACE_CTAS
#define PT_WF_TRACE_CTAS 5        // resident CTAs (256 threads) per SM the traversal kernel is compiled for
#endif
#define PT_WF_SHADOW_BIT 0x80000000u
#define PT_WF_FLAG_SHADOW 1u      // fr.w: a shadow ray flies for this bounce (cs holds its pending light term)
#define PT_WF_FLAG_EXTEND 2u      //       a bounce ray flies

// queue counters: one per 256 bytes (counters 32 bytes apart share an L2 atomic unit)
#define PT_WF_CTR_STRIDE 64       // in uint32
enum { WF_C_RAY_TAIL = 0, WF_C_RAY_HEAD, WF_C_HIT_TAIL, WF_C_HIT_HEAD, WF_C_MISS_TAIL, WF_C_MISS_HEAD, WF_C_KINDS };
#define PT_WF_NUM_CTRS (WF_C_KINDS * (PT_WF_MAX_BOUNCE + 1) + 1)      // + the head kernel's item counter
__device__ __forceinline__ uint32_t* wf_ctr(const WaveBuffers& wf, int kind, int b) {
    return wf.counters + (size_t)(kind * (PT_WF_MAX_BOUNCE + 1) + b) * PT_WF_CTR_STRIDE;
}
__device__ __forceinline__ uint32_t* wf_item_ctr(const WaveBuffers& wf) {
    return wf.counters + (size_t)(WF_C_KINDS * (PT_WF_MAX_BOUNCE + 1)) * PT_WF_CTR_STRIDE;
}

// warp-aggregated append: all 32 lanes call this; lanes with pred write id
__device__ __forceinline__ void wf_push(uint32_t* queue, uint32_t* tail, bool pred, uint32_t id) {
    const uint32_t m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(m) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(tail, (uint32_t)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (pred) queue[base + __popc(m & ((1u << lane) - 1u))] = id;
}
// two entries per lane at most (the shadow ray and the bounce ray of one shaded hit) with ONE atomic per warp
__device__ __forceinline__ void wf_push2(uint32_t* queue, uint32_t* tail, bool p0, uint32_t id0, bool p1, uint32_t id1) {
    const uint32_t m0 = __ballot_sync(0xffffffffu, p0), m1 = __ballot_sync(0xffffffffu, p1);
    if ((m0 | m1) == 0) return;
    const int lane = threadIdx.x & 31;
    const uint32_t n0 = __popc(m0);
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(tail, n0 + (uint32_t)__popc(m1));
    base = __shfl_sync(0xffffffffu, base, 0);
    const uint32_t lt = (1u << lane) - 1u;
    if (p0) queue[base + __popc(m0 & lt)] = id0;
    if (p1) queue[base + n0 + __popc(m1 & lt)] = id1;
}

// a warp claims the next n entries of a work list; returns the first index (may be >= the list's length)
__device__ __forceinline__ uint32_t wf_claim(uint32_t* head, uint32_t n) {
    uint32_t base = 0;
    if ((threadIdx.x & 31) == 0) base = atomicAdd(head, n);
    return __shfl_sync(0xffffffffu, base, 0);
}

__device__ __forceinline__ void wf_flush_rays(uint32_t nrays, unsigned long long* ray_counter) {
    for (int o = 16; o > 0; o >>= 1) nrays += __shfl_xor_sync(0xffffffffu, nrays, o);
    if ((threadIdx.x & 31) == 0 && nrays) atomicAdd(ray_counter, (unsigned long long)nrays);
}

__device__ __forceinline__ float4 f4(V3 v, float w) { return make_float4(v.x, v.y, v.z, w); }
__device__ __forceinline__ V3 xyz(float4 v) { return mk3(v.x, v.y, v.z); }

// the path is finished: colour = Le0 + Lo (P:862-864; Le0 = emissive of the PRIMARY hit's material, whose id rides in lo.w)
template <class Cfg>
__device__ __forceinline__ void wf_finish(const RenderParams& p, uint32_t slot, V3 Lo, int mat0) {
    typename MatOf<Cfg::RAW>::T m0;
    load_material<Cfg::RAW>(p.sc, mat0, m0);
    p.wf.lo[slot] = f4(m0.emissive + Lo, 1.0f);
}

// ---- bounce `b` of a path at a closest hit: the body of shade_sample()'s "closest hit" branch ------------------------
// In: the ray (o, d) that hit, its TraceResult, the path's history / Lo BEFORE this hit and, for b > 0, the f_r, cosw, pdf of
// the ray.  Parks the path (state arrays) and reports which rays fly next; finishes the path when none does.
struct WfNext { bool shadow, extend, to_miss; };
template <class Cfg, int VARIANT>
__device__ __forceinline__ WfNext wf_shade_hit(const RenderParams& p, uint32_t slot, int px, int py, uint32_t frame, const float* __restrict__ sobol_row,
                                               int b, V3 o, V3 d, const TraceResult& tr, V3 history, V3 Lo, V3 f_r, float cosw, float pdf,
                                               uint32_t seed, int mat0, Counter<Cfg::COUNT>& cnt) {
    typedef typename Cfg::M M;
    const DevScene& sc = p.sc;
    const WaveBuffers& wf = p.wf;
    WfNext nx; nx.shadow = nx.extend = nx.to_miss = false;
    Hit hit;
    int mat_id;
    finalize_hit<Cfg::RAW>(sc, o, d, tr, hit, mat_id);
    cnt.shaded();
    typename MatOf<Cfg::RAW>::T mat;
    load_material<Cfg::RAW>(sc, mat_id, mat);
    set_ext(mat, sc, mat_id);
    if (b == 0) {
        mat0 = mat_id;                                                                            // Le0, P:862
    } else {
        Lo = Lo + history * mat.emissive * f_r * cosw / pdf;                                      // P:833-834
        history = history * (f_r * cosw / pdf);                                                   // P:837
    }
    if (b >= p.max_bounce) { wf_finish<Cfg>(p, slot, Lo, mat0); return nx; }
    const V3 V = -hit.viewDir, N = hit.normal;
    o = hit.hitPoint;
    V3 L;
    bool want_shadow = false, next_ok = true;
    if (VARIANT == 2) {
        // light sampling through the hdrCache, P:770-793: the term is computed NOW (nothing in it depends on the shadow ray) and parked
        const float e1 = rnd(seed);
        const float e2 = rnd(seed);
        V3 Ld;
        uint32_t light_idx = 0;
        if (Cfg::LIGHT_TAB) {
            light_idx = (uint32_t)nearest_texel(e2, sc.hdrH) * (uint32_t)sc.hdrW + (uint32_t)nearest_texel(e1, sc.hdrW);
            const float4 r0 = ldg4(sc.light_tab + 2 * (size_t)light_idx);
            Ld = mk3(r0.x, r0.y, r0.z);
        } else {
            Ld = SampleHdr<M>(sc, e1, e2);
        }
        cnt.texel(1u);
        const float ndl = dot(N, Ld);
        want_shadow = ndl > 0.0f;
        float su, sv;
        sobol_pair(sobol_row, (uint32_t)b, frame, su, sv);                                        // P:795-816
        cranley_patterson(px, py, su, sv);
        const float xi_3 = rnd(seed);
        L = SampleBRDF<M>(su, sv, xi_3, V, N, mat);
        cosw = dot(N, L);
        next_ok = cosw > 0.0f;
        V3 fl = mk3(0, 0, 0);
        float pdf_b = 0.0f;
#pragma unroll 1
        for (int k = 0; k < 2; k++) {
            if (k == 0 ? !want_shadow : !next_ok) continue;
            const V3 Lk = k == 0 ? Ld : L;
            const V3 f = BRDF_Evaluate_iso<M>(V, N, Lk, mat);
            const float pf = BRDF_Pdf<M>(V, N, Lk, mat);
            if (k == 0) { fl = f; pdf_b = pf; }
            else { f_r = f; pdf = pf; }
        }
        if (next_ok && pdf <= 0.0f) next_ok = false;
        if (want_shadow) {
            V3 color;
            float pdf_light;
            if (Cfg::LIGHT_TAB) {
                const float4 e = ldg4(sc.light_tab + 2 * (size_t)light_idx + 1);
                color = mk3(e.x, e.y, e.z); pdf_light = e.w;
            } else {
                hdr_color_pdf<M>(sc, Ld, color, pdf_light);
            }
            cnt.texel(2u);
            const float mis_weight = misMixWeight(pdf_light, pdf_b);
            const V3 Cs = (history * mis_weight) * color * fl * ndl / pdf_light;                  // P:781-792
            wf.cs[slot] = f4(Cs, 0.0f);
            wf.sh_d[slot] = f4(Ld, 0.0f);
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
    if (!want_shadow && !next_ok) { wf_finish<Cfg>(p, slot, Lo, mat0); return nx; }
    wf.lo[slot] = f4(Lo, __int_as_float(mat0));
    wf.fr[slot] = f4(f_r, __uint_as_float((want_shadow ? PT_WF_FLAG_SHADOW : 0u) | (next_ok ? PT_WF_FLAG_EXTEND : 0u)));
    wf.ray_o[slot] = f4(o, pdf);                                   // origin of both the shadow ray and the bounce ray
    if (next_ok) {
        wf.ray_d[slot] = f4(L, cosw);
        wf.hist[slot] = f4(history, __uint_as_float(seed));
    }
    nx.shadow = want_shadow; nx.extend = next_ok; nx.to_miss = want_shadow && !next_ok;
    return nx;
}

// ---- head: generate + primary ray + bounce 0 ----------------------------------------------------------------------
// work item = (8x4 pixel tile, frame of the wave), claimed dynamically like the persistent kernel's items
template <class Cfg, int VARIANT>
__global__ void __launch_bounds__(256) wf_head(const __grid_constant__ RenderParams p, uint32_t f0, uint32_t nf) {
    const WaveBuffers& wf = p.wf;
    const int lane = threadIdx.x & 31;
    const TileMap tm = make_tile_map(p.x0, p.y0, p.x1, p.y1, p.band_world, p.band_rank, p.band_diag);
    const uint32_t ntiles = tm.count();
    const uint32_t nitems = ntiles * nf;
    uint32_t nrays = 0;
    Counter<Cfg::COUNT> cnt;
    for (;;) {
        const uint32_t item = wf_claim(wf_item_ctr(wf), 1u);
        if (item >= nitems) break;
        const uint32_t fl = item / ntiles, tile = item - fl * ntiles;
        int tx, strip;
        tm.locate(tile, tx, strip);
        const int px = p.x0 + tx * 8 + (lane & 7);
        const int py = p.y0 + strip * 4 + (lane >> 3);
        const bool valid = px < p.x1 && py < p.y1;
        const uint32_t frame = p.first_frame + (f0 + fl) * p.frame_stride;
        const size_t idx = (size_t)py * p.width + px;
        const uint32_t slot = fl * wf.npix + (uint32_t)idx;
        WfNext nx; nx.shadow = nx.extend = nx.to_miss = false;
        if (valid) {
            uint32_t seed = ((uint32_t)px * 1973u + (uint32_t)py * 9277u + frame * 26699u) | 1u;         // P:73-76
            V3 o, d;
            generate_ray(p, px, py, seed, o, d);
            const TraceResult tr = trace<Cfg>(p.sc, o, d, false, cnt, p.eye_far != 0);
            nrays++;
            if (f0 + fl + 1u == p.n_frames && p.primary_ids) p.primary_ids[idx] = tr.tri;
            if (tr.tri < 0) {
                cnt.texel(1u);
                wf.lo[slot] = f4(sampleHdr<typename Cfg::M>(p.sc, d), 1.0f);                             // P:857-859
            } else {
                nx = wf_shade_hit<Cfg, VARIANT>(p, slot, px, py, frame, p.sobol_tab ? p.sobol_tab + (size_t)(f0 + fl) * 8 : nullptr, 0, o, d, tr,
                                                mk3(1, 1, 1), mk3(0, 0, 0), mk3(0, 0, 0), 0.0f, 1.0f, seed, 0, cnt);
            }
        }
        wf_push2(wf.q_ray, wf_ctr(wf, WF_C_RAY_TAIL, 0), nx.shadow, slot | PT_WF_SHADOW_BIT, nx.extend, slot);
        if (VARIANT == 2) wf_push(wf.q_miss[0], wf_ctr(wf, WF_C_MISS_TAIL, 0), nx.to_miss, slot);
    }
    wf_flush_rays(nrays, p.ray_counter);
    flush_counts(cnt, p.work_counters);
}

// ---- trace: the rays of bounce b -----------------------------------------------------------------------------------
// Per-warp staging of finished paths (shared memory): results leave for Q_hit / Q_miss 32 at a time, one atomic per
// 32 entries instead of one per finishing event.
struct WfStage {
    uint32_t* buf;       // this warp's 64 entries
    uint32_t n;          // warp-uniform fill
};
__device__ __forceinline__ void wf_stage_add(WfStage& s, bool pred, uint32_t id, uint32_t* queue, uint32_t* tail) {
    const uint32_t m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return;
    const int lane = threadIdx.x & 31;
    if (pred) s.buf[s.n + __popc(m & ((1u << lane) - 1u))] = id;
    s.n += __popc(m);
    __syncwarp();
    if (s.n >= 32u) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(tail, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        queue[base + lane] = s.buf[lane];
        const uint32_t rest = s.n - 32u;
        const uint32_t carry = (uint32_t)lane < rest ? s.buf[32 + lane] : 0u;
        __syncwarp();
        if ((uint32_t)lane < rest) s.buf[lane] = carry;
        s.n = rest;
        __syncwarp();
    }
}
__device__ __forceinline__ void wf_stage_flush(WfStage& s, uint32_t* queue, uint32_t* tail) {
    if (s.n == 0) return;
    const int lane = threadIdx.x & 31;
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(tail, s.n);
    base = __shfl_sync(0xffffffffu, base, 0);
    if ((uint32_t)lane < s.n) queue[base + lane] = s.buf[lane];
    s.n = 0;
}

// Reference-topology configurations (PT_MODE_PARITY_WIDE): one ray per lane per round, no refill (a correctness mode).
template <class Cfg>
__device__ __forceinline__ void wf_trace_simple(const RenderParams& p, int b, uint32_t count, WfStage& sh, WfStage& sm, uint32_t& nrays,
                                                Counter<Cfg::COUNT>& cnt) {
    const WaveBuffers& wf = p.wf;
    const int lane = threadIdx.x & 31;
    for (;;) {
        const uint32_t base = wf_claim(wf_ctr(wf, WF_C_RAY_HEAD, b), (uint32_t)PT_WF_CHUNK);
        if (base >= count) break;
#pragma unroll 1
        for (uint32_t k = 0; k < PT_WF_CHUNK; k += 32) {
            const uint32_t i = base + k + lane;
            bool hit = false, miss = false;
            uint32_t slot = 0;
            if (i < count) {
                const uint32_t e = wf.q_ray[i];
                slot = e & ~PT_WF_SHADOW_BIT;
                const bool shadow = (e & PT_WF_SHADOW_BIT) != 0;
                const V3 o = xyz(wf.ray_o[slot]), d = xyz(shadow ? wf.sh_d[slot] : wf.ray_d[slot]);
                const TraceResult tr = trace<Cfg>(p.sc, o, d, shadow, cnt);
                nrays++;
                if (shadow) {
                    if (tr.tri >= 0) wf.cs[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
                } else if (tr.tri >= 0) {
                    wf.hit[slot] = make_int2(tr.tri | (tr.inside ? (int)0x80000000 : 0), __float_as_int(tr.t));
                    hit = true;
                } else {
                    miss = true;
                }
            }
            wf_stage_add(sh, hit, slot, wf.q_hit, wf_ctr(wf, WF_C_HIT_TAIL, b));
            wf_stage_add(sm, miss, slot, wf.q_miss[b & 1], wf_ctr(wf, WF_C_MISS_TAIL, b));
        }
    }
}

// Production topology: trace_fast's while-while loop (pt_trace.cuh — same tests, same order, same results per ray) with the
// ray state per LANE and a refill point at the top of every round: lanes whose ray is finished take the next rays of the
// warp's chunk of the queue.  A lane without a ray carries ref = PT_REF_NONE through both loops.
template <class Cfg>
__device__ __forceinline__ void wf_trace_refill(const RenderParams& p, int b, uint32_t count, WfStage& sh, WfStage& sm, uint32_t* smem_stack,
                                                uint32_t& nrays, Counter<Cfg::COUNT>& cnt) {
    const WaveBuffers& wf = p.wf;
    const DevScene& sc = p.sc;
    const int lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t* const head = wf_ctr(wf, WF_C_RAY_HEAD, b);
    uint32_t chunk_next = 0, chunk_end = 0;            // warp-uniform: the part of the queue this warp still owns
    bool more = true;                                  // the queue may hold unclaimed chunks
    // per-lane ray
    bool have = false;
    uint32_t entry = 0;
    Ray r;
    r.o = r.d = r.inv = mk3(0, 0, 0);
    CullRay cr;
    cr.inv = cr.noi = mk3(0, 0, 0);
    TraceResult res; res.tri = -1; res.t = PT_INF; res.inside = false;
    float cull = PT_INF;
    bool tie = false;
    uint32_t steps = 0;
    TravStack<PT_STACK> stk;
    stk.init(smem_stack);
    uint32_t ref = PT_REF_NONE;
    for (;;) {
        // ---- refill ----
        {
            uint32_t idle = __ballot_sync(0xffffffffu, !have);
            const bool can_fetch = more || chunk_next < chunk_end;
            if (idle == 0xffffffffu || (can_fetch && (uint32_t)__popc(idle) >= (uint32_t)PT_WF_REFILL_MIN)) {
                for (;;) {
                    if (chunk_next >= chunk_end) {
                        if (!more) break;
                        const uint32_t base = wf_claim(head, (uint32_t)PT_WF_CHUNK);
                        if (base >= count) { more = false; break; }
                        chunk_next = base;
                        chunk_end = min(base + (uint32_t)PT_WF_CHUNK, count);
                    }
                    const uint32_t n_idle = (uint32_t)__popc(idle);
                    const uint32_t take = min(n_idle, chunk_end - chunk_next);
                    const uint32_t rank = (uint32_t)__popc(idle & lt);
                    if (!have && rank < take) {
                        entry = wf.q_ray[chunk_next + rank];
                        const uint32_t slot = entry & ~PT_WF_SHADOW_BIT;
                        const float4 o4 = wf.ray_o[slot];
                        const float4 d4 = (entry & PT_WF_SHADOW_BIT) ? wf.sh_d[slot] : wf.ray_d[slot];
                        r.o = xyz(o4); r.d = xyz(d4);
                        cr = make_cull_ray(r.o, mk3(rcp_approx(r.d.x), rcp_approx(r.d.y), rcp_approx(r.d.z)));
                        res.tri = -1; res.t = PT_INF; res.inside = false;
                        cull = PT_INF; tie = false; steps = 0;
                        stk.clear();
                        ref = sc.root_fast;
                        have = true;
                        nrays++;
                    }
                    chunk_next += take;
                    if (take == n_idle) break;
                    idle = __ballot_sync(0xffffffffu, !have);
                }
                if (__ballot_sync(0xffffffffu, have) == 0u) break;      // every lane idle and nothing left to fetch
            }
        }
        const bool anyhit = (entry & PT_WF_SHADOW_BIT) != 0;
        // ---- one round of trace_fast: descend to a leaf, test it, pop ----
        while (ref_count(ref) == 0) {
            const float4* w = sc.wnodes_fast + 4 * (size_t)ref;
            const float4 q0 = ldg4(w), q1 = ldg4(w + 1), q2 = ldg4(w + 2), q3 = ldg4(w + 3);
            cnt.inner();
            if (Cfg::COUNT) steps++;
            float e1, e2;
            const bool g1 = slab_cull_fused(cr, mk3(q0.x, q0.y, q0.z), mk3(q0.w, q1.x, q1.y), cull, e1);
            const bool g2 = slab_cull_fused(cr, mk3(q1.z, q1.w, q2.x), mk3(q2.y, q2.z, q2.w), cull, e2);
            const uint32_t lref = __float_as_uint(q3.x), rref = __float_as_uint(q3.y);
            if (g1 && g2) {
                const bool left_first = e1 <= e2;
                if (stk.full()) { *sc.error_flag = PT_ERR_STACK; stk.clear(); ref = PT_REF_NONE; break; }
                stk.push(left_first ? rref : lref);
                ref = left_first ? lref : rref;
            } else if (g1 || g2) {
                ref = g1 ? lref : rref;
            } else {
                ref = PT_REF_NONE;
            }
        }
        if (ref != PT_REF_NONE) {
            const uint32_t first = ref_index(ref), end = first + ref_count(ref);
            for (uint32_t i = first; i < end; i++) {
                const float4* g = sc.trigeo_fast + 3 * (size_t)i;
                const float4 a = ldg4(g), bq = ldg4(g + 1), c = ldg4(g + 2);
                cnt.tri();
                if (Cfg::COUNT) steps++;
                float tt; bool ins;
                if (hit_triangle(r, mk3(a.x, a.y, a.z), mk3(bq.x, bq.y, bq.z), mk3(c.x, c.y, c.z), mk3(a.w, bq.w, c.w), res.t, tt, ins,
                                 !anyhit && res.tri >= 0)) {
                    cnt.cand();
                    const int2 meta = __ldg(sc.trimeta + i);
                    if (meta.y != 1) {
                        const float* nd = sc.nodes_raw + 12 * (size_t)meta.y;
                        float e;
                        r.inv = mk3(s_rcp(r.d.x), s_rcp(r.d.y), s_rcp(r.d.z));
                        if (!(hit_aabb(r, raw3(nd + 6), raw3(nd + 9), e) > 0.0f)) continue;
                    }
                    if (!anyhit && tt == res.t) { tie = true; continue; }
                    res.t = tt; res.tri = meta.x; res.inside = ins;
                    if (anyhit) { stk.clear(); break; }
                    cull = s_mul(tt, PT_CULL_SLACK);
                }
            }
        }
        // ---- pop, or finish this lane's ray ----
        bool fin = false;
        if (have) {
            if (stk.empty()) { fin = true; ref = PT_REF_NONE; }
            else ref = stk.pop();
        }
        if (__ballot_sync(0xffffffffu, fin) != 0u) {
            const uint32_t slot = entry & ~PT_WF_SHADOW_BIT;
            bool hit = false, miss = false;
            if (fin) {
                if (Cfg::COUNT) atomicAdd(sc.ray_histogram + (steps ? 32 - __clz(steps) : 0), 1ull);
                if (!anyhit && tie) { cnt.tie(); res = trace_wide_fallback(sc, r.o, r.d); }     // exact tie: let the reference order decide
                if (anyhit) {
                    if (res.tri >= 0) wf.cs[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
                } else if (res.tri >= 0) {
                    wf.hit[slot] = make_int2(res.tri | (res.inside ? (int)0x80000000 : 0), __float_as_int(res.t));
                    hit = true;
                } else {
                    miss = true;
                }
                have = false;
            }
            wf_stage_add(sh, hit, slot, wf.q_hit, wf_ctr(wf, WF_C_HIT_TAIL, b));
            wf_stage_add(sm, miss, slot, wf.q_miss[b & 1], wf_ctr(wf, WF_C_MISS_TAIL, b));
        }
    }
}

template <class Cfg>
__global__ void __launch_bounds__(256, PT_WF_TRACE_CTAS) wf_trace(const __grid_constant__ RenderParams p, int b) {
    const WaveBuffers& wf = p.wf;
    __shared__ uint32_t s_stage[2][8][64];
#if PT_SMEM_STACK > 0
    __shared__ uint32_t s_stack[PT_SMEM_STACK * PT_STACK_STRIDE];
    uint32_t* const smem_stack = s_stack;
#else
    uint32_t* const smem_stack = nullptr;
#endif
    const int warp = threadIdx.x >> 5;
    WfStage sh, sm;
    sh.buf = s_stage[0][warp]; sh.n = 0;
    sm.buf = s_stage[1][warp]; sm.n = 0;
    const uint32_t count = *wf_ctr(wf, WF_C_RAY_TAIL, b);
    uint32_t nrays = 0;
    Counter<Cfg::COUNT> cnt;
    if (Cfg::FAST_TREE) wf_trace_refill<Cfg>(p, b, count, sh, sm, smem_stack, nrays, cnt);
    else wf_trace_simple<Cfg>(p, b, count, sh, sm, nrays, cnt);
    wf_stage_flush(sh, wf.q_hit, wf_ctr(wf, WF_C_HIT_TAIL, b));
    wf_stage_flush(sm, wf.q_miss[b & 1], wf_ctr(wf, WF_C_MISS_TAIL, b));
    wf_flush_rays(nrays, p.ray_counter);
    flush_counts(cnt, p.work_counters);
}

// ---- shade: what trace(b-1) resolved ---------------------------------------------------------------------------------
// phase 1: Q_hit[b-1] — the bounce ray hit: emissive term, history, then bounce b;  phase 2: Q_miss[b-1] — it escaped (or
// never flew): environment term, path done.  Both lists are compact: full warps in either phase.
template <class Cfg, int VARIANT>
__global__ void __launch_bounds__(256) wf_shade(const __grid_constant__ RenderParams p, uint32_t f0, int b) {
    typedef typename Cfg::M M;
    const WaveBuffers& wf = p.wf;
    const DevScene& sc = p.sc;
    const int lane = threadIdx.x & 31;
    Counter<Cfg::COUNT> cnt;
    {
        const uint32_t count = *wf_ctr(wf, WF_C_HIT_TAIL, b - 1);
        for (;;) {
            const uint32_t base = wf_claim(wf_ctr(wf, WF_C_HIT_HEAD, b - 1), 64u);
            if (base >= count) break;
#pragma unroll 1
            for (uint32_t k = 0; k < 64u; k += 32) {
                const uint32_t i = base + k + lane;
                WfNext nx; nx.shadow = nx.extend = nx.to_miss = false;
                uint32_t slot = 0;
                if (i < count) {
                    slot = wf.q_hit[i];
                    const float4 ro = wf.ray_o[slot], rd = wf.ray_d[slot], hs = wf.hist[slot], fr = wf.fr[slot], lo4 = wf.lo[slot];
                    const int2 hr = wf.hit[slot];
                    TraceResult tr;
                    tr.tri = hr.x & 0x7fffffff; tr.inside = hr.x < 0; tr.t = __int_as_float(hr.y);
                    V3 Lo = xyz(lo4);
                    if (__float_as_uint(fr.w) & PT_WF_FLAG_SHADOW) Lo = Lo + xyz(wf.cs[slot]);            // the light term of bounce b-1, in order
                    const uint32_t fl = slot / wf.npix, idx = slot - fl * wf.npix;
                    const int px = (int)(idx % (uint32_t)p.width), py = (int)(idx / (uint32_t)p.width);
                    const uint32_t frame = p.first_frame + (f0 + fl) * p.frame_stride;
                    nx = wf_shade_hit<Cfg, VARIANT>(p, slot, px, py, frame, p.sobol_tab ? p.sobol_tab + (size_t)(f0 + fl) * 8 : nullptr, b, xyz(ro), xyz(rd), tr,
                                                    xyz(hs), Lo, xyz(fr), rd.w, ro.w, __float_as_uint(hs.w), __float_as_int(lo4.w), cnt);
                }
                wf_push2(wf.q_ray, wf_ctr(wf, WF_C_RAY_TAIL, b), nx.shadow, slot | PT_WF_SHADOW_BIT, nx.extend, slot);
                if (VARIANT == 2) wf_push(wf.q_miss[b & 1], wf_ctr(wf, WF_C_MISS_TAIL, b), nx.to_miss, slot);
            }
        }
    }
    {
        const uint32_t count = *wf_ctr(wf, WF_C_MISS_TAIL, b - 1);
        for (;;) {
            const uint32_t base = wf_claim(wf_ctr(wf, WF_C_MISS_HEAD, b - 1), 64u);
            if (base >= count) break;
#pragma unroll 1
            for (uint32_t k = 0; k < 64u; k += 32) {
                const uint32_t i = base + k + lane;
                if (i >= count) continue;
                const uint32_t slot = wf.q_miss[(b - 1) & 1][i];
                const float4 fr = wf.fr[slot], lo4 = wf.lo[slot];
                const uint32_t flags = __float_as_uint(fr.w);
                V3 Lo = xyz(lo4);
                if (flags & PT_WF_FLAG_SHADOW) Lo = Lo + xyz(wf.cs[slot]);
                if (flags & PT_WF_FLAG_EXTEND) {
                    const float4 ro = wf.ray_o[slot], rd = wf.ray_d[slot], hs = wf.hist[slot];
                    const V3 d = xyz(rd), history = xyz(hs), f_r = xyz(fr);
                    const float cosw = rd.w, pdf = ro.w;
                    float u, w;
                    spherical_uv(d, u, w);
                    const float4 c = tex_nearest(sc.hdr4, sc.hdrW, sc.hdrH, u, w);
                    cnt.texel(VARIANT == 2 ? 2u : 1u);
                    if (VARIANT == 2) {                                                                   // P:819-830
                        const V3 color = mk3(c.x, c.y, c.z);
                        const float pdf_light = hdr_pdf<M>(sc, u, w);
                        const float mis_weight = misMixWeight(pdf, pdf_light);
                        Lo = Lo + (history * mis_weight) * color * f_r * cosw / pdf;
                    } else {                                                                              // O:349-353, D:465-469
                        const V3 sky = mk3(fminf(c.x, 10.0f), fminf(c.y, 10.0f), fminf(c.z, 10.0f));
                        Lo = Lo + history * sky * f_r * cosw / pdf;
                    }
                }
                wf_finish<Cfg>(p, slot, Lo, __float_as_int(lo4.w));
            }
        }
    }
    flush_counts(cnt, p.work_counters);
}

// ---- accumulate: the wave's nf colours of every owned pixel, folded in frame order ---------------------------------------
template <class Cfg>
__global__ void __launch_bounds__(256) wf_accumulate(const __grid_constant__ RenderParams p, uint32_t f0, uint32_t nf) {
    const WaveBuffers& wf = p.wf;
    const int lane = threadIdx.x & 31;
    const TileMap tm = make_tile_map(p.x0, p.y0, p.x1, p.y1, p.band_world, p.band_rank, p.band_diag);
    const uint32_t ntiles = tm.count();
    for (uint32_t tile = blockIdx.x * 8u + (threadIdx.x >> 5); tile < ntiles; tile += gridDim.x * 8u) {
        int tx, strip;
        tm.locate(tile, tx, strip);
        const int px = p.x0 + tx * 8 + (lane & 7);
        const int py = p.y0 + strip * 4 + (lane >> 3);
        if (px >= p.x1 || py >= p.y1) continue;
        const size_t idx = (size_t)py * p.width + px;
        float4 acc = p.accum[idx];
        for (uint32_t fl = 0; fl < nf; fl++) {
            const float4 c = wf.lo[(size_t)fl * wf.npix + idx];
            if (p.accumulate_sum) {
                acc.x = s_add(acc.x, c.x); acc.y = s_add(acc.y, c.y); acc.z = s_add(acc.z, c.z);
            } else {                                      // mix(lastColor, color, 1/(frameCounter+1)), P:868-869
                const float a = s_div(1.0f, (float)(p.first_frame + (f0 + fl) * p.frame_stride + 1u));
                const float bq = s_sub(1.0f, a);
                acc.x = s_add(s_mul(acc.x, bq), s_mul(c.x, a));
                acc.y = s_add(s_mul(acc.y, bq), s_mul(c.y, a));
                acc.z = s_add(s_mul(acc.z, bq), s_mul(c.z, a));
            }
        }
        acc.w = 1.0f;
        p.accum[idx] = acc;
    }
}

// ---- host side: wave after wave on the context's stream ---------------------------------------------------------------
// launches per wave (the counter memset is not a kernel): head + max_bounce x (trace + shade) + accumulate
__host__ inline int wavefront_launches_per_wave(int max_bounce) { return 2 + 2 * max_bounce; }
__host__ inline uint32_t wavefront_frames_per_wave(const WaveBuffers& wf) { return (uint32_t)(wf.cap / wf.npix > 0 ? wf.cap / wf.npix : 1); }

template <class Cfg, int VARIANT>
cudaError_t launch_wavefront_variant(const RenderParams& p, int num_sms, cudaStream_t stream) {
    const WaveBuffers& wf = p.wf;
    const int mb = p.max_bounce;
    const uint32_t F = wavefront_frames_per_wave(wf);
    if (VARIANT == 2 && p.sobol_tab)
        sobol_table_kernel<<<(p.n_frames * 8 + 255) / 256, 256, 0, stream>>>(const_cast<float*>(p.sobol_tab), p.first_frame, p.n_frames, p.frame_stride);
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wf_trace<Cfg>, 256, 0);
    if (e != cudaSuccess) return e;
    const int grid_trace = num_sms * (per_sm > 0 ? per_sm : 1), grid = num_sms * 5;
    for (uint32_t f0 = 0; f0 < p.n_frames; f0 += F) {
        const uint32_t nf = p.n_frames - f0 < F ? p.n_frames - f0 : F;
        e = cudaMemsetAsync(wf.counters, 0, (size_t)PT_WF_NUM_CTRS * PT_WF_CTR_STRIDE * sizeof(uint32_t), stream);
        if (e != cudaSuccess) return e;
        wf_head<Cfg, VARIANT><<<grid, 256, 0, stream>>>(p, f0, nf);
        for (int b = 0; b < mb; b++) {
            wf_trace<Cfg><<<grid_trace, 256, 0, stream>>>(p, b);
            wf_shade<Cfg, VARIANT><<<grid, 256, 0, stream>>>(p, f0, b + 1);
        }
        wf_accumulate<Cfg><<<grid, 256, 0, stream>>>(p, f0, nf);
    }
    return cudaGetLastError();
}

template <class Cfg>
cudaError_t launch_wavefront(const RenderParams& p, int variant, int num_sms, cudaStream_t stream) {
    if (variant == 0) return launch_wavefront_variant<Cfg, 0>(p, num_sms, stream);
    if (variant == 1) return launch_wavefront_variant<Cfg, 1>(p, num_sms, stream);
    return launch_wavefront_variant<Cfg, 2>(p, num_sms, stream);
}

}  // namespace pt
