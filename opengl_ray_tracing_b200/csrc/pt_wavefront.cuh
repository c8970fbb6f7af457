// pt_wavefront.cuh — the wavefront pipeline: the same per-sample arithmetic as shade_sample() (pt_shade.cuh), cut at
// every secondary hitBVH call into kernels that hand rays and paths to each other through queues in HBM.
//
// Round-2 design (the round-1 form processed ONE frame per wave with one ray per lane and no refill, and lost 9x to the
// persistent kernel; profiles/r01_wavefront.md):
//   * a WAVE is F frames of the whole region at once.  Only the paths that survive their primary ray get a path slot
//     (compact ids from one atomic counter), so F adapts to the scene: the host reads the survivor count of every wave and
//     sizes the next one to fill the path capacity (64 M slots by default) — 32 frames per wave where every pixel hits
//     geometry, hundreds where most of the image is sky.  Every launch then carries tens of millions of rays;
//   * secondary rays — shadow and bounce rays of one bounce together — go through ONE queue into a TRAVERSAL-ONLY
//     persistent kernel (wf_trace) that holds nothing but ray state per lane: lanes whose ray is finished write its result
//     and take the next rays of the warp's chunk of the queue INSIDE the while-while loop (__ballot_sync / __popc ranks,
//     one atomicAdd per chunk of rays), as soon as PT_WF_REFILL_MIN lanes are free, so the node loop keeps running with
//     full warps instead of the 7-10 of 32 lanes the megakernel averages on mesh rays (north star items 1 and 4);
//   * primary rays are coherent (8x4 pixel tiles) and need no queue: wf_head generates, traces and shades them like the
//     persistent kernel does, then parks the path;
//   * shading runs over COMPACTED queues (wf_classify splits the traced rays into paths that hit / escaped; wf_shade runs
//     the ~1.5 k instructions of Disney BRDF code with all 32 lanes of a warp on the same branch).
//
// One wave, bounce b = 0 .. max_bounce-1 (launches: 2 + 3 max_bounce):
//   wf_head         ray generation P:846-850 + primary hitBVH; miss -> sampleHdr, sample done; hit -> bounce 0 is shaded:
//                   light sample -> shadow ray + its MIS-weighted term Cs (pending), BRDF sample -> bounce ray, f_r, pdf
//   wf_trace(b)     every ray of bounce b: result (triangle, t) per queue entry; any-hit for shadow rays
//   wf_classify(b)  occluded shadow ray -> Cs := 0; bounce ray -> hit record + Q_hit, or Q_miss
//   wf_shade(b+1)   Q_hit: Lo += Cs, emissive term, history (P:833-837), then bounce b+1 as above;
//                   Q_miss: Lo += Cs, environment term (MIS-weighted in variant 2, P:819-830), sample done
//   wf_accumulate   per pixel, the wave's F colours folded into the framebuffer IN FRAME ORDER (P:868-869)
//
// Exactness: a path's radiance terms are added to its Lo in the reference's order — the light term of bounce b is
// computed when bounce b is shaded, parked in `cs`, zeroed by an occluded shadow ray (x + 0 is exact; Lo is never -0)
// and added by whichever stage touches the path next, BEFORE that stage's own term — and the running mean sees the
// frames in order, so under the parity arithmetic the pipeline is bit-identical to the megakernel and to the oracle.
#pragma once
#include <algorithm>

#include "pt_shade.cuh"

namespace pt {

#ifndef PT_WF_CHUNK
#define PT_WF_CHUNK 256           // queue entries a warp claims per atomic
#endif
#ifndef PT_WF_REFILL_MIN
#define PT_WF_REFILL_MIN 16       // wf_trace: free lanes of a warp before it stops to store results and fetch new rays (none busy: always);
                                  // measured 2 / 4 / 8 / 16 on disney-1M: 7.9 / 8.3 / 8.5 / 8.7 Grays/s
#endif
#ifndef PT_WF_TRACE_CTAS
#define PT_WF_TRACE_CTAS 3        // resident CTAs (256 threads) per SM the traversal kernel is compiled for: 3 = 80 registers, no spills;
                                  // 4 (64 registers, 150 B of spills) -4 %, 5 (48 registers) -10 % on disney-1M
#endif
#ifndef PT_WF_SMEM_STACK
#define PT_WF_SMEM_STACK 0        // wf_trace: traversal-stack entries per thread kept in shared memory (TravStack, pt_trace.cuh)
#endif
#ifndef PT_WF_NODE_BIAS
#define PT_WF_NODE_BIAS 0         // wf_trace's vote: a node step runs when (lanes at nodes) + bias >= (lanes at triangles)
#endif
#ifndef PT_WF_HEAD_CTAS
#define PT_WF_HEAD_CTAS 4         // same for wf_head / wf_shade (64 registers)
#endif
#define PT_WF_SHADOW_BIT 0x80000000u
#define PT_WF_FLAG_SHADOW 1u      // fr.w: a shadow ray flies for this bounce (cs holds its pending light term)
#define PT_WF_FLAG_EXTEND 2u      //       a bounce ray flies

// queue counters: one per 256 bytes (counters 32 bytes apart share an L2 atomic unit)
#define PT_WF_CTR_STRIDE 64       // in uint32
enum { WF_C_RAY_TAIL = 0, WF_C_RAY_HEAD, WF_C_HIT_TAIL, WF_C_HIT_HEAD, WF_C_MISS_TAIL, WF_C_MISS_HEAD, WF_C_KINDS };
#define PT_WF_NUM_CTRS (WF_C_KINDS * (PT_WF_MAX_BOUNCE + 1) + 4)      // + the head kernel's item counter, path allocator, overflow flag, wave ray count
__device__ __forceinline__ uint32_t* wf_ctr(const WaveBuffers& wf, int kind, int b) {
    return wf.counters + (size_t)(kind * (PT_WF_MAX_BOUNCE + 1) + b) * PT_WF_CTR_STRIDE;
}
__host__ __device__ __forceinline__ uint32_t* wf_misc_ctr(const WaveBuffers& wf, int k) {      // 0 item head, 1 paths allocated, 2 overflow, 3 rays of wf_head
    return wf.counters + (size_t)(WF_C_KINDS * (PT_WF_MAX_BOUNCE + 1) + k) * PT_WF_CTR_STRIDE;
}
__device__ __forceinline__ uint32_t* wf_item_ctr(const WaveBuffers& wf) { return wf_misc_ctr(wf, 0); }

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
__device__ __forceinline__ void wf_finish(const RenderParams& p, uint32_t islot, V3 Lo, int mat0) {
    typename MatOf<Cfg::RAW>::T m0;
    load_material<Cfg::RAW>(p.sc, mat0, m0);
    p.wf.img[islot] = f4(m0.emissive + Lo, 1.0f);
}

// ---- bounce `b` of a path at a closest hit: the body of shade_sample()'s "closest hit" branch ------------------------
// In: the ray (o, d) that hit, its TraceResult, the path's history / Lo BEFORE this hit and, for b > 0, the f_r, cosw, pdf of
// the ray.  Parks the path (state arrays, index `slot` = path id) and reports which rays fly next; finishes the sample
// (image slot `islot`) when none does.
struct WfNext { bool shadow, extend, to_miss; };
template <class Cfg, int VARIANT>
__device__ __forceinline__ WfNext wf_shade_hit(const RenderParams& p, uint32_t slot, uint32_t islot, int px, int py, uint32_t frame, const float* __restrict__ sobol_row,
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
    if (b >= p.max_bounce) { wf_finish<Cfg>(p, islot, Lo, mat0); return nx; }
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
            Ld = SampleHdr<M, Cfg::ENV_CDF>(sc, e1, e2);
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
                hdr_color_pdf<M, Cfg::ENV_CDF>(sc, Ld, color, pdf_light);
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
        L = toNormalHemisphere(mk3(r * M::cos_dir(phi), r * M::sin_dir(phi), z), N);
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
    if (!want_shadow && !next_ok) { wf_finish<Cfg>(p, islot, Lo, mat0); return nx; }
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
// work item = (8x4 pixel tile, frame of the wave), claimed dynamically like the persistent kernel's items.  Paths that hit
// get a path id from one warp-aggregated atomic; running out of ids sets the overflow flag (the host repeats the wave
// with fewer frames).
template <class Cfg, int VARIANT>
__global__ void __launch_bounds__(256, PT_WF_HEAD_CTAS) wf_head(const __grid_constant__ RenderParams p, uint32_t f0, uint32_t nf) {
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
        const uint32_t islot = (uint32_t)PT_IDX(p.sc, (size_t)fl * wf.npix + idx, valid ? wf.icap : ~(size_t)0);
        uint32_t seed = ((uint32_t)px * 1973u + (uint32_t)py * 9277u + frame * 26699u) | 1u;             // P:73-76
        V3 o = mk3(0, 0, 0), d = mk3(0, 0, 1);
        TraceResult tr; tr.tri = -1; tr.t = PT_INF; tr.inside = false;
        if (valid) {
            generate_ray(p, px, py, seed, o, d);
            tr = trace<Cfg>(p.sc, o, d, false, cnt, p.eye_far != 0);
            nrays++;
            if (f0 + fl + 1u == p.n_frames && p.primary_ids) p.primary_ids[idx] = tr.tri;
            if (tr.tri < 0) {
                cnt.texel(1u);
                wf.img[islot] = f4(sampleHdr<typename Cfg::M>(p.sc, d), 1.0f);                           // P:857-859
            }
        }
        // path ids for the lanes whose primary ray hit
        bool alive = valid && tr.tri >= 0;
        const uint32_t am = __ballot_sync(0xffffffffu, alive);
        uint32_t slot = 0;
        if (am) {
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(wf_misc_ctr(wf, 1), (uint32_t)__popc(am));
            base = __shfl_sync(0xffffffffu, base, 0);
            slot = base + (uint32_t)__popc(am & ((1u << lane) - 1u));
            if (alive && slot >= wf.pcap) {                                  // out of path slots: this wave is repeated with fewer frames
                *wf_misc_ctr(wf, 2) = 1u;
                alive = false;
            }
        }
        WfNext nx; nx.shadow = nx.extend = nx.to_miss = false;
        if (alive) {
            wf.pix[slot] = islot;
            nx = wf_shade_hit<Cfg, VARIANT>(p, slot, islot, px, py, frame, p.sobol_tab ? p.sobol_tab + (size_t)(f0 + fl) * 8 : nullptr, 0, o, d, tr,
                                            mk3(1, 1, 1), mk3(0, 0, 0), mk3(0, 0, 0), 0.0f, 1.0f, seed, 0, cnt);
        }
        wf_push2(wf.q_ray, wf_ctr(wf, WF_C_RAY_TAIL, 0), nx.shadow, slot | PT_WF_SHADOW_BIT, nx.extend, slot);
        if (VARIANT == 2) wf_push(wf.q_miss[0], wf_ctr(wf, WF_C_MISS_TAIL, 0), nx.to_miss, slot);
    }
    // primary rays of this wave: added to the context's ray counter by wf_accumulate, i.e. only once the wave is accepted
    for (int o2 = 16; o2 > 0; o2 >>= 1) nrays += __shfl_xor_sync(0xffffffffu, nrays, o2);
    if (lane == 0 && nrays) atomicAdd(wf_misc_ctr(wf, 3), nrays);
    flush_counts(cnt, p.work_counters);
}

// ---- trace: the rays of bounce b -----------------------------------------------------------------------------------
// Output: res[i] for queue entry i — nothing else leaves this kernel (wf_classify turns results into queues), so a finished
// ray costs its lane one 8-byte store.
__device__ __forceinline__ int2 wf_pack_result(const TraceResult& r) {
    return make_int2(r.tri >= 0 ? (r.tri | (r.inside ? (int)0x80000000 : 0)) : -1, __float_as_int(r.t));
}

// Reference-topology configurations (PT_MODE_PARITY_WIDE): one ray per lane per round, no refill (a correctness mode).
template <class Cfg>
__device__ __forceinline__ void wf_trace_simple(const RenderParams& p, int b, uint32_t count, uint32_t& nrays, Counter<Cfg::COUNT>& cnt) {
    const WaveBuffers& wf = p.wf;
    const int lane = threadIdx.x & 31;
    for (;;) {
        const uint32_t base = wf_claim(wf_ctr(wf, WF_C_RAY_HEAD, b), (uint32_t)PT_WF_CHUNK);
        if (base >= count) break;
#pragma unroll 1
        for (uint32_t k = 0; k < PT_WF_CHUNK; k += 32) {
            const uint32_t i = base + k + lane;
            if (i >= count) continue;
            const uint32_t e = wf.q_ray[i], slot = e & ~PT_WF_SHADOW_BIT;
            const bool shadow = (e & PT_WF_SHADOW_BIT) != 0;
            const V3 o = xyz(wf.ray_o[slot]), d = xyz(shadow ? wf.sh_d[slot] : wf.ray_d[slot]);
            const TraceResult tr = trace<Cfg>(p.sc, o, d, shadow, cnt);
            nrays++;
            wf.res[i] = wf_pack_result(tr);
        }
    }
}

// Production topology: trace_fast's tests in trace_fast's order per ray (pt_trace.cuh: same result per ray, bit for bit), but
// scheduled for a warp whose 32 lanes hold 32 UNRELATED rays:
//   * every lane is in one of two modes — at an inner NODE (one 64-byte fetch + two slab tests), or at a TRIangle of a leaf
//     (one 48-byte fetch + hitTriangle).  Each turn the warp VOTES (__ballot_sync / __popc) and runs the step the majority
//     wants; the minority idles for that turn only.  (A while-while loop — descend until every lane has a leaf, then test —
//     makes the whole warp wait for its freshest ray's ~25-level descent from the root: measured 5.5 of 32 lanes per
//     instruction with refill, profiles/r02_wavefront.md.)
//   * lanes whose ray is finished park their result; once PT_WF_REFILL_MIN lanes are free (or no lane is busy) they store the
//     results and all free lanes take the next rays of the warp's chunk of the queue together — one pass through the ~70
//     instructions of result store + ray set-up per PT_WF_REFILL_MIN rays.
#define PT_WF_MODE_NONE 0
#define PT_WF_MODE_NODE 1
#define PT_WF_MODE_TRI 2
#define PT_WF_MODE_DONE 3
template <class Cfg>
__device__ __forceinline__ void wf_trace_refill(const RenderParams& p, int b, uint32_t count, uint32_t* smem_stack, uint32_t& nrays,
                                                Counter<Cfg::COUNT>& cnt) {
    const WaveBuffers& wf = p.wf;
    const DevScene& sc = p.sc;
    const int lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t* const head = wf_ctr(wf, WF_C_RAY_HEAD, b);
    // chunk = rays a warp claims per atomic: PT_WF_CHUNK for long queues, less (>= 32) when the queue holds fewer than ~4
    // chunks per resident warp, so that a short queue (late bounces) still spreads over every SM
    const uint32_t chunk = max(32u, min((uint32_t)PT_WF_CHUNK, (count / (gridDim.x * 8u * 4u)) & ~31u));
    uint32_t chunk_next = 0, chunk_end = 0;            // warp-uniform: the part of the queue this warp still owns
    bool more = true;                                  // the queue may hold unclaimed chunks
    int mode = PT_WF_MODE_NONE;
    uint32_t qi = 0;                                   // queue position of this lane's ray
    bool anyhit = false;
    Ray r;
    r.o = r.d = r.inv = mk3(0, 0, 0);
    CullRay cr;
    cr.inv = cr.noi = mk3(0, 0, 0);
    TraceResult res; res.tri = -1; res.t = PT_INF; res.inside = false;
    float cull = PT_INF;
    bool tie = false;
    uint32_t steps = 0;
    TravStack<PT_STACK, PT_WF_SMEM_STACK> stk;
    stk.init(smem_stack);
    uint32_t ref = PT_REF_NONE;                        // NODE: the inner node to fetch
    uint32_t ti = 0, tend = 0;                         // TRI: the next triangle of the leaf, and the leaf's end
    for (;;) {
        const uint32_t m_node = __ballot_sync(0xffffffffu, mode == PT_WF_MODE_NODE), m_tri = __ballot_sync(0xffffffffu, mode == PT_WF_MODE_TRI);
        const uint32_t busy = m_node | m_tri;
        // ---- store finished results, refill ----
        {
            const uint32_t nfree = 32u - (uint32_t)__popc(busy);
            const bool can_fetch = more || chunk_next < chunk_end;
            if (busy == 0u || (nfree >= (uint32_t)PT_WF_REFILL_MIN && can_fetch)) {      // (queue exhausted: results wait for the warp's last ray)
                if (mode == PT_WF_MODE_DONE) {
                    if (Cfg::COUNT) atomicAdd(sc.ray_histogram + (steps ? 32 - __clz(steps) : 0), 1ull);
                    if (!anyhit && tie) { cnt.tie(); res = trace_wide_fallback(sc, r.o, r.d); }   // exact tie: let the reference order decide
                    wf.res[qi] = wf_pack_result(res);
                    mode = PT_WF_MODE_NONE;
                }
                uint32_t idle = ~busy;                 // every lane that is not in flight is free now
                while (can_fetch) {
                    if (chunk_next >= chunk_end) {
                        if (!more) break;
                        const uint32_t base = wf_claim(head, chunk);
                        if (base >= count) { more = false; break; }
                        chunk_next = base;
                        chunk_end = min(base + chunk, count);
                    }
                    const uint32_t n_idle = (uint32_t)__popc(idle);
                    const uint32_t take = min(n_idle, chunk_end - chunk_next);
                    const uint32_t rank = (uint32_t)__popc(idle & lt);
                    if (mode == PT_WF_MODE_NONE && rank < take) {
                        qi = (uint32_t)PT_IDX(sc, chunk_next + rank, 2 * wf.pcap);
                        const uint32_t entry = wf.q_ray[qi], slot = (uint32_t)PT_IDX(sc, entry & ~PT_WF_SHADOW_BIT, wf.pcap);
                        anyhit = (entry & PT_WF_SHADOW_BIT) != 0;
                        const float4 o4 = wf.ray_o[slot];
                        const float4 d4 = anyhit ? wf.sh_d[slot] : wf.ray_d[slot];
                        r.o = xyz(o4); r.d = xyz(d4);
                        cr = make_cull_ray(r.o, mk3(rcp_approx(r.d.x), rcp_approx(r.d.y), rcp_approx(r.d.z)));
                        res.tri = -1; res.t = PT_INF; res.inside = false;
                        cull = PT_INF; tie = false; steps = 0;
                        stk.clear();
                        const uint32_t root = sc.root_fast;
                        if (ref_count(root) == 0) { ref = root; mode = PT_WF_MODE_NODE; }
                        else { ti = ref_index(root); tend = ti + ref_count(root); mode = PT_WF_MODE_TRI; }
                        nrays++;
                    }
                    chunk_next += take;
                    if (take == n_idle) break;
                    idle = __ballot_sync(0xffffffffu, mode == PT_WF_MODE_NONE);
                }
                if (__ballot_sync(0xffffffffu, mode == PT_WF_MODE_NODE || mode == PT_WF_MODE_TRI) == 0u) break;   // nothing in flight, nothing left
                continue;                                                                                        // vote again with the new rays
            }
        }
        uint32_t next = PT_REF_NONE;                   // where this lane goes after its step (NONE: pop)
        bool advance = false;
        if (__popc(m_node) + PT_WF_NODE_BIAS >= __popc(m_tri)) {
            // ---- node step: both child boxes and refs in one 64-byte record ----
            if (mode == PT_WF_MODE_NODE) {
                const float4* w = sc.wnodes_fast + 4 * (size_t)PT_IDX(sc, ref, sc.nFastInner);
                float4 q0, q1, q2, q3;
                ldg8(w, q0, q1); ldg8(w + 2, q2, q3);
                cnt.inner();
                if (Cfg::COUNT) steps++;
                float e1, e2;
                const bool g1 = slab_cull_fused(cr, mk3(q0.x, q0.y, q0.z), mk3(q0.w, q1.x, q1.y), cull, e1);
                const bool g2 = slab_cull_fused(cr, mk3(q1.z, q1.w, q2.x), mk3(q2.y, q2.z, q2.w), cull, e2);
                const uint32_t lref = __float_as_uint(q3.x), rref = __float_as_uint(q3.y);
                advance = true;
                if (g1 && g2) {
                    const bool left_first = e1 <= e2;
                    if (stk.full()) { *sc.error_flag = PT_ERR_STACK; stk.clear(); }
                    else { stk.push(left_first ? rref : lref); next = left_first ? lref : rref; }
                } else if (g1 || g2) {
                    next = g1 ? lref : rref;
                }
            }
        } else {
            // ---- triangle step: hitArray, one triangle per turn ----
            if (mode == PT_WF_MODE_TRI) {
                const uint32_t i = (uint32_t)PT_IDX(sc, ti, sc.nTriangles);
                ti++;
                const float4* g = sc.trigeo_fast + 4 * (size_t)i;
                float4 a, bq, c, m4;
                ldg8(g, a, bq); ldg8(g + 2, c, m4);
                cnt.tri();
                if (Cfg::COUNT) steps++;
                float tt; bool ins;
                advance = ti == tend;
                if (hit_triangle(r, mk3(a.x, a.y, a.z), mk3(bq.x, bq.y, bq.z), mk3(c.x, c.y, c.z), mk3(a.w, bq.w, c.w), res.t, tt, ins,
                                 !anyhit && res.tri >= 0)) {
                    cnt.cand();
                    const int2 meta = make_int2(__float_as_int(m4.x), __float_as_int(m4.y));
                    bool accept = true;
                    if (meta.y != 1) {                // the reference never box-tests its root (P:342-343)
                        const float* nd = sc.nodes_raw + 12 * (size_t)PT_IDX(sc, meta.y, sc.nNodes);
                        float e;
                        r.inv = mk3(s_rcp(r.d.x), s_rcp(r.d.y), s_rcp(r.d.z));
                        accept = hit_aabb(r, raw3(nd + 6), raw3(nd + 9), e) > 0.0f;
                    }
                    if (accept) {
                        if (!anyhit && tt == res.t) {
                            tie = true;
                        } else {
                            res.t = tt; res.tri = meta.x; res.inside = ins;
                            if (anyhit) { stk.clear(); advance = true; }   // a shadow ray needs no more than this
                            else cull = s_mul(tt, PT_CULL_SLACK);
                        }
                    }
                }
            }
        }
        // ---- where next: the child, or the next deferred subtree, or this ray is finished ----
        if (advance) {
            if (next == PT_REF_NONE) {
                if (stk.empty()) mode = PT_WF_MODE_DONE;
                else next = stk.pop();
            }
            if (mode != PT_WF_MODE_DONE) {
                if (ref_count(next) == 0) { ref = next; mode = PT_WF_MODE_NODE; }
                else { ti = ref_index(next); tend = ti + ref_count(next); mode = PT_WF_MODE_TRI; }
            }
        }
    }
}

template <class Cfg>
__global__ void __launch_bounds__(256, PT_WF_TRACE_CTAS) wf_trace(const __grid_constant__ RenderParams p, int b) {
    const WaveBuffers& wf = p.wf;
#if PT_WF_SMEM_STACK > 0
    __shared__ uint32_t s_stack[PT_WF_SMEM_STACK * PT_STACK_STRIDE];
    uint32_t* const smem_stack = s_stack;
#else
    uint32_t* const smem_stack = nullptr;
#endif
    const uint32_t count = *wf_ctr(wf, WF_C_RAY_TAIL, b);
    uint32_t nrays = 0;
    Counter<Cfg::COUNT> cnt;
    if (Cfg::FAST_TREE) wf_trace_refill<Cfg>(p, b, count, smem_stack, nrays, cnt);
    else wf_trace_simple<Cfg>(p, b, count, nrays, cnt);
    wf_flush_rays(nrays, p.ray_counter);
    flush_counts(cnt, p.work_counters);
}

// ---- classify: results of bounce b -> what the paths need -------------------------------------------------------------
// A streaming pass over the ray queue (12 bytes in per ray, full warps): an occluded shadow ray zeroes its path's pending
// light term; a bounce ray that hit leaves its hit record with the path and the path joins Q_hit, one that escaped Q_miss.
static __global__ void __launch_bounds__(256) wf_classify(const __grid_constant__ RenderParams p, int b) {
    const WaveBuffers& wf = p.wf;
    const uint32_t count = *wf_ctr(wf, WF_C_RAY_TAIL, b);
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i0 = blockIdx.x * blockDim.x; i0 < count; i0 += stride) {          // whole warps stay together for the ballots
        const uint32_t i = i0 + threadIdx.x;
        bool hit = false, miss = false;
        uint32_t slot = 0;
        if (i < count) {
            const uint32_t e = wf.q_ray[i];
            const int2 r = wf.res[i];
            slot = (uint32_t)PT_IDX(p.sc, e & ~PT_WF_SHADOW_BIT, wf.pcap);
            if (e & PT_WF_SHADOW_BIT) {
                if (r.x != -1) wf.cs[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
            } else if (r.x != -1) {
                wf.hit[slot] = r;
                hit = true;
            } else {
                miss = true;
            }
        }
        wf_push(wf.q_hit, wf_ctr(wf, WF_C_HIT_TAIL, b), hit, slot);
        wf_push(wf.q_miss[b & 1], wf_ctr(wf, WF_C_MISS_TAIL, b), miss, slot);
    }
}

// ---- shade: what trace(b-1) resolved ---------------------------------------------------------------------------------
// phase 1: Q_hit[b-1] — the bounce ray hit: emissive term, history, then bounce b;  phase 2: Q_miss[b-1] — it escaped (or
// never flew): environment term, path done.  Both lists are compact: full warps in either phase.
template <class Cfg, int VARIANT>
__global__ void __launch_bounds__(256, PT_WF_HEAD_CTAS) wf_shade(const __grid_constant__ RenderParams p, uint32_t f0, int b) {
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
                    slot = (uint32_t)PT_IDX(sc, wf.q_hit[i], wf.pcap);
                    const float4 ro = wf.ray_o[slot], rd = wf.ray_d[slot], hs = wf.hist[slot], fr = wf.fr[slot], lo4 = wf.lo[slot];
                    const int2 hr = wf.hit[slot];
                    TraceResult tr;
                    tr.tri = hr.x & 0x7fffffff; tr.inside = hr.x < 0; tr.t = __int_as_float(hr.y);
                    V3 Lo = xyz(lo4);
                    if (__float_as_uint(fr.w) & PT_WF_FLAG_SHADOW) Lo = Lo + xyz(wf.cs[slot]);            // the light term of bounce b-1, in order
                    const uint32_t islot = (uint32_t)PT_IDX(sc, wf.pix[slot], wf.icap);
                    const uint32_t fl = islot / wf.npix, idx = islot - fl * wf.npix;
                    const int px = (int)(idx % (uint32_t)p.width), py = (int)(idx / (uint32_t)p.width);
                    const uint32_t frame = p.first_frame + (f0 + fl) * p.frame_stride;
                    nx = wf_shade_hit<Cfg, VARIANT>(p, slot, islot, px, py, frame, p.sobol_tab ? p.sobol_tab + (size_t)(f0 + fl) * 8 : nullptr, b, xyz(ro), xyz(rd), tr,
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
                const uint32_t slot = (uint32_t)PT_IDX(sc, wf.q_miss[(b - 1) & 1][i], wf.pcap);
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
                        const float pdf_light = hdr_pdf<M, Cfg::ENV_CDF>(sc, u, w);
                        const float mis_weight = misMixWeight(pdf, pdf_light);
                        Lo = Lo + (history * mis_weight) * color * f_r * cosw / pdf;
                    } else {                                                                              // O:349-353, D:465-469
                        const V3 sky = mk3(fminf(c.x, 10.0f), fminf(c.y, 10.0f), fminf(c.z, 10.0f));
                        Lo = Lo + history * sky * f_r * cosw / pdf;
                    }
                }
                wf_finish<Cfg>(p, (uint32_t)PT_IDX(sc, wf.pix[slot], wf.icap), Lo, __float_as_int(lo4.w));
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
            const float4 c = wf.img[(size_t)fl * wf.npix + idx];
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
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(p.ray_counter, (unsigned long long)*wf_misc_ctr(wf, 3));   // the wave's primary rays
}

// ---- host side: wave after wave on the context's stream ---------------------------------------------------------------
// launches per wave (the counter memset is not a kernel): head + max_bounce x (trace + classify + shade) + accumulate
__host__ inline int wavefront_launches_per_wave(int max_bounce) { return 2 + 3 * max_bounce; }

// Frames per wave.  Safe start: every owned pixel of every frame could survive, so F0 = pcap / owned pixels (and never more
// image slots than icap).  After each wave the host knows the survivor fraction and sizes the next wave to fill 3/4 of the
// path slots; a wave that still runs out of slots (overflow flag) is repeated with half the frames — nothing has been
// folded into the framebuffer at that point.  One stream synchronisation per wave (tens of ms of GPU work).
struct WaveStats { unsigned long long waves = 0, retries = 0, frames = 0; };
template <class Cfg, int VARIANT>
cudaError_t launch_wavefront_variant(const RenderParams& p, int num_sms, cudaStream_t stream, WaveStats* ws) {
    const WaveBuffers& wf = p.wf;
    const int mb = p.max_bounce;
    const TileMap tm = make_tile_map(p.x0, p.y0, p.x1, p.y1, p.band_world, p.band_rank, p.band_diag);
    const size_t owned = (size_t)tm.count() * 32;                            // upper bound of the pixels this launch renders per frame
    if (owned == 0) return cudaSuccess;
    const uint32_t f_img = (uint32_t)(wf.icap / wf.npix);                    // frames the image-slot array holds
    uint32_t F = (uint32_t)std::min<size_t>(f_img, std::max<size_t>(1, wf.pcap / owned));
    if (VARIANT == 2 && p.sobol_tab)
        sobol_table_kernel<<<(p.n_frames * 8 + 255) / 256, 256, 0, stream>>>(const_cast<float*>(p.sobol_tab), p.first_frame, p.n_frames, p.frame_stride);
    int per_sm = 0, per_sm_head = 0, per_sm_shade = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wf_trace<Cfg>, 256, 0);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_head, wf_head<Cfg, VARIANT>, 256, 0);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_shade, wf_shade<Cfg, VARIANT>, 256, 0);
    if (e != cudaSuccess) return e;
    const int grid_trace = num_sms * (per_sm > 0 ? per_sm : 1), grid_head = num_sms * (per_sm_head > 0 ? per_sm_head : 1),
              grid_shade = num_sms * (per_sm_shade > 0 ? per_sm_shade : 1), grid = num_sms * 8;
    uint32_t f0 = 0;
    while (f0 < p.n_frames) {
        const uint32_t nf = std::min(p.n_frames - f0, F);
        e = cudaMemsetAsync(wf.counters, 0, (size_t)PT_WF_NUM_CTRS * PT_WF_CTR_STRIDE * sizeof(uint32_t), stream);
        if (e != cudaSuccess) return e;
        wf_head<Cfg, VARIANT><<<grid_head, 256, 0, stream>>>(p, f0, nf);
        uint32_t paths = 0, overflow = 0;
        e = cudaMemcpyAsync(&paths, wf_misc_ctr(wf, 1), sizeof paths, cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(&overflow, wf_misc_ctr(wf, 2), sizeof overflow, cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) return e;
        if (overflow) {                                                       // only reachable after F grew on an optimistic estimate
            F = std::max(1u, nf / 2u);
            if (ws) ws->retries++;
            continue;
        }
        for (int b = 0; b < mb; b++) {
            wf_trace<Cfg><<<grid_trace, 256, 0, stream>>>(p, b);
            wf_classify<<<grid, 256, 0, stream>>>(p, b);
            wf_shade<Cfg, VARIANT><<<grid_shade, 256, 0, stream>>>(p, f0, b + 1);
        }
        wf_accumulate<Cfg><<<grid, 256, 0, stream>>>(p, f0, nf);
        if (ws) { ws->waves++; ws->frames += nf; }
        f0 += nf;
        // next wave: fill 3/4 (WaveBuffers.fill) of the path slots at the survivor rate just measured
        const double per_frame = std::max(1.0, (double)paths / (double)nf);
        const double want = (double)wf.fill * (double)wf.pcap / per_frame;
        F = (uint32_t)std::min<double>((double)f_img, std::max(1.0, want));
    }
    return cudaGetLastError();
}

template <class Cfg>
cudaError_t launch_wavefront(const RenderParams& p, int variant, int num_sms, cudaStream_t stream, WaveStats* ws) {
    if (variant == 0) return launch_wavefront_variant<Cfg, 0>(p, num_sms, stream, ws);
    if (variant == 1) return launch_wavefront_variant<Cfg, 1>(p, num_sms, stream, ws);
    return launch_wavefront_variant<Cfg, 2>(p, num_sms, stream, ws);
}

}  // namespace pt
