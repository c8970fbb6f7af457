// pt_trace.cuh — ray/box, ray/triangle and the two BVH traversals (the hitBVH of the reference shaders).
//
// Reference: ImportanceSampling_LowDiscrepancySequence/shaders/pass1.fsh ("P:") hitTriangle 251-301,
// hitAABB 303-316, hitArray 319-332, hitBVH 335-382 (same code in the other two shaders).
//
// Two traversals with IDENTICAL results:
//   trace_ref  — the reference algorithm on the reference layout: int stack[256], three 48-byte node
//                fetches per inner pop, every leaf whose box is touched is fully tested, no culling.
//   trace_wide — same near-first order, on 64-byte wide nodes (both child boxes in the parent: one fetch
//                per inner node), packed child refs with leaves folded into their parent, closest-hit
//                culling against the current best distance, optional any-hit exit for shadow rays, and
//                48-byte triangle records with the face normal precomputed at upload.
// Both keep the reference's strict `<` closest-hit rule and leaf/triangle visiting order, so ties are
// resolved the same way; all tests run in un-fused IEEE arithmetic (pt_common.cuh "strict").
#pragma once
#include "pt_common.cuh"

namespace pt {

#define PT_INF 2147483648.0f     // float(INF), P:24

struct Ray {
    V3 o, d, inv;                // inv = 1/d, computed once per ray (the reference recomputes it per box)
};
__device__ __forceinline__ Ray make_ray(V3 o, V3 d) {
    Ray r; r.o = o; r.d = d;
    r.inv = mk3(s_rcp(d.x), s_rcp(d.y), s_rcp(d.z));
    return r;
}

// hitAABB, P:303-316.  Returns the reference's value (entry distance, or EXIT distance when the origin is
// inside, or -1) and the entry distance clamped to 0 for culling.
__device__ __forceinline__ float hit_aabb(const Ray& r, V3 AA, V3 BB, float& entry) {
    float fx = s_mul(s_sub(BB.x, r.o.x), r.inv.x), nx = s_mul(s_sub(AA.x, r.o.x), r.inv.x);
    float fy = s_mul(s_sub(BB.y, r.o.y), r.inv.y), ny = s_mul(s_sub(AA.y, r.o.y), r.inv.y);
    float fz = s_mul(s_sub(BB.z, r.o.z), r.inv.z), nz = s_mul(s_sub(AA.z, r.o.z), r.inv.z);
    float t1 = fminf(fmaxf(fx, nx), fminf(fmaxf(fy, ny), fmaxf(fz, nz)));
    float t0 = fmaxf(fminf(fx, nx), fmaxf(fminf(fy, ny), fminf(fz, nz)));
    entry = fmaxf(t0, 0.0f);
    return (t1 >= t0) ? ((t0 > 0.0f) ? t0 : t1) : -1.0f;
}

// hitTriangle, P:251-301, up to the hit decision.  N is the unit face normal normalize(cross(p2-p1,p3-p1))
// BEFORE the reference's back-face flip: t and the inside test are invariant under N -> -N in IEEE
// arithmetic (negation commutes exactly with *, + and /), so the flip only sets isInside.
// `best` implements the strict closest-hit rule: a candidate must have t < best to matter (t == best also
// qualifies when allow_tie is set — see trace_wide's early far-leaf test).
__device__ __forceinline__ bool hit_triangle(const Ray& r, V3 p1, V3 p2, V3 p3, V3 N, float best, float& t_out,
                                             bool& inside, bool allow_tie = false) {
    float Nd = s_dot(N, r.d);
    if (fabsf(Nd) < 0.00001f) return false;
    const float num = s_sub(s_dot(N, p1), s_dot(r.o, N));
    // t = num / Nd can only reach 0.0005 if num and Nd have the same sign and num is not 0 (|Nd| >= 1e-5, so a product that
    // underflows to 0 means t < 1e-35): everything else is rejected WITHOUT the division — same decision, and it keeps the
    // exact-zero numerators of rays that start on the plane of the triangle (every secondary ray leaving a flat floor) off
    // the IEEE divide's slow path (8 M warp-level slow-path calls per 64-frame launch on is-1M, profiles/r01_render_persistent.md).
    if (!(s_mul(num, Nd) > 0.0f) && !(num != num)) return false;
    float t = s_div(num, Nd);
    if (t < 0.0005f) return false;
    if (!(t < best) && !(allow_tie && t == best)) return false;
    V3 P = s_point(r.o, r.d, t);
    float s1 = s_dot(s_cross(s_sub3(p2, p1), s_sub3(P, p1)), N);
    float s2 = s_dot(s_cross(s_sub3(p3, p2), s_sub3(P, p2)), N);
    float s3 = s_dot(s_cross(s_sub3(p1, p3), s_sub3(P, p3)), N);
    bool pos = (s1 > 0.0f) && (s2 > 0.0f) && (s3 > 0.0f);
    bool ngt = (s1 < 0.0f) && (s2 < 0.0f) && (s3 < 0.0f);
    if (!(pos || ngt)) return false;
    t_out = t;
    inside = Nd > 0.0f;
    return true;
}

// Culling test of the production traversal: the decision "hitAABB(...) > 0 and entry <= cull",
//   hitAABB > 0  <=>  t1 >= t0 and t1 > 0   (t0 > 0 implies t1 > 0; otherwise the returned value is t1)
// with hit_aabb's NaN-dropping min/max nesting.
// The PRODUCTION tree's boxes only cull (padded by 2^-17 of the scene radius, pt_build.cpp): one
// fused multiply-add per plane, t = box * (1/d) - o * (1/d), instead of a subtraction and a multiplication.  The fused
// form is off by a few ulps of the ORIGIN's coordinates (6e-8 |o| in space), some fifty times less than the padding, so
// "passes" stays a superset of "hitTriangle accepts"; what the reference's own boxes decide is tested exactly, per
// candidate, with hit_aabb.  1/d is clamped to +-1e18 here: with an infinite 1/d the fused form gives inf - inf = NaN on
// BOTH planes of a slab the origin lies in, and the NaN-dropping min/max would then reject a ray that passes.
struct CullRay { V3 inv, noi; };                      // clamped 1/d and -(o * that)
// The culling reciprocal is the SFU's (rcp.approx: one instruction, 1 ulp) rather than the correctly rounded one: a
// relative 2^-23 on every slab distance, again far inside the padding and the 2^-16 slack on `cull`.  The exact 1/d of the
// reference's hitAABB is computed only for the rare candidate that reaches the leaf-box test.
__device__ __forceinline__ float rcp_approx(float a) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r;
}
__device__ __forceinline__ CullRay make_cull_ray(V3 o, V3 inv) {
    CullRay c;
    c.inv = mk3(fminf(fmaxf(inv.x, -1e18f), 1e18f), fminf(fmaxf(inv.y, -1e18f), 1e18f), fminf(fmaxf(inv.z, -1e18f), 1e18f));
    c.noi = mk3(-s_mul(o.x, c.inv.x), -s_mul(o.y, c.inv.y), -s_mul(o.z, c.inv.z));
    return c;
}
__device__ __forceinline__ bool slab_cull_fused(const CullRay& c, V3 AA, V3 BB, float cull, float& entry) {
    float fx = __fmaf_rn(BB.x, c.inv.x, c.noi.x), nx = __fmaf_rn(AA.x, c.inv.x, c.noi.x);
    float fy = __fmaf_rn(BB.y, c.inv.y, c.noi.y), ny = __fmaf_rn(AA.y, c.inv.y, c.noi.y);
    float fz = __fmaf_rn(BB.z, c.inv.z, c.noi.z), nz = __fmaf_rn(AA.z, c.inv.z, c.noi.z);
    float t1 = fminf(fmaxf(fx, nx), fminf(fmaxf(fy, ny), fmaxf(fz, nz)));
    float t0 = fmaxf(fminf(fx, nx), fmaxf(fminf(fy, ny), fminf(fz, nz)));
    entry = fmaxf(t0, 0.0f);
    return (t1 >= t0) && (t1 > 0.0f) && (entry <= cull);
}

struct TraceResult {
    int tri;                     // -1 = miss
    float t;
    bool inside;
};

// ---- the reference traversal on the reference layout -------------------------------------------
__device__ __forceinline__ V3 raw3(const float* p) { return mk3(ldg1(p), ldg1(p + 1), ldg1(p + 2)); }

static __device__ __noinline__ TraceResult trace_ref(const DevScene& sc, V3 origin, V3 dir) {
    Ray r = make_ray(origin, dir);
    TraceResult res; res.tri = -1; res.t = PT_INF; res.inside = false;
    int stack[256];                                   // P:340
    int sp = 0;
    stack[sp++] = 1;
    while (sp > 0) {
        int top = stack[--sp];
        const float* nd = sc.nodes_raw + 12 * (size_t)top;
        int left = (int)ldg1(nd), right = (int)ldg1(nd + 1), n = (int)ldg1(nd + 3), index = (int)ldg1(nd + 4);
        if (n > 0) {                                  // hitArray over [index, index+n-1]
            for (int i = index; i < index + n; i++) {
                const float* t = sc.tris_raw + 36 * (size_t)i;
                V3 p1 = raw3(t), p2 = raw3(t + 3), p3 = raw3(t + 6);
                V3 N = s_normalize(s_cross(s_sub3(p2, p1), s_sub3(p3, p1)));
                float tt; bool ins;
                if (hit_triangle(r, p1, p2, p3, N, res.t, tt, ins)) { res.t = tt; res.tri = i; res.inside = ins; }
            }
            continue;
        }
        float d1 = PT_INF, d2 = PT_INF, e;
        if (left > 0) { const float* c = sc.nodes_raw + 12 * (size_t)left; d1 = hit_aabb(r, raw3(c + 6), raw3(c + 9), e); }
        if (right > 0) { const float* c = sc.nodes_raw + 12 * (size_t)right; d2 = hit_aabb(r, raw3(c + 6), raw3(c + 9), e); }
        if (sp + 2 > 256) { *sc.error_flag = PT_ERR_STACK; break; }   // the reference would write past stack[256] here
        if (d1 > 0.0f && d2 > 0.0f) {
            if (d1 < d2) { stack[sp++] = right; stack[sp++] = left; }
            else { stack[sp++] = left; stack[sp++] = right; }
        } else if (d1 > 0.0f) stack[sp++] = left;
        else if (d2 > 0.0f) stack[sp++] = right;
    }
    return res;
}

// ---- the wide traversal --------------------------------------------------------------------------
// A child is skipped when its box ENTRY distance exceeds best*(1+2^-16): the slab entry and the plane
// distance t of a triangle on the box face are different roundings of the same number, so a bare
// `entry > best` could, in principle, cull the very triangle that defines best's successor.
#define PT_CULL_SLACK 1.0000153f
// Deferred-subtree stack.  Only INNER nodes are ever deferred (see below), which keeps the need small even on
// the reference builder's degenerate SAH chains (tree depth 6781 on the 1M-triangle scene; deepest stack seen
// by the oracle's model of this traversal: 59).  Overflow is detected, flagged and reported as PT_ERANGE.
#define PT_STACK 128

struct WideState {
    TraceResult res;
    float cull;
    int early_sp;                // >= 0: res came from a far leaf tested ahead of its turn at that stack depth
    bool tie;                    // trace_fast only: two different triangles hit at exactly the same distance
};

// hitArray over one leaf.  `early`: this leaf is a FAR child being tested before the near subtree (its turn in
// the reference's order comes later).  Exact-distance ties must still go to whoever the reference visits first:
// an early hit needs strictly smaller t, and while an early hit is the best, in-turn candidates win ties.
template <bool ANYHIT>
__device__ __forceinline__ bool wide_leaf(const DevScene& sc, const Ray& r, uint32_t ref, bool early, int sp, WideState& st) {
    const uint32_t first = ref_index(ref), cnt = ref_count(ref);
    for (uint32_t i = first; i < first + cnt; i++) {
        const float4* g = sc.trigeo + 3 * (size_t)i;
        float4 a = ldg4(g), b = ldg4(g + 1), c = ldg4(g + 2);
        float tt; bool ins;
        const bool allow_tie = !early && st.early_sp >= 0;
        if (hit_triangle(r, mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), mk3(c.x, c.y, c.z), mk3(a.w, b.w, c.w), st.res.t, tt, ins,
                         allow_tie)) {
            st.res.t = tt; st.res.tri = (int)i; st.res.inside = ins;
            if (ANYHIT) return true;
            st.cull = s_mul(tt, PT_CULL_SLACK);
            st.early_sp = early ? sp : -1;
        }
    }
    return false;
}

template <bool ANYHIT, int STACK>
__device__ __forceinline__ TraceResult trace_wide(const DevScene& sc, V3 origin, V3 dir) {
    Ray r = make_ray(origin, dir);
    WideState st;
    st.res.tri = -1; st.res.t = PT_INF; st.res.inside = false;
    st.cull = PT_INF; st.early_sp = -1; st.tie = false;
    uint32_t stack_ref[STACK];
    float stack_t[STACK];
    int sp = 0;
    uint32_t ref = sc.root_ref;
    if (ref_count(ref)) {                             // the whole scene is one leaf
        wide_leaf<ANYHIT>(sc, r, ref, false, 0, st);
        return st.res;
    }
    while (true) {
        // ref is an inner node: one 64-byte record holds both child boxes and both child refs
        const float4* w = sc.wnodes + 4 * (size_t)ref_index(ref);
        float4 q0 = ldg4(w), q1 = ldg4(w + 1), q2 = ldg4(w + 2), q3 = ldg4(w + 3);
        float e1, e2;
        float d1 = hit_aabb(r, mk3(q0.x, q0.y, q0.z), mk3(q0.w, q1.x, q1.y), e1);
        float d2 = hit_aabb(r, mk3(q1.z, q1.w, q2.x), mk3(q2.y, q2.z, q2.w), e2);
        const uint32_t lref = __float_as_uint(q3.x), rref = __float_as_uint(q3.y);
        const bool h1 = d1 > 0.0f, h2 = d2 > 0.0f;    // the reference's "push" conditions (P:367-379)
        uint32_t next = PT_REF_NONE;
        if (h1 && h2) {
            const bool left_first = d1 < d2;          // near child first (P:368)
            const uint32_t nref = left_first ? lref : rref, fref = left_first ? rref : lref;
            const float ne = left_first ? e1 : e2, fe = left_first ? e2 : e1;
            const bool nleaf = ref_count(nref) != 0, fleaf = ref_count(fref) != 0;
            if (nleaf) {
                if (ne <= st.cull && wide_leaf<ANYHIT>(sc, r, nref, false, sp, st)) return st.res;
                if (fleaf) { if (fe <= st.cull && wide_leaf<ANYHIT>(sc, r, fref, false, sp, st)) return st.res; }
                else if (fe <= st.cull) next = fref;
            } else if (fleaf) {
                // far leaf, near subtree: test the leaf NOW rather than defer it
                if (fe <= st.cull && wide_leaf<ANYHIT>(sc, r, fref, true, sp, st)) return st.res;
                if (ne <= st.cull) next = nref;
            } else {
                const bool ng = ne <= st.cull, fg = fe <= st.cull;
                if (ng && fg) {
                    if (sp >= STACK) { *sc.error_flag = PT_ERR_STACK; break; }
                    stack_ref[sp] = fref; stack_t[sp] = fe; sp++;
                    next = nref;
                } else if (ng) next = nref;
                else if (fg) next = fref;
            }
        } else if (h1 || h2) {
            const uint32_t cref = h1 ? lref : rref;
            const float ce = h1 ? e1 : e2;
            if (ce <= st.cull) {
                if (ref_count(cref)) { if (wide_leaf<ANYHIT>(sc, r, cref, false, sp, st)) return st.res; }
                else next = cref;
            }
        }
        if (next != PT_REF_NONE) { ref = next; continue; }
        // pop the next deferred subtree still worth visiting
        bool found = false;
        while (sp > 0) {
            sp--;
            if (st.early_sp > sp) st.early_sp = -1;   // the early leaf's own turn has passed: ties are strict again
            if (stack_t[sp] <= st.cull) { ref = stack_ref[sp]; found = true; break; }
        }
        if (!found) break;
    }
    return st.res;
}

// ---- the fast-topology traversal -------------------------------------------------------------------
// Same node format, over the binned-SAH tree pt_build.cpp builds over the individual triangles (boxes padded,
// triangles re-ordered).  The padded boxes only CULL; what the reference's tree means for the result is applied
// exactly, per candidate: a triangle that passes hitTriangle counts only if the ray also passes the reference's
// hitAABB(...) > 0 on the box of the reference leaf that holds it (P:356-379; that implies all its ancestors).
// The root is the one node the reference never box-tests (P:342-343).  The visiting order here is
// nearest-entry-first (not the reference's), which can only matter for an exact tie between two triangles: that is
// detected (st.tie) and the ray is re-traced on the reference topology.
// Work counters of the fast traversal (DESIGN.md "Roofline"): compiled in only for the counting configuration.
struct TraceCount {
    uint32_t inner, tris, cand, ties, shaded, texels;   // inner nodes fetched, triangles tested, candidates that reached the leaf-box test, tie re-traces
};
template <bool ON> struct Counter {
    TraceCount c;
    __device__ __forceinline__ Counter() { c.inner = c.tris = c.cand = c.ties = c.shaded = c.texels = 0; }
    __device__ __forceinline__ void inner() { c.inner++; }
    __device__ __forceinline__ void tri() { c.tris++; }
    __device__ __forceinline__ void cand() { c.cand++; }
    __device__ __forceinline__ void tie() { c.ties++; }
    __device__ __forceinline__ void shaded() { c.shaded++; }     // closest hits whose shading data is fetched
    __device__ __forceinline__ void texel(uint32_t n) { c.texels += n; }   // environment map / hdrCache texels fetched
};
template <> struct Counter<false> {
    __device__ __forceinline__ void inner() {}
    __device__ __forceinline__ void tri() {}
    __device__ __forceinline__ void cand() {}
    __device__ __forceinline__ void tie() {}
    __device__ __forceinline__ void shaded() {}
    __device__ __forceinline__ void texel(uint32_t) {}
};

// ---- the deferred-subtree stack of trace_fast (the reference's `int stack[256]`, P:340) -----------------------
// PT_SMEM_STACK = entries per thread kept in SHARED memory, laid out [level][thread] (consecutive lanes hit consecutive
// banks: conflict-free); deeper entries overflow into a local-memory array, so the total capacity stays PT_STACK.
// PT_SMEM_STACK = 0 keeps the whole stack in local memory (L1-resident).  PT_STACK_TOS = 1 additionally holds the top
// entry in a register, which takes the stack load off the pop -> node-address dependency chain.  Measured: DESIGN.md §9.
#ifndef PT_SMEM_STACK
#define PT_SMEM_STACK 0
#endif
// Triangle records of trace_fast (the persistent kernel and the megakernel): PT_TRI64 = 0: 48-byte records (three 16-byte loads)
// + the (index, leaf) pair fetched only for the 0.25 candidates per ray; 1: the 64-byte records wf_trace reads (two LDG.E.256).
// The 64-byte form costs these kernels 33 % more triangle bytes through an L1 that already misses half of the time:
// is-1M 21.3 vs 22.0 Grays/s in the same box (round-1 build as the yardstick), so they keep the 48-byte form.
// trace_fast's stack when nothing of it lives in shared memory or a register: a plain local array and an int, written out —
// NOT the TravStack template with SMEM = 0, although that is the same program: ptxas lays the loop out differently and the
// persistent kernel loses 1.3 % on is-1M (47.41 vs 46.84 ms per 256-frame launch, same box, tools/r02_call11.sh).  "Code shape
// is worth per cents" (DESIGN.md section 9).
#ifndef PT_PLAIN_STACK
#define PT_PLAIN_STACK (PT_SMEM_STACK == 0 && !PT_STACK_TOS)
#endif
#ifndef PT_TRI64
#define PT_TRI64 0
#endif
#ifndef PT_NODE256
#define PT_NODE256 0                  // trace_fast: wide nodes by two 32-byte loads (1) or four 16-byte loads (0): 0 is +0.4 % here (wf_trace: 256-bit, +4 %)
#endif
#ifndef PT_FAR_EYE
#define PT_FAR_EYE 1                  // primary rays of an eye beyond 8 scene radii take the reference topology (pt_shade.cuh trace<>)
#endif
#ifndef PT_STACK_TOS
#define PT_STACK_TOS 0
#endif
#define PT_STACK_STRIDE 256           // threads per CTA of every kernel that traverses
template <int STACK, int SMEM>
struct TravStack {
    uint32_t local[STACK > SMEM ? STACK - SMEM : 1];
    uint32_t* sm;
#if PT_STACK_TOS
    uint32_t tos;
#endif
    int sp;
    __device__ __forceinline__ void init(uint32_t* smem_base) {
        sp = 0;
        sm = SMEM > 0 ? smem_base + threadIdx.x : nullptr;
    }
    __device__ __forceinline__ void put(int k, uint32_t v) {
        if (SMEM > 0 && k < SMEM) { sm[k * PT_STACK_STRIDE] = v; return; }
        if (STACK > SMEM) local[k - SMEM] = v;
    }
    __device__ __forceinline__ uint32_t get(int k) const {
        if (SMEM > 0 && k < SMEM) return sm[k * PT_STACK_STRIDE];
        return STACK > SMEM ? local[k - SMEM] : PT_REF_NONE;
    }
    __device__ __forceinline__ bool full() const { return sp >= STACK; }
    __device__ __forceinline__ bool empty() const { return sp == 0; }
    __device__ __forceinline__ void clear() { sp = 0; }
    __device__ __forceinline__ void push(uint32_t v) {
#if PT_STACK_TOS
        if (sp > 0) put(sp - 1, tos);
        tos = v; sp++;
#else
        put(sp++, v);
#endif
    }
    __device__ __forceinline__ uint32_t pop() {
#if PT_STACK_TOS
        const uint32_t v = tos;
        if (--sp > 0) tos = get(sp - 1);
        return v;
#else
        return get(--sp);
#endif
    }
};

// `anyhit` is a run-time flag (shadow rays stop at the first accepted triangle): ONE copy of this loop serves every
// ray of a path.  While-while form (one node-test site, one triangle-test site: the hot loop is ~250 SASS instructions and stays
// in the instruction cache): descend inner nodes — nearer child first, farther child pushed — until the current
// ref is a leaf; test its triangles; pop.  Leaves travel through the stack like inner nodes.
// Loop shape: ONE exit and ONE pop site.  A lane that is done (stack empty, or a shadow ray that found its hit: sp = 0)
// carries ref = PT_REF_NONE — whose count bits are set, so it also falls out of the node loop — down to the single `break`
// at the bottom; a node whose children both miss leaves the node loop the same way and pops where the leaves pop.  With
// `return`s, a `goto` and a second pop loop inside the nested loops the compiler kept a separate path per site and the
// lanes of a warp reconverged later: same instructions per ray, is-1M +3.6 %, disney-1M +11 % (tools/ab_run.sh).
// The stack (TravStack above: shared memory and/or local memory) holds refs only: a deferred subtree is culled by the box tests of its own node when it is popped (a deferred
// LEAF by hit_triangle's t < best), not by an entry distance kept beside the ref — half the stack traffic and no pop loop:
// is-1M +1.3 %, disney-1M +2.7 %.
template <int STACK, class CNT>
__device__ __forceinline__ TraceResult trace_fast(const DevScene& sc, V3 origin, V3 dir, const bool anyhit, CNT& cnt) {
    Ray r; r.o = origin; r.d = dir;                   // r.inv stays unset until a candidate needs the exact 1/d
    const CullRay cr = make_cull_ray(origin, mk3(rcp_approx(dir.x), rcp_approx(dir.y), rcp_approx(dir.z)));
    TraceResult res; res.tri = -1; res.t = PT_INF; res.inside = false;
    float cull = PT_INF;
    bool tie = false;
#if PT_SMEM_STACK > 0
    __shared__ uint32_t s_stack[PT_SMEM_STACK * PT_STACK_STRIDE];
    uint32_t* const smem_base = s_stack;
#else
    uint32_t* const smem_base = nullptr;
#endif
#if PT_PLAIN_STACK
    uint32_t stack_ref[STACK];                        // the round-1 form, verbatim: a plain local array and an int
    int sp = 0;
    (void)smem_base;
#else
    TravStack<STACK, PT_SMEM_STACK> stk;
    stk.init(smem_base);
#endif
    uint32_t ref = sc.root_fast;
    while (true) {
        while (ref_count(ref) == 0) {                 // inner node: both child boxes and refs in one 64-byte record
            const float4* w = sc.wnodes_fast + 4 * (size_t)PT_IDX(sc, ref, sc.nFastInner);
#if PT_NODE256
            float4 q0, q1, q2, q3;
            ldg8(w, q0, q1); ldg8(w + 2, q2, q3);
#else
            const float4 q0 = ldg4(w), q1 = ldg4(w + 1), q2 = ldg4(w + 2), q3 = ldg4(w + 3);
#endif
            cnt.inner();
            float e1, e2;
            const bool g1 = slab_cull_fused(cr, mk3(q0.x, q0.y, q0.z), mk3(q0.w, q1.x, q1.y), cull, e1);
            const bool g2 = slab_cull_fused(cr, mk3(q1.z, q1.w, q2.x), mk3(q2.y, q2.z, q2.w), cull, e2);
            const uint32_t lref = __float_as_uint(q3.x), rref = __float_as_uint(q3.y);
            if (g1 && g2) {
                const bool left_first = e1 <= e2;
#if PT_PLAIN_STACK
                if (sp >= STACK) { *sc.error_flag = PT_ERR_STACK; return res; }
                stack_ref[sp++] = left_first ? rref : lref;
#else
                if (stk.full()) { *sc.error_flag = PT_ERR_STACK; return res; }
                stk.push(left_first ? rref : lref);
#endif
                ref = left_first ? lref : rref;
            } else if (g1 || g2) {
                ref = g1 ? lref : rref;
            } else {
                ref = PT_REF_NONE;                    // both children missed (the count bits of NONE are set: leaves the node loop)
            }
        }
        if (ref != PT_REF_NONE) {                     // leaf: hitArray over its (re-ordered) triangles
            const uint32_t first = ref_index(ref), end = first + ref_count(ref);
            for (uint32_t i = first; i < end; i++) {
#if PT_TRI64
                const float4* g = sc.trigeo_fast + 4 * (size_t)PT_IDX(sc, i, sc.nTriangles);
                float4 a, b, c, m4;
                ldg8(g, a, b); ldg8(g + 2, c, m4);
#else
                const float4* g = sc.trigeo_fast48 + 3 * (size_t)PT_IDX(sc, i, sc.nTriangles);
                const float4 a = ldg4(g), b = ldg4(g + 1), c = ldg4(g + 2);
#endif
                cnt.tri();
                float tt; bool ins;
                if (hit_triangle(r, mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), mk3(c.x, c.y, c.z), mk3(a.w, b.w, c.w), res.t, tt, ins,
                                 !anyhit && res.tri >= 0)) {
                    cnt.cand();
#if PT_TRI64
                    const int2 meta = make_int2(__float_as_int(m4.x), __float_as_int(m4.y));
#else
                    const int2 meta = __ldg(sc.trimeta + i);
#endif
                    if (meta.y != 1) {                // the reference never box-tests its root (P:342-343)
                        const float* nd = sc.nodes_raw + 12 * (size_t)PT_IDX(sc, meta.y, sc.nNodes);
                        float e;
                        r.inv = mk3(s_rcp(dir.x), s_rcp(dir.y), s_rcp(dir.z));
                        if (!(hit_aabb(r, raw3(nd + 6), raw3(nd + 9), e) > 0.0f)) continue;
                    }
                    if (!anyhit && tt == res.t) { tie = true; continue; }
                    res.t = tt; res.tri = meta.x; res.inside = ins;
#if PT_PLAIN_STACK
                    if (anyhit) { sp = 0; break; }    // a shadow ray needs no more than this
#else
                    if (anyhit) { stk.clear(); break; }   // a shadow ray needs no more than this
#endif
                    cull = s_mul(tt, PT_CULL_SLACK);
                }
            }
        }
#if PT_PLAIN_STACK
        if (sp == 0) break;                           // the ONE pop site (after a leaf, and after a node whose children both missed)
        ref = stack_ref[--sp];                        // and the one exit: this lane's ray is finished when nothing is deferred
#else
        if (stk.empty()) break;
        ref = stk.pop();
#endif
    }
    if (!anyhit && tie) res.tri = -2;                 // caller: re-trace on the reference topology
    return res;
}

// out-of-line so that the (practically never taken) tie path costs no registers in the hot loop
static __device__ __noinline__ TraceResult trace_wide_fallback(const DevScene& sc, V3 origin, V3 dir) {
    return trace_wide<false, PT_STACK>(sc, origin, dir);
}

// a / b with the IEEE result, taking the common exact-zero numerator (horizontal triangles: the xy-only barycentrics
// of P:293-295 degenerate) around the divide's slow path: 0 / b = 0 * b (same signed zero) for every b except 0 and NaN/inf
__device__ __forceinline__ float div_z(float a, float b) {
    if (a == 0.0f && b != 0.0f && fabsf(b) <= 3.0e38f) return s_mul(a, b);
    return s_div(a, b);
}

// ---- what the shader's HitResult carries for the winning triangle (P:285-299) -------------------
template <bool RAW_LAYOUT>
__device__ __forceinline__ void finalize_hit(const DevScene& sc, V3 origin, V3 dir, const TraceResult& tr, Hit& h,
                                             int& mat_id) {
    h.isHit = tr.tri >= 0;
    h.tri = tr.tri;
    h.distance = tr.t;
    h.isInside = tr.inside;
    h.viewDir = dir;
    mat_id = 0;
    if (!h.isHit) return;
    V3 p1, p2, p3, n1, n2, n3;
    if (RAW_LAYOUT) {
        const float* t = sc.tris_raw + 36 * (size_t)tr.tri;
        p1 = raw3(t); p2 = raw3(t + 3); p3 = raw3(t + 6); n1 = raw3(t + 9); n2 = raw3(t + 12); n3 = raw3(t + 15);
        mat_id = tr.tri;                              // material read from the triangle record itself
    } else {
        const float4* g = sc.trigeo + 3 * (size_t)PT_IDX(sc, tr.tri, sc.nTriangles);
        const float4* s = sc.trishade + 3 * (size_t)PT_IDX(sc, tr.tri, sc.nTriangles);
        float4 a = ldg4(g), b = ldg4(g + 1), c = ldg4(g + 2), u = ldg4(s), v = ldg4(s + 1), w = ldg4(s + 2);
        p1 = mk3(a.x, a.y, a.z); p2 = mk3(b.x, b.y, b.z); p3 = mk3(c.x, c.y, c.z);
        n1 = mk3(u.x, u.y, u.z); n2 = mk3(v.x, v.y, v.z); n3 = mk3(w.x, w.y, w.z);
        mat_id = __float_as_int(u.w);
    }
    V3 P = s_point(origin, dir, tr.t);
    h.hitPoint = P;
    // P:293-297 — barycentrics from x and y only, with the author's 5e-5 offsets
    float alpha = div_z(-(P.x - p2.x) * (p3.y - p2.y) + (P.y - p2.y) * (p3.x - p2.x),
                        -(p1.x - p2.x - 0.00005f) * (p3.y - p2.y + 0.00005f) + (p1.y - p2.y + 0.00005f) * (p3.x - p2.x + 0.00005f));
    float beta = div_z(-(P.x - p3.x) * (p1.y - p3.y) + (P.y - p3.y) * (p1.x - p3.x),
                       -(p2.x - p3.x - 0.00005f) * (p1.y - p3.y + 0.00005f) + (p2.y - p3.y + 0.00005f) * (p1.x - p3.x + 0.00005f));
    float gama = 1.0f - alpha - beta;
    V3 Ns = normalize(n1 * alpha + n2 * beta + n3 * gama);
    h.normal = tr.inside ? -Ns : Ns;
}

template <bool RAW_LAYOUT>
__device__ __forceinline__ void load_material(const DevScene& sc, int mat_id, Material& m) {
    if (RAW_LAYOUT) {                                 // getMaterial, P:207-232: texels 6..11 of the triangle
        const float* t = sc.tris_raw + 36 * (size_t)mat_id + 18;
        m.emissive = raw3(t); m.baseColor = raw3(t + 3);
        m.subsurface = ldg1(t + 6); m.metallic = ldg1(t + 7); m.specular = ldg1(t + 8);
        m.specularTint = ldg1(t + 9); m.roughness = ldg1(t + 10); m.anisotropic = ldg1(t + 11);
        m.sheen = ldg1(t + 12); m.sheenTint = ldg1(t + 13); m.clearcoat = ldg1(t + 14); m.clearcoatGloss = ldg1(t + 15);
    } else {
        const float4* p = sc.materials + 8 * (size_t)PT_IDX(sc, mat_id, sc.nMaterials);      // 128-byte records: 4 float4 as uploaded, 4 derived (pt_shade.cuh)
        float4 a = ldg4(p), b = ldg4(p + 1), c = ldg4(p + 2), d = ldg4(p + 3);
        m.emissive = mk3(a.x, a.y, a.z); m.subsurface = a.w;
        m.baseColor = mk3(b.x, b.y, b.z); m.metallic = b.w;
        m.specular = c.x; m.specularTint = c.y; m.roughness = c.z; m.anisotropic = c.w;
        m.sheen = d.x; m.sheenTint = d.y; m.clearcoat = d.z; m.clearcoatGloss = d.w;
    }
}

}  // namespace pt
