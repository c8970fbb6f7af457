/* pt.h — C ABI of libpt.so: the B200 (sm_100a) CUDA path tracer that replaces the reference's
 * OpenGL dispatch of its per-pixel pathTracing() fragment shaders.
 *
 * The reference has no plugin/FFI layer; its "interface" to the hot path is the set of GL calls that
 * hand buffers and uniforms to pass1.fsh and draw the full-screen quad.  Each entry point below names
 * the reference call sites it replaces.  Citations: "M:" = ImportanceSampling_LowDiscrepancySequence/
 * main.cpp, "P:" = ImportanceSampling_LowDiscrepancySequence/shaders/pass1.fsh (OpenglRayTracing/ and
 * DisneyBRDF/ carry the same host code and the two smaller shaders).
 *
 * Conventions: plain C, no torch / C++ types.  Every function returns PT_OK (0) or a negative PT_E*
 * code; pt_last_error() returns a thread-local message.  Nothing here calls exit() (the reference
 * does, M:176, M:270) and no exception crosses the boundary.  Host arrays are caller-owned and copied
 * during the call.  A context is bound to one CUDA device and one stream; calls on one context are not
 * thread-safe (the reference is single-threaded GLUT); different contexts may be driven from different
 * threads/processes (one per GPU).  There is no CPU fallback: without a CUDA device pt_create fails.
 */
#ifndef PT_H
#define PT_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PT_OK          0
#define PT_EINVAL     -1   /* bad argument */
#define PT_ECUDA      -2   /* a CUDA runtime call failed (message has the CUDA error string) */
#define PT_ENOSCENE   -3   /* render/readback before pt_upload_scene */
#define PT_ERANGE     -4   /* index not representable in the reference's float encoding (>= 2^24), M:817-818 */
#define PT_ENOMEM     -5
#define PT_EFORMAT    -6   /* malformed node/triangle array (child or leaf range out of bounds) */

/* which fragment shader of the reference the kernels reproduce */
#define PT_VARIANT_LAMBERT        0   /* OpenglRayTracing/shaders/pass1.fsh   — maxBounce 2 (O:385) */
#define PT_VARIANT_DISNEY_UNIFORM 1   /* DisneyBRDF/shaders/pass1.fsh         — maxBounce 5 (D:502) */
#define PT_VARIANT_DISNEY_IS_MIS  2   /* ImportanceSampling_…/shaders/pass1.fsh — maxBounce 2 (P:861) */

/* arithmetic mode */
#define PT_MODE_FAST          0   /* production: culled traversal of a binned-SAH topology over the reference's leaves
                                     (exact ties re-traced in reference order); CUDA fast math + FMA in shading;
                                     exact (IEEE, un-fused) ray generation and ray/box/triangle tests */
#define PT_MODE_PARITY_REF    1   /* reference layout (48 B nodes, 144 B triangles), reference traversal without
                                     culling, no FMA, include/ptmath.h transcendental functions: bit-identical to
                                     the CPU oracle */
#define PT_MODE_PARITY_WIDE   2   /* the wide-node data layout + culled traversal of the REFERENCE topology under the
                                     parity arithmetic: bit-identical to the oracle */
#define PT_MODE_PARITY_FAST   3   /* PT_MODE_FAST's traversal (binned-SAH topology, tie fallback) under the parity
                                     arithmetic: bit-identical to the oracle too — proves that neither the layout nor
                                     the topology change any result */

#define PT_MODE_FAST_COUNT    4   /* PT_MODE_FAST with traversal work counters compiled in (PtStats.inner_visits …): the
                                     source of the roofline's algorithmic bytes; slower, never the timed path */

/* how a sample is scheduled on the GPU (same arithmetic, same results) */
#define PT_PIPELINE_MEGAKERNEL 0  /* one thread per pixel runs the whole path (ray generation .. accumulation) */
#define PT_PIPELINE_WAVEFRONT  1  /* generate / shade / shadow / extend / accumulate kernels with compacted path queues
                                     in HBM (DESIGN.md "Pipelines"); not available for PT_MODE_PARITY_REF */
#define PT_PIPELINE_PERSISTENT 2  /* persistent warps claim (8x4 tile, frame) items frame-major and accumulate in frame order */

/* how variant 2 turns two random numbers into an environment light sample (SampleHdr, P:573-585) */
#define PT_ENV_TABLE       0   /* the reference's: NEAREST lookup in the precomputed hdrCache table (calculateHdrCache, M:555-652) */
#define PT_ENV_CDF_SEARCH  1   /* SURVEY.md 8f.3, opt-in, NOT the reference's image: binary search of the marginal / conditional CDFs
                                  (the same fp32 tables calculateHdrCache inverts) per sample, on the continuous random numbers
                                  instead of the table's W x H grid of them; needs no hdrCache (33.6 MB) and no light table (67 MB
                                  at 2048x1024).  Variant 2, persistent pipeline, PT_MODE_FAST or PT_MODE_PARITY_FAST. */

typedef struct PtConfig {
    int variant;        /* PT_VARIANT_* */
    int width, height;  /* framebuffer size; pass1.width/height M:82-83, uniforms M:871-872 */
    int max_bounce;     /* <= 0: the variant's default (2 / 5 / 2); variant 2 supports <= 4 (8 Sobol dims, P:92) */
    int device;         /* CUDA device ordinal */
    int mode;           /* PT_MODE_* */
    int pipeline;       /* PT_PIPELINE_* */
    int env_sampler;    /* PT_ENV_* (0 = the reference's) */
} PtConfig;

typedef struct PtStats {
    unsigned long long rays;          /* hitBVH calls (primary + shadow + bounce) since the last reset */
    unsigned long long samples;       /* pixel-samples rendered since the last reset */
    double render_ms;                 /* device time of pt_render_spp launches since the last reset (CUDA events) */
    unsigned long long launches;      /* kernels launched by this context since the last reset */
    int max_stack_depth;              /* deepest traversal stack the uploaded BVH can need */
    int n_inner_nodes;                /* wide nodes built at upload */
    int n_materials;                  /* distinct materials found at upload */
    int fast_tree_depth;              /* depth of the traversal topology built at upload (pt_build.cpp) */
    int n_fast_inner;                 /* its inner-node count */
    double upload_build_ms;           /* host time spent building that topology in the last pt_upload_scene */
    unsigned long long inner_visits;  /* PT_MODE_FAST_COUNT only: inner nodes fetched since the last reset */
    unsigned long long tri_tests;     /*   triangles tested */
    unsigned long long leaf_candidates; /* triangle hits that went through the reference-leaf box test */
    unsigned long long tie_retraces;  /*   rays re-traced on the reference topology because of an exact-distance tie */
    unsigned long long shaded_hits;   /*   closest hits whose normals + material were fetched */
    unsigned long long env_texels;    /*   environment-map and hdrCache texels fetched */
} PtStats;

typedef struct pt_ctx pt_ctx;

/* glutCreateWindow + getShaderProgram + the three RGBA32F attachments (M:746-751, M:862-866): allocates the
   fp32 accumulation framebuffer (zeroed) and the context's stream. */
int pt_create(const PtConfig* cfg, pt_ctx** out);
int pt_destroy(pt_ctx* ctx);

/* glBufferData + glTexBuffer(GL_RGB32F) x2 (M:825-840) and glTexImage2D x2 + hdrResolution (M:846-853).
   tris:  nTriangles x 36 floats, Triangle_encoded (M:50-60).
   nodes: nNodes x 12 floats, BVHNode_encoded (M:68-72); node 0 is the dummy, root is node 1 (M:780-788).
   hdr:   hdrH x hdrW x 3 floats, rows in file order (sampled NEAREST + CLAMP_TO_EDGE, M:185-195).
   hdrCache: same shape, calculateHdrCache output (M:555-652); may be NULL: variants 0 and 1 do not read it, and for
   variant 2 the table is then computed on the device (same arithmetic, same table: pt_calculate_hdr_cache).
   Validates the tree, then repacks on the device into the traversal layout (DESIGN.md "HBM layout"). */
int pt_upload_scene(pt_ctx* ctx, const float* tris, int nTriangles, const float* nodes, int nNodes,
                    const float* hdr, const float* hdrCache, int hdrW, int hdrH);

/* calculateHdrCache (M:555-652) on a CUDA device: hdr and cache are host arrays of height*width*3 floats, cache receives the
   table (x/w, y/h, pdf) bit for bit as the reference's host loop produces it (its sequential fp32 sums are kept sequential,
   the searches run in parallel).  Needs no context.  pt_upload_scene runs the same kernels when variant 2 is handed
   hdrCache = NULL. */
int pt_calculate_hdr_cache(int device, const float* hdr, int width, int height, float* cache);

/* No counterpart in the reference.  pt_calculate_hdr_cache for callers that
   cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): d_hdr and d_cache are DEVICE
   pointers the caller owns (height*width*3 floats each, on `device`), the temporaries live in a static 2.1 MB device array
   inside libpt.so (held by every GPU that loads the library until exit), and the call performs no cudaMalloc / cudaFree / host
   or library allocation.  Limits: width <= 4096 and width * height <= 262,144 texels (512x512, 1024x256 ...); larger maps are
   rejected with PT_ERANGE.  Same table bit for bit.  Asynchronous on `cuda_stream` (a cudaStream_t, NULL = the default stream):
   synchronise that stream before reading d_cache.  All calls on a device share the static array (it keeps the last call's
   contents; nothing in it needs clearing): calls that overlap on the GPU race — serialise them.  The first call on a device
   resolves the array's device address (cudaGetSymbolAddress); make that first call outside graph capture. */
int pt_calculate_hdr_cache_noalloc(int device, const float* d_hdr, int width, int height, float* d_cache, void* cuda_stream);

/* glUniform3fv(eye) + glUniformMatrix4fv(cameraRotate, GL_FALSE) (M:678-679): cameraRotate is column-major. */
int pt_set_camera(pt_ctx* ctx, const float eye[3], const float cameraRotate[16]);

/* Restricts rendering to the pixel rectangle [x0,x1) x [y0,y1) (y = 0 is the bottom row) — the tile split of
   multi-GPU runs.  Default: the whole image.  (No reference equivalent: glViewport is always full, M:125.) */
int pt_set_region(pt_ctx* ctx, int x0, int y0, int x1, int y1);

/* Interleaved tile split in ONE launch: the image rows are cut into bands of 8 rows (one CTA tile row); this
   context renders band b iff b % world == rank.  world = 1 (default) renders every band.  Applies on top of the
   region rectangle.  Bands interleave so that sky rows and geometry rows spread evenly over the GPUs. */
int pt_set_band_interleave(pt_ctx* ctx, int world, int rank);

/* The finer form of the same split, also ONE launch: the region is cut into tiles of 8 x 4 pixels (column tx, 4-row
   strip s, both counted from the region's lower left corner); this context renders tile (tx, s) iff
   (tx + s) % world == rank.  Every rank gets an even sample of the whole image, which balances the ranks better than
   whole bands do (measured at 8 GPUs).  The last call of pt_set_band_interleave /
   pt_set_tile_interleave wins. */
int pt_set_tile_interleave(pt_ctx* ctx, int world, int rank);

/* pass1.draw() + pass2.draw() for frameCounter = first_frame .. first_frame + n_frames - 1 (M:680, M:704-705):
   one sample per pixel per frame, running mean mix(last, colour, 1/(frame+1)) (P:868-869) kept in the fp32
   framebuffer.  Asynchronous on the context's stream. */
int pt_render_spp(pt_ctx* ctx, unsigned first_frame, unsigned n_frames);

/* Sample-split multi-GPU form: renders frameCounter = first_frame + k*stride, k = 0 .. n_frames-1.
   accumulate = PT_ACC_MEAN: the reference's running mean (needs stride 1 and contiguous calls from frame 0);
   accumulate = PT_ACC_SUM: framebuffer += colour, so that per-GPU partial sums can be added (NCCL reduce) and
   divided by the total frame count afterwards (pt_scale_framebuffer). */
#define PT_ACC_MEAN 0
#define PT_ACC_SUM  1
int pt_render_frames(pt_ctx* ctx, unsigned first_frame, unsigned n_frames, unsigned stride, int accumulate);

/* Zeroes the framebuffer (frameCounter = 0 after a camera move, M:718, M:737, makes the first mix() discard it). */
int pt_clear(pt_ctx* ctx);

/* New (the reference never reads pixels back): W*H float4, row 0 = bottom row (GL window origin). Synchronises. */
int pt_readback(pt_ctx* ctx, float* rgba, size_t bytes);
/* pt_readback without the wait: snapshots the framebuffer as of this call (device-to-device, stream-ordered) and copies the
   snapshot to `rgba_pinned` (pinned host memory) on a second stream, so that the next pt_render_spp overlaps the transfer.
   The buffer is valid after pt_synchronize(ctx).  One copy in flight per context: a second call waits (on the device) for
   the first.  No reference equivalent. */
int pt_readback_async(pt_ctx* ctx, float* rgba_pinned, size_t bytes);
/* pass3.draw() (M:706): the display pass pass3.fsh:14-24 — toneMapping(colour, 1.5) — fused with the window's 8-bit
   conversion on the device; W*H RGBA8, row 0 = TOP row (image order), alpha 255.  A quarter of pt_readback's bytes
   cross PCIe.  Synchronises. */
int pt_readback_display(pt_ctx* ctx, unsigned char* rgba8, size_t bytes);
/* pt_readback_display without the wait, for a display() loop (M:659-709) that draws frame k while frame k + 1 renders: the
   display pass runs now (stream-ordered: the image as of this call) into a staging image and the copy to `rgba8_pinned`
   (pinned host memory) rides on a second stream under the next pt_render_spp.  The buffer is valid after
   pt_synchronize(ctx); one copy in flight per context, as for pt_readback_async.  No reference equivalent. */
int pt_readback_display_async(pt_ctx* ctx, unsigned char* rgba8_pinned, size_t bytes);
/* Parity hook: triangle index (into the uploaded array) of the primary hit of the LAST frame rendered, -1 = miss. */
int pt_readback_primary_ids(pt_ctx* ctx, int* ids, size_t bytes);

/* Profiling hook (no reference equivalent): when enabled, every render launch also writes, per pixel, the SM
   cycles (clock64) its thread spent on that pixel in that launch — a cost heat map for finding stragglers.
   Off by default; costs one pointer test per pixel when off. */
int pt_enable_cost_map(pt_ctx* ctx, int on);
int pt_readback_cost_map(pt_ctx* ctx, unsigned* cycles, size_t bytes);   /* W*H uint32, row 0 = bottom. Synchronises. */

/* Multi-GPU plumbing: the device pointer of the framebuffer (W*H float4) and the context's cudaStream_t, so a
   host (torch.distributed / NCCL) can reduce or gather in place without a host round trip. */
int pt_framebuffer_device_ptr(pt_ctx* ctx, void** dptr, size_t* bytes);
int pt_stream(pt_ctx* ctx, void** cuda_stream);
/* Multiplies the framebuffer's rgb by `scale` — the epilogue after a sample-split NCCL sum. */
int pt_scale_framebuffer(pt_ctx* ctx, float scale);

/* ---- multi-GPU: one context per GPU, scene replicated, the per-GPU framebuffers combined on demand ----------
   No reference equivalent (one GPU, no readback).  The contexts may live in one process (one thread per GPU, as ptdrive
   --gpus N does) or in one process per GPU (torchrun).  NCCL is loaded at pt_comm_init with dlopen("libnccl.so.2").
   pt_comm_get_unique_id: rank 0 obtains the 128-byte rendezvous id (ncclGetUniqueId) and hands it to the other ranks by any
   means.  pt_comm_init: collective over the `world` contexts; also maps every peer's framebuffer and rank 0's output
   images into this context's address space (CUDA IPC across processes, peer access inside one) for the fused combine. */
#define PT_COMM_ID_BYTES 128
int pt_comm_get_unique_id(void* id, size_t bytes);
int pt_comm_init(pt_ctx* ctx, int world, int rank, const void* id, size_t bytes);
int pt_comm_info(pt_ctx* ctx, int* world, int* rank, int* peer_memory);    /* peer_memory = 1: the one-kernel NVLink path is active */
int pt_comm_destroy(pt_ctx* ctx);
/* Collective: combines the framebuffers of all ranks into ONE image on rank 0.  flags (the same on every rank):
     PT_COMBINE_SUM       sample split: every rank rendered its frames with PT_ACC_SUM; result rgb = sum / total_frames
     PT_COMBINE_DISJOINT  tile split: every rank rendered its own tiles (pt_set_tile_interleave / band / region) as the
                          running mean; the sum of disjoint regions is exact (x + 0), i.e. bit-identical to one GPU
     | PT_COMBINE_DISPLAY also produce the 8-bit display image (the pass3 tone map, shaders/pass3.fsh:14-24, in the same kernel)
   Alpha = 1 (P:871).  One kernel per GPU: rank r loads slice r of all framebuffers over NVLink, applies the scale (and the
   tone map) and stores into rank 0's image, between two NCCL barriers on the contexts' streams.  Rank 0 may pass rgba_host
   (W*H float4, row 0 = bottom) and/or rgba8_host (W*H RGBA8, row 0 = top; needs PT_COMBINE_DISPLAY) to receive the result,
   and then synchronises; with both NULL the call is asynchronous on every rank.  Each rank's own framebuffer is left
   untouched, so rendering can continue and be combined again.  world = 1 runs the same epilogue on the one framebuffer. */
#define PT_COMBINE_SUM      0
#define PT_COMBINE_DISJOINT 1
#define PT_COMBINE_DISPLAY  4
#define PT_COMBINE_ASYNC    8   /* rank 0's host copy rides on a second stream and overlaps whatever the context renders next; the host
                                   buffers (pinned memory, or the copy is not asynchronous) are valid after pt_synchronize(ctx) */
#define PT_COMBINE_PIPELINED 16 /* the whole combine leaves the render stream (needs peer memory and world > 1; otherwise the flag is ignored
                                   and the call behaves as without it, plus PT_COMBINE_ASYNC): the call snapshots this rank's framebuffer
                                   (device to device, two alternating slots), announces it to the peers with a flag store over NVLink and
                                   returns; a second stream runs ONE small kernel that waits for every rank's announcement inside the kernel
                                   (no NCCL kernel, no host), reduces this rank's slice of the snapshots and announces completion; rank 0
                                   then copies the image to rgba_host / rgba8_host (pinned).  That kernel fits beside the resident render
                                   kernel on every SM, so the combine of step k runs under the rendering of step k + 1 and a rank's render
                                   stream waits for a peer only when that peer is two combines behind.  Same bits as the synchronous
                                   combine.  The host buffers are valid after pt_synchronize(ctx); every rank must pt_synchronize (and the
                                   ranks must meet, e.g. in a barrier of the host's) before any of them is destroyed. */
int pt_combine(pt_ctx* ctx, int flags, unsigned total_frames, float* rgba_host, unsigned char* rgba8_host);
/* Makes the context's render stream wait (on the device, not the host) until every rank's slice of every pipelined combine
   issued so far has landed in rank 0's image.  Collective in the sense that all ranks must have issued those combines. */
int pt_combine_fence(pt_ctx* ctx);

int pt_synchronize(pt_ctx* ctx);
int pt_get_stats(pt_ctx* ctx, PtStats* out);
int pt_reset_stats(pt_ctx* ctx);
/* PT_MODE_FAST_COUNT: log2 histogram of traversal steps (inner nodes + triangles) per ray since the last reset;
   bucket k counts rays with 2^(k-1) <= steps < 2^k (bucket 0: no step).  n >= 33. */
int pt_get_ray_histogram(pt_ctx* ctx, unsigned long long* buckets, int n);

/* ---- BasicRayTracingWithC++/main.cpp ("B:") on the GPU: SURVEY.md 8f.4, BASELINE.json config 1 ----------------------
   The reference's CPU tracer has no entry point but main(); these three calls are its scene literal (B:304-353), its
   sample loop (B:355-432) and its imshow conversion (B:169-190).  Same arithmetic operation for operation; the random
   numbers come from a counter-based stream keyed by (seed, sample, row, column) instead of the reference's shared,
   random_device-seeded mt19937, so runs are reproducible (oracle: the reference's own source with that stream). */
typedef struct PtBasicShape {        /* class Triangle / class Sphere + struct Material, B:48-58, B:77-166 */
    int kind;                        /* 0 = Triangle(p1, p2, p3), 1 = Sphere(O = p1, R = radius) */
    float p1[3], p2[3], p3[3];
    double radius;
    float color[3];
    int isEmissive;
    double specularRate, roughness, refractRate, refractAngle, refractRoughness;
} PtBasicShape;
/* Writes up to `capacity` shapes of the reference's Cornell box; returns how many it has (17). */
int pt_basic_default_scene(PtBasicShape* shapes, int capacity);
/* image: height*width*3 doubles, row 0 = top: the per-pixel SUM over `samples` passes (what imshow receives).
   device_ms (optional): device time of the kernels.  At most 64 shapes (brute-force closest hit, like shoot(), B:192-205). */
int pt_basic_render(int device, const PtBasicShape* shapes, int n_shapes, int width, int height, int samples, unsigned seed,
                    double* image, double* device_ms);
/* imshow without the file: gamma 1/2.2, x255, clamp, truncate -> height*width*3 bytes (feed to pts_write_png). */
int pt_basic_imshow(const double* image, int width, int height, unsigned char* rgb8);
const char* pt_basic_last_error(void);

/* Measurement probe ("L2 peak must be measured on the box"): streams 16-byte ld.global.nc loads from every
   resident thread over a window of window_mb MiB (<= 64: L2-resident on B200, but beyond the reach of the SMs' L1s because
   the slice a CTA reads rotates from pass to pass), `passes` times, and reports the best of 5 launches in GB/s.  hbm_gbs
   (optional): the same kernel over 2 GiB, i.e. the DRAM read rate of the same access pattern.  Needs no context. */
int pt_measure_l2_read_gbs(int device, int window_mb, int passes, double* gbs, double* hbm_gbs);

/* sha256 (16 hex digits) over the SASS of the render kernels inside this library, computed at build time.  An ncu capture
   (profiles/traffic.json) names the hash of the build it was taken from; bench.py quotes the capture only on a match. */
const char* pt_kernel_sass_hash(void);

const char* pt_last_error(void);
/* "sm_100a" build tag + whether a CUDA device is visible; never fails (used by the CPU-only symbol test) */
const char* pt_build_info(void);

#ifdef __cplusplus
}
#endif
#endif /* PT_H */
