// pt_comm.cu — the multi-GPU combine behind the C ABI (include/pt.h: pt_comm_*, pt_combine).
//
// the scene is replicated, image tiles or sample indices are split over the GPUs, and the only exchange is
// the combine of the per-GPU fp32 framebuffers when an image is wanted.  The reference has one GPU and no such step
// (its display() loop, IS/main.cpp:659-709, draws what pass1 accumulated); this file is what lets the C++ caller of that
// loop (ptdrive, or the reference's own main() re-hosted) use N GPUs without Python.
//
// Shape: a reduce-scatter fused with its consumer, over NVLink peer memory, in ONE kernel per GPU:
//   * at pt_comm_init every rank learns a device pointer to every peer's framebuffer and to rank 0's output images
//     (cudaIpc handles exchanged with one ncclAllGather; plain pointers + cudaDeviceEnablePeerAccess inside one process);
//   * pt_combine: [NCCL barrier: every rank's render kernel is complete]  ->  k_combine_peer on every GPU: rank r owns
//     pixels [r n/W, (r+1) n/W); it LOADS that slice from all W framebuffers through NVLink (16-byte loads, fixed rank
//     order: the sum is deterministic), applies what the consumer needs — rgb x 1/total_frames for a sample split, alpha
//     = 1 (P:871), and for display bytes the pass3 tone map (shaders/pass3.fsh:14-24) — and STORES the result straight
//     into rank 0's image  ->  [NCCL barrier: all slices have landed]  ->  rank 0 copies to the host.
//     Per GPU: (W-1)/W of one framebuffer slice set in, one slice out — 29 MB in / 4 MB out at 1080p on 8 GPUs, instead
//     of 232 MB into rank 0 (reduce-to-root) followed by a separate scale pass.
//   * no peer memory (IPC refused, no P2P): ncclReduce to rank 0 + the same consume kernel there.
// NCCL is resolved with dlopen("libnccl.so.2") at pt_comm_init, so that single-GPU users of libpt.so do not need it and a
// host that already loaded NCCL (torch) shares its copy.
#include <dlfcn.h>
#include <nccl.h>
#include <unistd.h>

#include <cstring>
#include <string>
#include <vector>

#include "pt_internal.h"

namespace {

using namespace pt;

#define PT_COMM_MAX_WORLD 16

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string why;
};

NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.handle ? &api : nullptr;
    tried = true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
    }
    if (!api.handle) { api.why = std::string("dlopen(libnccl.so.2): ") + dlerror(); return nullptr; }
#define PT_SYM(field, name)                                                              \
    *(void**)(&api.field) = dlsym(api.handle, name);                                     \
    if (!api.field) { api.why = std::string("libnccl lacks ") + name; api.handle = nullptr; return nullptr; }
    PT_SYM(GetUniqueId, "ncclGetUniqueId")
    PT_SYM(CommInitRank, "ncclCommInitRank")
    PT_SYM(CommDestroy, "ncclCommDestroy")
    PT_SYM(AllGather, "ncclAllGather")
    PT_SYM(AllReduce, "ncclAllReduce")
    PT_SYM(Reduce, "ncclReduce")
    PT_SYM(GetErrorString, "ncclGetErrorString")
#undef PT_SYM
    return &api;
}

// what every rank tells the others at pt_comm_init
struct PeerInfo {
    cudaIpcMemHandle_t h_fb, h_f4, h_u8;
    unsigned long long p_fb, p_f4, p_u8;     // the same allocations as plain pointers (valid inside the exporting process)
    long long pid;
    int device, ok;
};

struct PeerTable {
    const float4* fb[PT_COMM_MAX_WORLD];
};

// pass3.fsh:14-24 on one pixel (the arithmetic of k_display in pt_api.cu, strict: identical bytes)
__device__ __forceinline__ uchar4 tone_map_u8(float4 c) {
    const float lum = s_add(s_add(s_mul(0.3f, c.x), s_mul(0.6f, c.y)), s_mul(0.1f, c.z));
    const float k = s_add(1.0f, s_div(lum, 1.5f));
    const float v[3] = {s_div(s_mul(c.x, 1.0f), k), s_div(s_mul(c.y, 1.0f), k), s_div(s_mul(c.z, 1.0f), k)};
    unsigned char q[3];
    for (int i = 0; i < 3; i++) {
        const float t = fminf(fmaxf(v[i], 0.0f), 1.0f);
        q[i] = (unsigned char)(int)s_add(s_mul(t, 255.0f), 0.5f);
    }
    return make_uchar4(q[0], q[1], q[2], 255);
}

// Pixels [begin, end) of the image: sum of the W framebuffers (rank order), consumer epilogue, stores into the output images
// (rank 0's, through peer pointers).  world == 1 with fb[0] = an already reduced buffer is the no-peer fallback's epilogue.
__global__ void __launch_bounds__(256) k_combine(const PeerTable t, int world, size_t begin, size_t end, float scale, int scale_rgb,
                                                 float4* __restrict__ out_f4, uchar4* __restrict__ out_u8, int width, int height) {
    for (size_t i = begin + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += (size_t)gridDim.x * blockDim.x) {
        float4 s = t.fb[0][i];
        for (int k = 1; k < world; k++) {
            const float4 v = t.fb[k][i];
            s.x = s_add(s.x, v.x); s.y = s_add(s.y, v.y); s.z = s_add(s.z, v.z);
        }
        if (scale_rgb) { s.x = s_mul(s.x, scale); s.y = s_mul(s.y, scale); s.z = s_mul(s.z, scale); }
        s.w = 1.0f;                                                       // P:871
        if (out_f4) out_f4[i] = s;
        if (out_u8) {
            const size_t y = i / (size_t)width, x = i - y * (size_t)width;
            out_u8[((size_t)height - 1 - y) * (size_t)width + x] = tone_map_u8(s);   // display rows run top to bottom
        }
    }
}

}  // namespace

struct PtComm {
    int world = 1, rank = 0;
    NcclApi* api = nullptr;
    ncclComm_t nccl = nullptr;
    bool peer = false;                       // every rank mapped every peer's buffers
    PeerTable table;                         // framebuffers of all ranks (own entry = own pointer)
    float4* out_f4 = nullptr;                // rank 0's combined image as THIS rank addresses it
    uchar4* out_u8 = nullptr;
    float4* my_f4 = nullptr;                 // this rank's allocations (rank 0's are the outputs)
    uchar4* my_u8 = nullptr;
    int* token = nullptr;                    // barrier payload
    float4* reduced = nullptr;               // fallback: ncclReduce target on rank 0
    std::vector<void*> opened;               // cudaIpcOpenMemHandle results
};

#define NCK(call)                                                                                          \
    do {                                                                                                   \
        ncclResult_t r_ = (call);                                                                          \
        if (r_ != ncclSuccess) return fail(PT_ECUDA, std::string(#call) + ": " + m->api->GetErrorString(r_)); \
    } while (0)

void pt_comm_release(pt_ctx* c) {
    PtComm* m = c->comm;
    if (!m) return;
    cudaSetDevice(c->cfg.device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (void* p : m->opened) cudaIpcCloseMemHandle(p);
    if (m->nccl && m->api) m->api->CommDestroy(m->nccl);
    cudaFree(m->my_f4); cudaFree(m->my_u8); cudaFree(m->token); cudaFree(m->reduced);
    delete m;
    c->comm = nullptr;
}

extern "C" {

int pt_comm_get_unique_id(void* id, size_t bytes) {
    if (!id || bytes < sizeof(ncclUniqueId)) return fail(PT_EINVAL, "pt_comm_get_unique_id: need a 128-byte buffer");
    NcclApi* api = nccl_api();
    if (!api) return fail(PT_ECUDA, "pt_comm_get_unique_id: NCCL is not available");
    ncclUniqueId u;
    ncclResult_t r = api->GetUniqueId(&u);
    if (r != ncclSuccess) return fail(PT_ECUDA, std::string("ncclGetUniqueId: ") + api->GetErrorString(r));
    std::memcpy(id, &u, sizeof u);
    return PT_OK;
}

int pt_comm_init(pt_ctx* c, int world, int rank, const void* id, size_t bytes) {
    if (!c || !id) return fail(PT_EINVAL, "pt_comm_init: null argument");
    if (world < 1 || world > PT_COMM_MAX_WORLD || rank < 0 || rank >= world) return fail(PT_EINVAL, "pt_comm_init: need 0 <= rank < world <= 16");
    if (bytes < sizeof(ncclUniqueId)) return fail(PT_EINVAL, "pt_comm_init: the id is 128 bytes (pt_comm_get_unique_id)");
    if (c->comm) return fail(PT_EINVAL, "pt_comm_init: this context already has a communicator");
    NcclApi* api = nccl_api();
    if (!api) return fail(PT_ECUDA, "pt_comm_init: NCCL is not available (libnccl.so.2 could not be loaded)");
    CK(cudaSetDevice(c->cfg.device));
    PtComm* m = new (std::nothrow) PtComm();
    if (!m) return fail(PT_ENOMEM, "pt_comm_init: out of host memory");
    m->world = world; m->rank = rank; m->api = api;
    c->comm = m;
    const size_t npx = (size_t)c->cfg.width * c->cfg.height;
    ncclUniqueId u;
    std::memcpy(&u, id, sizeof u);
    NCK(api->CommInitRank(&m->nccl, world, u, rank));
    CK(cudaMalloc(&m->token, sizeof(int)));
    CK(cudaMemset(m->token, 0, sizeof(int)));
    CK(cudaMalloc(&m->my_f4, npx * sizeof(float4)));
    CK(cudaMalloc(&m->my_u8, npx * sizeof(uchar4)));
    // exchange: who am I, where are my buffers
    PeerInfo mine;
    std::memset(&mine, 0, sizeof mine);
    mine.pid = (long long)getpid(); mine.device = c->cfg.device;
    mine.p_fb = (unsigned long long)c->accum; mine.p_f4 = (unsigned long long)m->my_f4; mine.p_u8 = (unsigned long long)m->my_u8;
    mine.ok = cudaIpcGetMemHandle(&mine.h_fb, c->accum) == cudaSuccess && cudaIpcGetMemHandle(&mine.h_f4, m->my_f4) == cudaSuccess &&
              cudaIpcGetMemHandle(&mine.h_u8, m->my_u8) == cudaSuccess;
    cudaGetLastError();
    PeerInfo* d_all = nullptr;
    CK(cudaMalloc(&d_all, sizeof(PeerInfo) * (size_t)world));
    CK(cudaMemcpy(d_all + rank, &mine, sizeof mine, cudaMemcpyHostToDevice));
    NCK(api->AllGather(d_all + rank, d_all, sizeof(PeerInfo), ncclChar, m->nccl, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    std::vector<PeerInfo> all((size_t)world);
    CK(cudaMemcpy(all.data(), d_all, sizeof(PeerInfo) * (size_t)world, cudaMemcpyDeviceToHost));
    cudaFree(d_all);
    // map every peer's framebuffer and rank 0's outputs
    int ok = 1;
    if (std::getenv("PT_COMM_NO_PEER")) ok = 0;
    for (int k = 0; k < world && ok; k++) {
        const PeerInfo& pi = all[(size_t)k];
        if (k == rank) { m->table.fb[k] = c->accum; if (k == 0) { m->out_f4 = m->my_f4; m->out_u8 = m->my_u8; } continue; }
        if (pi.pid == mine.pid) {                         // another context of this process: plain pointers
            if (pi.device != c->cfg.device) {
                int can = 0;
                if (cudaDeviceCanAccessPeer(&can, c->cfg.device, pi.device) != cudaSuccess || !can) { ok = 0; break; }
                const cudaError_t e = cudaDeviceEnablePeerAccess(pi.device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { ok = 0; break; }
                cudaGetLastError();
            }
            m->table.fb[k] = (const float4*)pi.p_fb;
            if (k == 0) { m->out_f4 = (float4*)pi.p_f4; m->out_u8 = (uchar4*)pi.p_u8; }
        } else {                                          // another process: CUDA IPC
            if (!pi.ok) { ok = 0; break; }
            void* p = nullptr;
            if (cudaIpcOpenMemHandle(&p, pi.h_fb, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; break; }
            m->opened.push_back(p); m->table.fb[k] = (const float4*)p;
            if (k == 0) {
                void *q = nullptr, *r2 = nullptr;
                if (cudaIpcOpenMemHandle(&q, pi.h_f4, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; break; }
                m->opened.push_back(q);
                if (cudaIpcOpenMemHandle(&r2, pi.h_u8, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; break; }
                m->opened.push_back(r2);
                m->out_f4 = (float4*)q; m->out_u8 = (uchar4*)r2;
            }
        }
    }
    cudaGetLastError();
    // all ranks must agree on the path: min over ranks of `ok`
    CK(cudaMemcpy(m->token, &ok, sizeof ok, cudaMemcpyHostToDevice));
    NCK(api->AllReduce(m->token, m->token, 1, ncclInt, ncclMin, m->nccl, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaMemcpy(&ok, m->token, sizeof ok, cudaMemcpyDeviceToHost));
    m->peer = ok != 0;
    if (!m->peer) {
        for (void* p : m->opened) cudaIpcCloseMemHandle(p);
        m->opened.clear();
        m->table.fb[0] = nullptr;
        m->out_f4 = m->my_f4; m->out_u8 = m->my_u8;       // rank 0 consumes locally after ncclReduce
        if (rank == 0) CK(cudaMalloc(&m->reduced, npx * sizeof(float4)));
    }
    return PT_OK;
}

int pt_comm_info(pt_ctx* c, int* world, int* rank, int* peer_memory) {
    if (!c) return fail(PT_EINVAL, "pt_comm_info: null context");
    if (world) *world = c->comm ? c->comm->world : 1;
    if (rank) *rank = c->comm ? c->comm->rank : 0;
    if (peer_memory) *peer_memory = c->comm && c->comm->peer ? 1 : 0;
    return PT_OK;
}

int pt_combine(pt_ctx* c, int flags, unsigned total_frames, float* rgba_host, unsigned char* rgba8_host) {
    if (!c) return fail(PT_EINVAL, "pt_combine: null context");
    const int kind = flags & 3;
    const bool display = (flags & PT_COMBINE_DISPLAY) != 0, async = (flags & PT_COMBINE_ASYNC) != 0;
    if ((flags & ~(3 | PT_COMBINE_DISPLAY | PT_COMBINE_ASYNC)) || (kind != PT_COMBINE_SUM && kind != PT_COMBINE_DISJOINT))
        return fail(PT_EINVAL, "pt_combine: flags = PT_COMBINE_SUM or PT_COMBINE_DISJOINT, optionally | PT_COMBINE_DISPLAY | PT_COMBINE_ASYNC");
    if (kind == PT_COMBINE_SUM && total_frames == 0) return fail(PT_EINVAL, "pt_combine: total_frames must be > 0 for PT_COMBINE_SUM");
    if (rgba8_host && !display) return fail(PT_EINVAL, "pt_combine: rgba8_host needs PT_COMBINE_DISPLAY (on every rank)");
    CK(cudaSetDevice(c->cfg.device));
    const size_t npx = (size_t)c->cfg.width * c->cfg.height;
    const int W = c->cfg.width, H = c->cfg.height;
    const float scale = kind == PT_COMBINE_SUM ? 1.0f / (float)total_frames : 1.0f;
    const int scale_rgb = kind == PT_COMBINE_SUM ? 1 : 0;
    PtComm* m = c->comm;
    if (!m) return fail(PT_EINVAL, "pt_combine: pt_comm_init first (world = 1 is allowed)");
    if (m->rank != 0 && (rgba_host || rgba8_host)) return fail(PT_EINVAL, "pt_combine: only rank 0 receives the image");
    const int grid = (c->num_sms > 0 ? c->num_sms : 148) * 4;
    // an asynchronous host copy of the PREVIOUS combined image may still read rank 0's output images: the kernels that
    // overwrite them (on every rank) come after barrier A, which rank 0 joins only once that copy is done
    if (m->rank == 0 && c->copy_pending) CK(cudaStreamWaitEvent(c->stream, c->ev_copied, 0));
    if (m->world == 1) {
        // one GPU: the same consumer epilogue over the context's own framebuffer
        PeerTable t; std::memset(&t, 0, sizeof t); t.fb[0] = c->accum;
        k_combine<<<grid, 256, 0, c->stream>>>(t, 1, 0, npx, scale, scale_rgb, m->my_f4, display ? m->my_u8 : nullptr, W, H);
        CK(cudaGetLastError());
        c->launches += 1;
    } else if (m->peer) {
        // barrier A: every rank's render kernel is complete (stream order on every rank + the collective)
        NCK(m->api->AllReduce(m->token, m->token, 1, ncclInt, ncclMax, m->nccl, c->stream));
        const size_t begin = npx * (size_t)m->rank / (size_t)m->world, end = npx * (size_t)(m->rank + 1) / (size_t)m->world;
        k_combine<<<grid, 256, 0, c->stream>>>(m->table, m->world, begin, end, scale, scale_rgb, m->out_f4, display ? m->out_u8 : nullptr, W, H);
        CK(cudaGetLastError());
        c->launches += 1;
        // barrier B: all slices have landed in rank 0's images, and nobody reads a framebuffer any more
        NCK(m->api->AllReduce(m->token, m->token, 1, ncclInt, ncclMax, m->nccl, c->stream));
    } else {
        NCK(m->api->Reduce(c->accum, m->reduced, npx * 4, ncclFloat, ncclSum, 0, m->nccl, c->stream));
        if (m->rank == 0) {
            PeerTable t; std::memset(&t, 0, sizeof t); t.fb[0] = m->reduced;
            k_combine<<<grid, 256, 0, c->stream>>>(t, 1, 0, npx, scale, scale_rgb, m->my_f4, display ? m->my_u8 : nullptr, W, H);
            CK(cudaGetLastError());
            c->launches += 1;
        }
    }
    if (m->rank == 0 && (rgba_host || rgba8_host)) {
        if (async) {
            // the copy rides on the context's copy stream and overlaps the next step's rendering; pt_synchronize waits for it
            int rc = pt_copy_setup(c);
            if (rc != PT_OK) return rc;
            bool first = true;
            if (rgba_host) { rc = pt_copy_enqueue(c, rgba_host, m->my_f4, npx * sizeof(float4), first); first = false; if (rc != PT_OK) return rc; }
            if (rgba8_host) { rc = pt_copy_enqueue(c, rgba8_host, m->my_u8, npx * 4, first); if (rc != PT_OK) return rc; }
            return PT_OK;
        }
        if (rgba_host) CK(cudaMemcpyAsync(rgba_host, m->my_f4, npx * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
        if (rgba8_host) CK(cudaMemcpyAsync(rgba8_host, m->my_u8, npx * 4, cudaMemcpyDeviceToHost, c->stream));
        CK(cudaStreamSynchronize(c->stream));
    }
    return PT_OK;
}

int pt_comm_destroy(pt_ctx* c) {
    if (!c) return fail(PT_EINVAL, "pt_comm_destroy: null context");
    pt_comm_release(c);
    return PT_OK;
}

}  // extern "C"
