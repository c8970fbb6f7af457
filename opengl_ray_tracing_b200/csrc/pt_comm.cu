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
//   * PT_COMBINE_PIPELINED (peer memory only): the same reduction OFF the render stream.  pt_combine snapshots the rank's
//     framebuffer (device-to-device, into one of two slots), announces it with a flag store into every peer's flag array,
//     and returns; a second stream runs k_combine_pipe, which waits INSIDE the kernel for all ranks' announcements (local
//     polls of flags the peers wrote over NVLink: no NCCL kernel, no host), reduces its slice of the SNAPSHOTS and announces
//     "done".  The kernel is 128 threads x <= 32 registers per CTA, one CTA per SM: exactly the 4096 registers the persistent
//     render kernel (5 x 256 threads x 48 registers) leaves free on an SM, so the combine of step k runs UNDER the render
//     of step k + 1 and no rank's render stream ever waits for a peer that is less than two steps behind.
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
    cudaIpcMemHandle_t h_fb, h_f4, h_u8, h_snap, h_flags;
    unsigned long long p_fb, p_f4, p_u8, p_snap, p_flags;     // the same allocations as plain pointers (valid inside the exporting process)
    long long pid;
    int device, ok;
};

struct PeerTable {
    const float4* fb[PT_COMM_MAX_WORLD];
};
// PT_COMBINE_PIPELINED: every rank's flag array as THIS rank addresses it.  Layout of one array (uint32):
//   [PT_FLAG_READY + k]  rank k's snapshot number `seq` is complete          (written by rank k, into every rank's array)
//   [PT_FLAG_DONE + k]   rank k has reduced and stored its slice of `seq`    (likewise)
//   [PT_FLAG_COUNT]      CTAs of this rank's k_combine_pipe that have finished (local)
//   [PT_FLAG_ERROR]      a wait timed out (local; read by pt_synchronize)
struct FlagTable {
    uint32_t* f[PT_COMM_MAX_WORLD];
};
#define PT_FLAG_READY 0
#define PT_FLAG_DONE 16
#define PT_FLAG_COUNT 32
#define PT_FLAG_ERROR 33
#define PT_FLAG_WORDS 64
#define PT_PIPE_TIMEOUT_NS 30000000000ull     // a peer that does not show up within 30 s: flag the error instead of spinning for ever

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// spins until flag >= seq (sequence numbers only grow; wrap-safe comparison)
__device__ __forceinline__ void wait_flag(const uint32_t* flag, uint32_t seq, uint32_t* err) {
    if ((int32_t)(ld_acquire_sys(flag) - seq) >= 0) return;
    const unsigned long long t0 = global_timer_ns();
    while ((int32_t)(ld_acquire_sys(flag) - seq) < 0) {
        __nanosleep(256);
        if (global_timer_ns() - t0 > PT_PIPE_TIMEOUT_NS) { *err = 1u; return; }
    }
}
// lane k tells rank k: `which`[rank] = seq (everything this stream did before is visible first)
__global__ void k_signal(const __grid_constant__ FlagTable ft, int world, int rank, int which, uint32_t seq) {
    const int k = (int)threadIdx.x;
    if (k < world) { __threadfence_system(); st_release_sys(ft.f[k] + which + rank, seq); }
}
// lane k waits for `which`[k] >= seq in this rank's own array
__global__ void k_wait(uint32_t* mine, int world, int which, uint32_t seq) {
    const int k = (int)threadIdx.x;
    if (k < world) wait_flag(mine + which + k, seq, mine + PT_FLAG_ERROR);
}

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

// The pipelined form of k_combine (see the file comment): waits in the kernel for every rank's snapshot `seq`, reduces pixels
// [begin, end) of the snapshots in slot `slot` with k_combine's arithmetic in k_combine's order (identical bits), stores into
// slot `slot` of rank 0's output images and, as the last CTA to finish, announces DONE[rank] = seq to every rank.
// 128 threads x <= 32 registers: co-resident with the persistent render kernel.
__global__ void __launch_bounds__(128, 16) k_combine_pipe(const __grid_constant__ PeerTable snaps, const __grid_constant__ FlagTable ft, int world, int rank, uint32_t seq,
                                                          size_t slot_off, size_t begin, size_t end, float scale, int scale_rgb,
                                                          float4* __restrict__ out_f4, uchar4* __restrict__ out_u8, int width, int height) {
    uint32_t* const mine = ft.f[rank];
    if (threadIdx.x < (unsigned)world) wait_flag(mine + PT_FLAG_READY + threadIdx.x, seq, mine + PT_FLAG_ERROR);
    __syncthreads();
    for (size_t i = begin + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += (size_t)gridDim.x * blockDim.x) {
        float4 s = __ldcv(snaps.fb[0] + slot_off + i);                     // .cv: a line of this slot from two steps ago must not be served again
        for (int k = 1; k < world; k++) {
            const float4 v = __ldcv(snaps.fb[k] + slot_off + i);
            s.x = s_add(s.x, v.x); s.y = s_add(s.y, v.y); s.z = s_add(s.z, v.z);
        }
        if (scale_rgb) { s.x = s_mul(s.x, scale); s.y = s_mul(s.y, scale); s.z = s_mul(s.z, scale); }
        s.w = 1.0f;                                                       // P:871
        if (out_f4) out_f4[slot_off + i] = s;
        if (out_u8) {
            const size_t y = i / (size_t)width, x = i - y * (size_t)width;
            out_u8[slot_off + ((size_t)height - 1 - y) * (size_t)width + x] = tone_map_u8(s);
        }
    }
    __threadfence_system();                                               // this thread's stores (rank 0's memory, over NVLink) before the count
    __syncthreads();
    __shared__ int s_last;
    if (threadIdx.x == 0) s_last = atomicAdd(mine + PT_FLAG_COUNT, 1u) == gridDim.x - 1;
    __syncthreads();
    if (s_last && threadIdx.x < (unsigned)world) {
        if (threadIdx.x == 0) mine[PT_FLAG_COUNT] = 0u;                   // stream order: the next k_combine_pipe of this rank starts after this one
        __threadfence_system();
        st_release_sys(ft.f[threadIdx.x] + PT_FLAG_DONE + rank, seq);
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
    // PT_COMBINE_PIPELINED: two snapshot slots per rank, two output slots on rank 0 (my_f4 / my_u8 hold 2 images), flags
    float4* snap = nullptr;
    uint32_t* flags = nullptr;
    PeerTable snaps;                         // every rank's snapshot allocation
    FlagTable ftab;                          // every rank's flag array
    uint32_t seq = 0;                        // pipelined combines issued so far (the same on every rank)
    cudaEvent_t ev_snap = nullptr, ev_pipe = nullptr, ev_out_copied[2] = {nullptr, nullptr};
    bool out_copy_pending[2] = {false, false};
    bool pipe_pending = false;               // the copy stream holds pipelined work the render stream has not waited for
};

#define NCK(call)                                                                                          \
    do {                                                                                                   \
        ncclResult_t r_ = (call);                                                                          \
        if (r_ != ncclSuccess) return fail(PT_ECUDA, std::string(#call) + ": " + m->api->GetErrorString(r_)); \
    } while (0)

// pt_synchronize: did a wait inside a pipelined combine give up (a peer that never announced its snapshot)?
int pt_comm_check(pt_ctx* c) {
    PtComm* m = c->comm;
    if (!m || m->seq == 0) return PT_OK;
    uint32_t err = 0;
    CK(cudaMemcpy(&err, m->flags + PT_FLAG_ERROR, sizeof err, cudaMemcpyDeviceToHost));
    if (err) {
        cudaMemset(m->flags + PT_FLAG_ERROR, 0, sizeof err);
        return fail(PT_ECUDA, "pt_combine | PT_COMBINE_PIPELINED: a rank did not announce its framebuffer within 30 s; the combined image is incomplete");
    }
    return PT_OK;
}

void pt_comm_release(pt_ctx* c) {
    PtComm* m = c->comm;
    if (!m) return;
    cudaSetDevice(c->cfg.device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->copy_stream) cudaStreamSynchronize(c->copy_stream);
    for (void* p : m->opened) cudaIpcCloseMemHandle(p);
    if (m->nccl && m->api) m->api->CommDestroy(m->nccl);
    cudaFree(m->my_f4); cudaFree(m->my_u8); cudaFree(m->token); cudaFree(m->reduced); cudaFree(m->snap); cudaFree(m->flags);
    if (m->ev_snap) cudaEventDestroy(m->ev_snap);
    if (m->ev_pipe) cudaEventDestroy(m->ev_pipe);
    for (cudaEvent_t e : m->ev_out_copied) if (e) cudaEventDestroy(e);
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
    CK(cudaMalloc(&m->my_f4, 2 * npx * sizeof(float4)));               // two slots: the pipelined combine alternates, everything else uses slot 0
    CK(cudaMalloc(&m->my_u8, 2 * npx * sizeof(uchar4)));
    CK(cudaMalloc(&m->snap, 2 * npx * sizeof(float4)));
    CK(cudaMalloc(&m->flags, PT_FLAG_WORDS * sizeof(uint32_t)));
    CK(cudaMemset(m->flags, 0, PT_FLAG_WORDS * sizeof(uint32_t)));
    CK(cudaEventCreateWithFlags(&m->ev_snap, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&m->ev_pipe, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&m->ev_out_copied[0], cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&m->ev_out_copied[1], cudaEventDisableTiming));
    std::memset(&m->snaps, 0, sizeof m->snaps);
    std::memset(&m->ftab, 0, sizeof m->ftab);
    // exchange: who am I, where are my buffers
    PeerInfo mine;
    std::memset(&mine, 0, sizeof mine);
    mine.pid = (long long)getpid(); mine.device = c->cfg.device;
    mine.p_fb = (unsigned long long)c->accum; mine.p_f4 = (unsigned long long)m->my_f4; mine.p_u8 = (unsigned long long)m->my_u8;
    mine.p_snap = (unsigned long long)m->snap; mine.p_flags = (unsigned long long)m->flags;
    mine.ok = cudaIpcGetMemHandle(&mine.h_fb, c->accum) == cudaSuccess && cudaIpcGetMemHandle(&mine.h_f4, m->my_f4) == cudaSuccess &&
              cudaIpcGetMemHandle(&mine.h_u8, m->my_u8) == cudaSuccess && cudaIpcGetMemHandle(&mine.h_snap, m->snap) == cudaSuccess &&
              cudaIpcGetMemHandle(&mine.h_flags, m->flags) == cudaSuccess;
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
        if (k == rank) {
            m->table.fb[k] = c->accum; m->snaps.fb[k] = m->snap; m->ftab.f[k] = m->flags;
            if (k == 0) { m->out_f4 = m->my_f4; m->out_u8 = m->my_u8; }
            continue;
        }
        if (pi.pid == mine.pid) {                         // another context of this process: plain pointers
            if (pi.device != c->cfg.device) {
                int can = 0;
                if (cudaDeviceCanAccessPeer(&can, c->cfg.device, pi.device) != cudaSuccess || !can) { ok = 0; break; }
                const cudaError_t e = cudaDeviceEnablePeerAccess(pi.device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { ok = 0; break; }
                cudaGetLastError();
            }
            m->table.fb[k] = (const float4*)pi.p_fb; m->snaps.fb[k] = (const float4*)pi.p_snap; m->ftab.f[k] = (uint32_t*)pi.p_flags;
            if (k == 0) { m->out_f4 = (float4*)pi.p_f4; m->out_u8 = (uchar4*)pi.p_u8; }
        } else {                                          // another process: CUDA IPC
            if (!pi.ok) { ok = 0; break; }
            void* p = nullptr;
            if (cudaIpcOpenMemHandle(&p, pi.h_fb, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; break; }
            m->opened.push_back(p); m->table.fb[k] = (const float4*)p;
            void *ps = nullptr, *pf = nullptr;
            if (cudaIpcOpenMemHandle(&ps, pi.h_snap, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; break; }
            m->opened.push_back(ps); m->snaps.fb[k] = (const float4*)ps;
            if (cudaIpcOpenMemHandle(&pf, pi.h_flags, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; break; }
            m->opened.push_back(pf); m->ftab.f[k] = (uint32_t*)pf;
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
    const bool display = (flags & PT_COMBINE_DISPLAY) != 0, async = (flags & (PT_COMBINE_ASYNC | PT_COMBINE_PIPELINED)) != 0;
    if ((flags & ~(3 | PT_COMBINE_DISPLAY | PT_COMBINE_ASYNC | PT_COMBINE_PIPELINED)) || (kind != PT_COMBINE_SUM && kind != PT_COMBINE_DISJOINT))
        return fail(PT_EINVAL, "pt_combine: flags = PT_COMBINE_SUM or PT_COMBINE_DISJOINT, optionally | PT_COMBINE_DISPLAY | PT_COMBINE_ASYNC | PT_COMBINE_PIPELINED");
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
    if ((flags & PT_COMBINE_PIPELINED) && m->world > 1 && m->peer) {
        // ---- the combine off the render stream (file comment).  Render stream: [wait: every rank is done with this slot's
        // previous snapshot] -> snapshot -> READY to all ranks.  Copy stream: k_combine_pipe (waits for all READYs in the kernel)
        // -> rank 0: [wait: all DONEs] -> host copy.
        int rc = pt_copy_setup(c);
        if (rc != PT_OK) return rc;
        const uint32_t seq = ++m->seq;
        const int slot = (int)(seq & 1u);
        const size_t slot_off = (size_t)slot * npx;
        if (seq >= 3) k_wait<<<1, 32, 0, c->stream>>>(m->flags, m->world, PT_FLAG_DONE, seq - 2);
        CK(cudaMemcpyAsync(m->snap + slot_off, c->accum, npx * sizeof(float4), cudaMemcpyDeviceToDevice, c->stream));
        // rank 0's host copy of this output slot (combine seq - 2) must be over before any rank stores into it again: the
        // peers' kernels wait for rank 0's READY, which is only given after that copy
        if (m->rank == 0 && m->out_copy_pending[slot]) CK(cudaStreamWaitEvent(c->stream, m->ev_out_copied[slot], 0));
        k_signal<<<1, 32, 0, c->stream>>>(m->ftab, m->world, m->rank, PT_FLAG_READY, seq);
        CK(cudaEventRecord(m->ev_snap, c->stream));
        CK(cudaStreamWaitEvent(c->copy_stream, m->ev_snap, 0));
        const size_t begin = npx * (size_t)m->rank / (size_t)m->world, end = npx * (size_t)(m->rank + 1) / (size_t)m->world;
        k_combine_pipe<<<c->num_sms > 0 ? c->num_sms : 148, 128, 0, c->copy_stream>>>(m->snaps, m->ftab, m->world, m->rank, seq, slot_off, begin, end,
                                                                                   scale, scale_rgb, m->out_f4, display ? m->out_u8 : nullptr, W, H);
        CK(cudaGetLastError());
        c->launches += seq >= 3 ? 3 : 2;
        if (m->rank == 0 && (rgba_host || rgba8_host)) {
            k_wait<<<1, 32, 0, c->copy_stream>>>(m->flags, m->world, PT_FLAG_DONE, seq);      // every slice has landed
            if (rgba_host) CK(cudaMemcpyAsync(rgba_host, m->my_f4 + slot_off, npx * sizeof(float4), cudaMemcpyDeviceToHost, c->copy_stream));
            if (rgba8_host) CK(cudaMemcpyAsync(rgba8_host, m->my_u8 + slot_off, npx * 4, cudaMemcpyDeviceToHost, c->copy_stream));
            CK(cudaEventRecord(m->ev_out_copied[slot], c->copy_stream));
            m->out_copy_pending[slot] = true;
            c->launches += 1;
        }
        CK(cudaGetLastError());
        c->copy_pending = true;                                           // pt_synchronize also drains the copy stream
        m->pipe_pending = true;
        return PT_OK;
    }
    if (m->pipe_pending) {
        // a synchronous combine after pipelined ones: this rank's pipelined work first (every rank does the same before it
        // joins barrier A, so nobody still stores into rank 0's slot 0 or reads it for a host copy)
        CK(cudaEventRecord(m->ev_pipe, c->copy_stream));
        CK(cudaStreamWaitEvent(c->stream, m->ev_pipe, 0));
        m->pipe_pending = false;
    }
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

int pt_combine_fence(pt_ctx* c) {
    if (!c) return fail(PT_EINVAL, "pt_combine_fence: null context");
    PtComm* m = c->comm;
    if (!m || !m->pipe_pending) return PT_OK;
    CK(cudaSetDevice(c->cfg.device));
    k_wait<<<1, 32, 0, c->copy_stream>>>(m->flags, m->world, PT_FLAG_DONE, m->seq);       // every rank's slice of the latest combine has landed
    CK(cudaGetLastError());
    CK(cudaEventRecord(m->ev_pipe, c->copy_stream));
    CK(cudaStreamWaitEvent(c->stream, m->ev_pipe, 0));
    c->launches += 1;
    m->pipe_pending = false;
    return PT_OK;
}

int pt_comm_destroy(pt_ctx* c) {
    if (!c) return fail(PT_EINVAL, "pt_comm_destroy: null context");
    pt_comm_release(c);
    return PT_OK;
}

}  // extern "C"
