// pt_internal.h — what the translation units of libpt.so share behind the C ABI: the context and the error helpers.
#pragma once
#include <cuda_runtime.h>

#include <string>

#include "../../include/pt.h"
#include "pt_common.cuh"

int pt_fail(int code, const std::string& msg);          // records the thread-local message pt_last_error() returns; returns code
#define fail pt_fail
#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return pt_fail(PT_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_));          \
    } while (0)

struct PtComm;                                           // pt_comm.cu: multi-GPU state of a context
void pt_comm_release(struct pt_ctx* c);                  // called by pt_destroy
int pt_comm_check(struct pt_ctx* c);                     // called by pt_synchronize: error flag of the pipelined combine
extern "C" int pt_copy_setup(struct pt_ctx* c);          // pt_api.cu (internal, not in pt.h): creates the copy stream + events on first use
extern "C" int pt_copy_enqueue(struct pt_ctx* c, void* host, const void* dev, size_t bytes, bool first);   // D2H on the copy stream, after ev_image

struct pt_ctx {
    PtConfig cfg;
    cudaStream_t stream = nullptr;
    // device timing of render launches: a ring of event pairs, folded into render_ms lazily (finished pairs) or at a
    // synchronising call, so that back-to-back pt_render_spp calls never wait for the previous kernel
    static const int NEV = 32;
    cudaEvent_t ev0[NEV] = {}, ev1[NEV] = {};
    int ev_head = 0, ev_tail = 0;            // pairs [ev_tail, ev_head) are pending
    // framebuffer
    float4* accum = nullptr;
    int* primary_ids = nullptr;
    unsigned long long* ray_counter = nullptr;
    int* error_flag = nullptr;
    unsigned* cost = nullptr;    // cost map, allocated by pt_enable_cost_map
    pt::WaveBuffers wf;          // wavefront path state, allocated at the first wavefront render
    void* wf_block = nullptr;
    uint32_t* sched = nullptr;   // persistent pipeline: work head
    uint32_t* tile_tab = nullptr;   // persistent pipeline: visiting order of a frame's tiles (RenderParams.tile_tab)
    size_t tile_cap = 0;
    float* sobol_tab = nullptr;  // variant 2: Sobol coordinates of the frames of one launch, [sobol_cap][8]
    size_t sobol_cap = 0;
    uchar4* display = nullptr;   // pt_readback_display staging
    int num_sms = 0;
    // scene
    bool has_scene = false;
    int fast_depth = 0, n_fast_inner = 0;
    float fast_pad = 0.0f;       // padding of the production tree's boxes = 2^-17 x (largest |coordinate| of the scene)
    double build_ms = 0.0;
    float *tris_raw = nullptr, *nodes_raw = nullptr;
    float4 *wnodes = nullptr, *wnodes_fast = nullptr, *trigeo_fast = nullptr, *trigeo = nullptr, *trishade = nullptr, *materials = nullptr, *hdr4 = nullptr, *cache4 = nullptr, *light_tab = nullptr;
    float4* trigeo_fast48 = nullptr;
    int2* trimeta = nullptr;
    float *env_pdf = nullptr, *env_cdf_x = nullptr, *env_cdf_y = nullptr;   // PT_ENV_CDF_SEARCH tables
    pt::DevScene sc;
    // camera + region
    float eye[3] = {0, 0, 4};
    float cam[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 4, 1};
    int x0 = 0, y0 = 0, x1 = 0, y1 = 0;
    int band_world = 1, band_rank = 0, band_diag = 0;
    // stats
    unsigned long long samples = 0, launches = 0;
    double render_ms = 0.0;
    unsigned long long wf_waves = 0, wf_retries = 0;   // wavefront pipeline: waves rendered / repeated since the last reset
    PtComm* comm = nullptr;      // pt_comm_init
    // asynchronous host copies (pt_readback_async, pt_combine | PT_COMBINE_ASYNC): a second stream carries the device-to-host
    // copy of a finished image while the context's stream already renders the next step
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_image = nullptr, ev_copied = nullptr;
    bool copy_pending = false;
    float4* snapshot = nullptr;  // pt_readback_async: the framebuffer as of the call (the next render overwrites accum)
};

