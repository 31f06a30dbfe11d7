// rtce_ctx.hpp — internal state of the C ABI (include/rtce.h): the context, its lanes, and the helpers the
// translation units of the API layer share (rtce_api.cu: lifetime / scene / view; rtce_api_trace.cu: the pass
// loop of src/main.rs:291-316 and the consumers of binding 0; rtce_comm.cu: the multi-GPU gather).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/rtce.h"
#include "rtce_host.hpp"
#include "rtce_kernels.cuh"

namespace rtce {

constexpr int kMaxGroups = 8;
constexpr unsigned kOverflowCap = 1u << 19;   // rays per frame group (38 MB each); beyond it long rays stay in their warp
constexpr int kDefaultInFlight = 8;           // RTCE_OPT_FRAMES_IN_FLIGHT default
constexpr int kAccumLanes = 4;                // chunks of rtce_trace_accumulate in flight, at least

// A lane is one self-contained launch chain: its own stream (plus one per frame group), its own
// window of frame slots in the pyramid allocation, task counters, overflow queues, camera buffer
// and captured graphs.  With RTCE_OPT_FRAMES_IN_FLIGHT > 1 successive rtce_trace_batch calls rotate
// over the lanes and overlap on the device — the latency-bound coarse levels and long-ray launches of
// one batch run under the issue-bound final level of another — and rtce_trace_accumulate rotates its
// chunks the same way.  Lanes are created EAGERLY (rtce_create / RTCE_OPT_FRAMES_IN_FLIGHT) and the
// launch graphs of a chunk shape are captured for every lane the first time the shape is seen (or by
// rtce_prepare): no allocation, stream creation or graph instantiation happens on a later call.
#ifndef RTCE_MAX_LANES
#define RTCE_MAX_LANES 8
#endif
constexpr int kMaxLanes = RTCE_MAX_LANES;
constexpr int kCamRing = 8;
struct ChunkGraph { int frames, slot0, lanes; uint64_t version; cudaGraphExec_t exec; uint64_t launches; };   // lanes: 0 = strict order
struct Lane {
    bool ready = false;
    bool pending = false;                  // traced since the context's stream last waited for it
    cudaStream_t st = nullptr;             // the lane's chain when it does not run on the context's stream
    cudaStream_t grp_stream[kMaxGroups] = {};   // one per frame group of a chunk, forked from / joined to the chain's stream
    cudaEvent_t ev_fork = nullptr, ev_join[kMaxGroups] = {};
    cudaEvent_t ev_done = nullptr;         // last chunk traced
    cudaEvent_t ev_free = nullptr;         // the consumer is done with the window
    // Cameras of the chunk being traced.  The strip and coarse-level kernels read them from device memory, not
    // from their parameters: their launch chain is replayed from a CUDA graph whose parameters are frozen.
    Cam* d_cams = nullptr;                 // [kMaxBatch]
    Cam* h_cams = nullptr;                 // pinned staging, [kCamRing][kMaxBatch]
    cudaEvent_t ev_cam[kCamRing] = {};     // staging entry consumed
    unsigned cam_seq = 0;
    std::vector<ChunkGraph> graphs;
    int* d_tickets = nullptr;    // [kMaxGroups][kMaxLevels][kTicketWords] {-, overflow count, overflow head, -, .., task counters}, zeroed per chunk
    float* d_oq = nullptr;       // overflow queues of the strip launches (kMaxGroups x kOverflowCap rays x 18 words)
};

struct CommState;   // rtce_comm.cu

}  // namespace rtce

struct rtce_ctx {
    int device = 0;
    float render_dist = 1000.0f;
    cudaStream_t own_stream = nullptr;   // created by rtce_create
    cudaStream_t stream = nullptr;       // the stream work is issued on (own_stream or the caller's)
    cudaStream_t copy_stream = nullptr;  // D2H copies of rtce_render_to_host
    int groups = 0;                      // RTCE_OPT_GROUPS (0 = auto)
    int strip_rows = 16;                 // RTCE_OPT_STRIP_ROWS
    int strip_ctas_per_sm = 0;           // RTCE_OPT_STRIP_CTAS_PER_SM (0 = every resident slot)
    int large_mode = 0;                  // RTCE_OPT_LARGE_KERNEL: 0 auto, 2 register kernel, 3 event-driven kernel
    int marcher = 3;                     // RTCE_OPT_MARCHER: 3 = the shipped algorithm, 1 / 2 = the author's alternatives
    bool far_cull = true;                // RTCE_OPT_FAR_MISS: resolve rays that start beyond every sphere without SDFs
    // diagnostics (rtce_debug_launch_timeline): an event pair around every kernel launch
    struct LaunchMark { cudaEvent_t e0, e1; int kind, level, group; };
    bool mark_launches = false;
    std::vector<LaunchMark> marks;
    cudaEvent_t ev_traced[2] = {nullptr, nullptr};
    cudaEvent_t ev_copied[2] = {nullptr, nullptr};
    // rtce_render_to_host(_async): which half of window 0 has a copy outstanding or finished on the copy stream; the
    // completion events of the last two asynchronous calls; whether copies may still be reading window 0
    bool half_used[2] = {false, false};
    cudaEvent_t ev_host_done[2] = {nullptr, nullptr};
    unsigned host_calls = 0;
    bool host_copies_pending = false;
    std::string err;

    // binding 1 (MutableData.objCount / objs)
    bool have_scene = false;
    int n_spheres = 0;
    float4 small[rtce::kSmallSceneMax];
    float4* d_spheres = nullptr;   // objCount > 8 only
    size_t d_spheres_cap = 0;
    void* large_state = nullptr;   // objCount > 8: scratch of the event-driven kernel (rtce_kernels_large_ev.cu)
    size_t large_state_bytes = 0;

    // binding 2 (ConstantBuffer) + binding 0 (the pyramid)
    bool have_view = false;
    rtce_constant_buffer cb{};
    int levels = 0;
    int lw[rtce::kMaxLevels]{}, lh[rtce::kMaxLevels]{};
    size_t loff[rtce::kMaxLevels]{};   // float offset of each level inside one slot
    size_t slot_stride = 0;      // floats per slot (all levels, 256-byte aligned each)
    int slots = 16;
    int slots_alloc = 0;
    float* d_pyr = nullptr;
    size_t d_pyr_floats = 0;
    int rect[4] = {0, 0, 0, 0};
    float* d_accum = nullptr;
    size_t d_accum_floats = 0;
    int accum_rows[2] = {-1, -1};        // rows of d_accum the last rtce_trace_accumulate wrote (others are zero)

    // presentation shading pass (fragment.glsl): the whole MutableData image, colour slots, staging
    bool have_materials = false;             // rtce_set_scene was given the 656-byte image (rtce_set_spheres carries no materials / lights)
    unsigned char md[RTCE_MUTABLE_DATA_SIZE] = {};
    float4* d_color = nullptr;               // [slots][h][w] RGBA32F
    size_t d_color_px = 0;                   // pixels per colour slot
    int color_slots = 0;
    uchar4* d_rgba8 = nullptr;               // [slots] RGBA8 UNORM, written by rtce_readback_color_rgba8 / rtce_render_rgba8_to_host
    rtce::Cam* d_shade_cams = nullptr;       // [kMaxBatch]
    rtce::Cam* h_shade_cams = nullptr;       // pinned
    cudaEvent_t ev_shade_cams = nullptr;
    unsigned long long* d_shade_counters = nullptr;   // {pixels, hit pixels, distance() calls, light evaluations}

    // options / counters
    bool count = false;
    int kernel_mode = 0;
    long long strip_min_rays = 400000;   // a level goes to the strip kernel when it has this many rays per launch
    unsigned long long* d_counters = nullptr;
    unsigned long long* d_dbg = nullptr;   // strip-kernel warp timeline of one level (rtce_debug_strip_timeline)
    int dbg_level = -1;
    // CUDA graphs of a chunk's launch chain (RTCE_OPT_GRAPH), valid while plan_version is unchanged
    bool use_graph = true;
    uint64_t plan_version = 1;             // bumped by every setter that changes what a launch chain looks like
    // lanes: independent launch chains over their own window of frame slots (see Lane)
    rtce::Lane lanes[rtce::kMaxLanes];
    int in_flight = rtce::kDefaultInFlight;   // RTCE_OPT_FRAMES_IN_FLIGHT
    int next_lane = 0, last_lane = 0;
    int lanes_active = 1;                  // lanes the current call rotates over
    int windows_alloc = 0;                 // slot windows the pyramid allocation holds
    cudaEvent_t ev_user = nullptr;         // "everything the caller enqueued so far" on `stream`
    uint64_t launches = 0;

    // multi-GPU gather (rtce_comm.cu)
    rtce::CommState* comm = nullptr;
};

namespace rtce {

int fail(rtce_ctx* c, int code, const char* what, cudaError_t e = cudaSuccess);

#define RTCE_CUDA(c, call)                                                      \
    do {                                                                        \
        cudaError_t e_ = (call);                                                \
        if (e_ != cudaSuccess) return ::rtce::fail((c), RTCE_ERR_CUDA, #call, e_); \
    } while (0)

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

int lane_init(rtce_ctx* c, int l);            // allocate what lane `l` needs (idempotent)
int flush_lanes(rtce_ctx* c);                 // the context's stream waits (device side) for every lane that traced
int quiesce(rtce_ctx* c);                     // everything issued so far has finished (host blocks)
int ensure_pyramid(rtce_ctx* c, int windows = 1);
void comm_destroy(rtce_ctx* c);               // rtce_comm.cu
int comm_wait_lane(rtce_ctx* c, int lane, cudaStream_t st);   // `st` waits for the exchange that last used lane's window
int comm_wait_all(rtce_ctx* c, cudaStream_t st);

// First float of frame slot `slot` of window `w`.
inline float* window_slot(const rtce_ctx* c, int w, int slot) {
    return c->d_pyr + ((size_t)w * (size_t)c->slots + (size_t)slot) * c->slot_stride;
}

}  // namespace rtce
