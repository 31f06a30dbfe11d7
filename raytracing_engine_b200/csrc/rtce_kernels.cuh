// rtce_kernels.cuh — launch interface between the C-ABI layer (rtce_api.cu) and the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtce {

// Per-device launch configuration is cached per DEVICE (a process may hold contexts on several GPUs): index = CUDA
// device ordinal.  Entries are written once with values that depend only on (kernel, device), so a race between two
// host threads writes the same values twice.
constexpr int kMaxDevices = 64;
inline int current_device_slot() {
    int dev = 0;
    cudaGetDevice(&dev);
    return dev >= 0 && dev < kMaxDevices ? dev : 0;
}
inline int device_sm_count() {
    static int sms[kMaxDevices] = {};
    const int d = current_device_slot();
    if (!sms[d]) {
        int n = 148;
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, d);
        sms[d] = n > 0 ? n : 148;
    }
    return sms[d];
}

constexpr int kMaxBatch = 32;        // frames per launch (camera table lives in kernel params)
constexpr int kSmallSceneMax = 8;    // MAX_OBJECTS of the reference shader (utilities.glsl:3)
constexpr int kMaxLevels = 16;

// Camera part of PushConstantData (compute.glsl:7-12): rot = quaternion [x,y,z,w], pos.
struct Cam {
    float rot[4];
    float pos[3];
    float pad;
};

// Everything one level-dispatch of compute.glsl:main needs, passed by value in the
// kernel parameter (constant) bank: the per-level push constants (main.rs:301-307), the
// ConstantBuffer ratio (main.rs:610), RENDER_DIST (compute.glsl:32), the objs[] part of
// MutableData (compute.glsl:17-24) and the camera of every frame of the batch.
struct LevelParams {
    float* out;              // level image of frame slot 0
    const float* parent;     // level-1 image of frame slot 0 (unused at level 0)
    size_t slot_stride;      // floats between consecutive frame slots
    unsigned long long* counters;   // {R, E, S} or nullptr
    int w, pw;               // row pitch (= width) of this level and of the parent level
    int x0, y0, x1, y1;      // rectangle of this level to compute
    int iter;                // pc.iter
    int n_spheres;
    float isx, isy;          // pc.imageSize
    float ratx, raty;        // cs.ratio
    float threshold;         // 1.4142135 * 8 * imageSize.x  (compute.glsl:75)
    float render_dist;
    float4 sph[kSmallSceneMax];   // {pos.xyz, size}
    const Cam* cams;         // the same cameras in device memory (k_trace_level_pix reads these: launches that
                             // are replayed from a CUDA graph cannot carry per-step values in their parameters)
    Cam cam[kMaxBatch];
};

// Per-level kernel, objCount <= 8 (one thread per pixel, frames in grid.z).
cudaError_t launch_trace_level_small(const LevelParams& p, int frames, bool count, cudaStream_t st);
// Same dispatch for the coarse, latency-bound levels: 8x4 pixels per warp, branch-free march trips,
// short correctly rounded sqrt.  Requires RENDER_DIST > 0.
cudaError_t launch_trace_level_pix(const LevelParams& p, int frames, bool count, cudaStream_t st);
// The author's alternative marchers (tracing_algorithms.txt: algorithm 1, algorithm 2) in the same frame; objCount 1..8.
cudaError_t launch_trace_level_alt(const LevelParams& p, int frames, bool count, int marcher, cudaStream_t st);

// Large scenes (objCount > 8, extension of utilities.glsl:3): spheres in global memory.
struct LevelParamsLarge {
    float* out;
    const float* parent;
    size_t slot_stride;
    unsigned long long* counters;
    const float4* spheres;   // n x {pos.xyz, size}, 16-byte aligned
    int w, pw;
    int x0, y0, x1, y1;
    int iter;
    int n_spheres;
    float isx, isy, ratx, raty, threshold, render_dist;
    const int* ray_list;          // nullptr = every ray of the rectangle; else ray indices (frame-major over the rectangle) ...
    const unsigned* ray_count;    // ... and their number, both in device memory (register kernel only: the event-driven
                                  // kernel's fallback list)
    Cam cam[kMaxBatch];
};
// The per-ray cache of lower bounds in registers (rtce_kernels_large_reg.cu); n_spheres <= large_reg_capacity().
cudaError_t launch_trace_level_large_reg(const LevelParamsLarge& p, int frames, bool count, cudaStream_t st);
int large_reg_capacity();
// Event-driven form (rtce_kernels_large_ev.cu): one ray per warp, exact bounds only for the objects near in "time";
// rays it cannot hold are traced by the register kernel from a fallback list (launched behind it on the same stream).
// `scratch` = large_ev_scratch_bytes(n_spheres, max_rays) bytes of device memory, reusable by the next launch on the stream.
size_t large_ev_scratch_bytes(int n_spheres, long long max_rays_per_launch, int* warps_out);
cudaError_t launch_trace_level_large_ev(LevelParamsLarge p, int frames, bool count, void* scratch, long long max_rays,
                                        cudaStream_t st);

// ---- strip kernel (objCount <= 8): ONE level of a frame batch, persistent warps ---------------
// The levels that hold almost all rays (the last three: 75 % + 19 % + 5 %) are each traced by one
// launch of persistent, barrier-free warps.  A warp task is a 32-wide x `rows`-high tile of the
// level; a lane keeps one image column for the whole task.  Rays that need more than two march
// trips are parked in a per-warp shared-memory queue that persists across the warp's tasks and
// is marched with lane refill (DESIGN.md "Strip kernel").
#ifndef RTCE_STRIP_THREADS
#define RTCE_STRIP_THREADS 256
#endif
constexpr int kStripThreads = RTCE_STRIP_THREADS;
// Task tickets of ONE strip launch.  A single global counter hands out at most one ticket per ~0.72 ns however many warps
// ask (same-address atomics serialise in L2: tools/probes/atomic_ticket_probe.cu, 130 560 tickets = 93 us, more than
// half of the final-level launch), so the ticket space is STRIPED over kTicketCounters counters, one 128-byte line
// each: counter k hands out tickets k, k + K, k + 2K, ... — all counters advance through the ticket space together,
// which keeps the global order (hot tiles first) — a warp draws from its home counter (warp index mod K) and, once
// that is exhausted, from the others in turn.
constexpr int kTicketCounters = 16;
constexpr int kTicketStride = 32;     // ints between counters
constexpr int kTicketWords = 32 + kTicketCounters * kTicketStride;   // per launch: {-, overflow count, overflow head, -, ...} + the counters
constexpr int kStripMaxRows = 32;      // rows per task (row-term table size)

struct StripParams {
    float* out;               // level image of frame slot 0
    const float* parent;      // level-1 image of frame slot 0; nullptr at level 0 (start = 1.0)
    size_t slot_stride;       // floats between frame slots (frames * slot_stride < 2^32: queue offsets are 32-bit)
    unsigned long long* counters;
    int* tickets;             // zero-initialised kTicketWords block of this launch (counters from word 32)
    float* oq;                // overflow queue: rays that outlive kMaxAge trips in a warp's tail go here (nullptr = none)
    unsigned* oq_count;       // zero-initialised number of rays in oq
    unsigned oq_cap;          // capacity of oq in rays; word w of ray k is oq[w * oq_cap + k]
    unsigned long long* dbg;  // nullptr, or 4 words per warp: globaltimer at start / tickets exhausted / exit, tasks done
    int w, pw;                // row pitch of this level and of the parent level
    int x0, y0, x1, y1;       // region of interest at this level
    int ntx, nty;             // tasks per frame: 32-wide columns x `rows`-high rows covering the ROI
    unsigned long long magic_tiles, magic_ntx;   // floor(2^40 / d) + 1 for d = ntx * nty and d = ntx (0: divide): ticket decode
                                                 // without integer division, exact while tasks < 2^24 and d < 2^16
    int rows;                 // rows per task, 1..kStripMaxRows
    int cta_share;            // launch 1/cta_share of the resident CTA slots (>= 1)
    int ctas_per_sm;          // > 0: launch at most this many CTAs per SM (leaves room for other lanes' launches)
    int two_pass;             // 1: hot tiles first, the rest in a second walk of the tickets (a launch that owns the device)
    int frames;
    int n_spheres;
    float isx, isy, ratx, raty, threshold, render_dist;
    float4 sph[kSmallSceneMax];
    const Cam* cams;          // cameras of the `frames` frames, device memory, 32-byte records
};
cudaError_t launch_trace_strip(const StripParams& p, bool count, cudaStream_t st);

// Second half of a strip launch: marches the rays of its overflow queue to the end.  The long rays
// (cones grazing a sphere: 10 .. 100+ march trips, 0.2 % of all rays but a quarter of all trips,
// clustered in a few tiles) would otherwise serialise in the warp that found them; here every warp
// of the device draws them 32 at a time, with lane refill.
struct OverflowParams {
    float* out;                 // level image of frame slot 0 (queue entries carry offsets from here)
    const float* oq;
    const unsigned* oq_count;
    unsigned* oq_head;          // zero-initialised consumer cursor
    unsigned oq_cap;
    unsigned long long* counters;
    int n_spheres;
    float threshold, render_dist;
    float4 sph[kSmallSceneMax];
};
cudaError_t launch_march_overflow(const OverflowParams& p, bool count, cudaStream_t st);

// ---- presentation shading pass (fragment.glsl:123-187; objCount, lightCount <= 8) -----------------
struct ShadeParams {
    const float* depth;       // final pyramid level of frame slot 0 (binding 0)
    float4* out;              // RGBA32F colour image of colour slot 0
    size_t depth_stride;      // floats between frame slots
    size_t out_stride;        // float4 between colour slots
    unsigned long long* counters;   // {pixels, hit pixels, distance() calls, light evaluations} or nullptr
    const Cam* cams;          // push constants rot / pos of every frame, device memory
    int w, h;                 // final level image (pitch = w)
    int out_w;                // pitch of the colour image = window width (cs.view.x <= w)
    int x0, y0, x1, y1;       // rectangle of window pixels to shade
    int n_obj, n_light;
    float vw, vh, ratx, raty, render_dist;   // cs.view, cs.ratio, RENDER_DIST
    float4 obj[kSmallSceneMax];    // objs[i] = {pos.xyz, size}
    float4 matc[kSmallSceneMax];   // mats[i] = {color.rgb, shine}
    float amb[kSmallSceneMax];     // mats[i].ambient
    float4 lpos[kSmallSceneMax];   // lights[i].pos
    float4 lcol[kSmallSceneMax];   // lights[i].color
};
cudaError_t launch_shade(const ShadeParams& p, int frames, bool count, cudaStream_t st);
// RGBA32F -> RGBA8 UNORM (clamp, x255, round to nearest even)
cudaError_t launch_rgba8(const float4* in, uchar4* out, size_t n, cudaStream_t st);

// Accumulation extension: acc = (...((first ? 0 : acc) + img[0]) + img[1]) ... + img[frames-1], the images
// `slot_stride` floats apart, added in frame order in one pass; and the final scale by 1/frames.
cudaError_t launch_accumulate(float* acc, const float* img, size_t slot_stride, int frames, size_t n, bool first,
                              cudaStream_t st);
cudaError_t launch_scale(float* acc, size_t n, float divisor, cudaStream_t st);

}  // namespace rtce
