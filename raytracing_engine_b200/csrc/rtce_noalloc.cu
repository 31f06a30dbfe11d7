// rtce_noalloc.cu — rtce_trace_noalloc (include/rtce.h): the strip-kernel pass with NO device allocation.
//
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The ordinary path
// (rtce_create / rtce_set_view / rtce_trace_batch) allocates the scratch of the strip kernel — task tickets, overflow
// queue, the cameras in device memory — per lane with cudaMalloc.  Here the same scratch is ONE file-scope
// `static __device__` array of fixed size: every GPU that loads librtce.so holds kNoAllocBytes (about 2.4 MB) from
// module load until the process exits, and this entry point calls no cudaMalloc / cudaFree / allocating routine.
//
//   * In host code `g_noalloc_scratch` is a host-side stand-in for the symbol, not a device address.  The device
//     address comes from cudaGetSymbolAddress, asked once per device with that device current (noalloc_scratch_of),
//     and that pointer is what the kernels get in their parameters.  Device code of this file (k_noalloc_setup)
//     names the array directly.
//   * Every call shares the array and the array keeps the last call's contents: the task tickets / queue counters of
//     the levels and the cameras are rewritten by k_noalloc_setup at the head of every call, on the caller's stream
//     (the queue body needs no clearing: only entries below the freshly zeroed count are ever read).
//   * Two calls that overlap on one GPU (two streams, two host threads) RACE on the array: keep calls on a device in
//     stream order.  Calls on different GPUs are independent (one copy of the array per device).
//
// Same kernels, same arithmetic and therefore the same bits as rtce_trace_batch with RTCE_OPT_STRIP_MIN_RAYS = 0
// (tests/test_zz_no_alloc.py compares them level by level); the far-miss shortcut is off (Cam::pad = +inf), which
// changes no value.
#include <cstdio>
#include <cstring>
#include <limits>
#include <mutex>

#include "../../include/rtce.h"
#include "rtce_host.hpp"
#include "rtce_kernels.cuh"

namespace rtce {

constexpr int kNoAllocFrames = RTCE_NOALLOC_MAX_FRAMES;
constexpr unsigned kNoAllocOqCap = 1u << 15;   // rays of the overflow queue; beyond it long rays stay in their warp (exact either way)
constexpr int kOqWords = 10 + kSmallSceneMax;  // words per queued ray (rtce_kernels_strip.cu)

struct NoAllocScratch {
    int tickets[kMaxLevels][kTicketWords];     // {-, overflow count, overflow head, -, .., task counters} per level
    Cam cams[kNoAllocFrames];                  // the frames' cameras, read by the strip kernel from device memory
    float oq[(size_t)kNoAllocOqCap * kOqWords];
};
constexpr size_t kNoAllocBytes = sizeof(NoAllocScratch);
static_assert(kNoAllocBytes < (size_t)4 << 20, "the static scratch is meant to stay modest");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

struct NoAllocSetup {
    Cam cams[kNoAllocFrames];
    int frames;
};

// Head of every call: what must start at zero (tickets, queue counters) and the cameras.  Names the array directly.
__global__ void k_noalloc_setup(const __grid_constant__ NoAllocSetup S) {
    NoAllocScratch* sc = reinterpret_cast<NoAllocScratch*>(g_noalloc_scratch);
    const int t = threadIdx.x;
    for (int k = t; k < kMaxLevels * kTicketWords; k += blockDim.x) (&sc->tickets[0][0])[k] = 0;
    constexpr int kCamWords = (int)(sizeof(Cam) / sizeof(float));
    for (int k = t; k < S.frames * kCamWords; k += blockDim.x)
        reinterpret_cast<float*>(sc->cams)[k] = reinterpret_cast<const float*>(S.cams)[k];
}

// Device address of the array on the CURRENT device; asked once per device.
static cudaError_t noalloc_scratch_of(NoAllocScratch** out) {
    static std::mutex mu;
    static NoAllocScratch* addr[kMaxDevices] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        addr[dev] = static_cast<NoAllocScratch*>(p);
    }
    *out = addr[dev];
    return cudaSuccess;
}

static int noalloc_fail(char* err, size_t err_len, int code, const char* what, cudaError_t e = cudaSuccess) {
    if (err && err_len) {
        if (e != cudaSuccess) snprintf(err, err_len, "rtce_trace_noalloc: %s: %s", what, cudaGetErrorString(e));
        else snprintf(err, err_len, "rtce_trace_noalloc: %s", what);
    }
    return code;
}

}  // namespace rtce

using namespace rtce;

extern "C" int rtce_trace_noalloc(int device, void* cuda_stream, const float* spheres, int n_spheres,
                                  const rtce_constant_buffer* cb, int levels, float render_dist,
                                  const rtce_push_constants* pcs, int frames, float* d_pyramid, size_t frame_stride,
                                  char* err, size_t err_len) {
    if (err && err_len) err[0] = 0;
    if (!cb || !pcs || !d_pyramid || (n_spheres > 0 && !spheres)) return noalloc_fail(err, err_len, RTCE_ERR_INVALID, "null argument");
    if (n_spheres < 0 || frames < 1) return noalloc_fail(err, err_len, RTCE_ERR_INVALID, "n_spheres < 0 or frames < 1");
    if (n_spheres > kSmallSceneMax)
        return noalloc_fail(err, err_len, RTCE_ERR_LIMIT, "more than 8 spheres (MAX_OBJECTS, utilities.glsl:3): use rtce_trace_batch");
    if (frames > kNoAllocFrames) {
        char msg[160];
        snprintf(msg, sizeof msg, "%d frames exceed RTCE_NOALLOC_MAX_FRAMES = %d (the static scratch holds that many cameras)", frames,
                 kNoAllocFrames);
        return noalloc_fail(err, err_len, RTCE_ERR_LIMIT, msg);
    }
    if (!(render_dist > 0.0f)) return noalloc_fail(err, err_len, RTCE_ERR_INVALID, "render_dist must be > 0 (the strip kernel assumes it)");
    if (levels == 0) levels = host_level_count(cb->view[0], false);
    if (levels < 1) return noalloc_fail(err, err_len, RTCE_ERR_INVALID, "levels");
    if (levels > kMaxLevels) return noalloc_fail(err, err_len, RTCE_ERR_LIMIT, "levels exceed RTCE_MAX_LEVELS (the static scratch holds that many ticket blocks)");
    int lw[kMaxLevels], lh[kMaxLevels];
    size_t loff[kMaxLevels], total = 0;
    for (int i = 0; i < levels; i++) {
        host_level_dims(levels, cb->view[0], cb->view[1], i, &lw[i], &lh[i]);
        if (lw[i] <= 0 || lh[i] <= 0) return noalloc_fail(err, err_len, RTCE_ERR_INVALID, "empty level");
        loff[i] = total;
        total += (size_t)lw[i] * (size_t)lh[i];
    }
    if (frame_stride < total) return noalloc_fail(err, err_len, RTCE_ERR_INVALID, "frame_stride smaller than the packed pyramid of one frame");
    if ((unsigned long long)frames * frame_stride >= (1ull << 32))
        return noalloc_fail(err, err_len, RTCE_ERR_LIMIT, "frames * frame_stride must stay below 2^32 floats (32-bit queue offsets)");

    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e != cudaSuccess) return noalloc_fail(err, err_len, RTCE_ERR_CUDA, "cudaGetDevice", e);
    if (prev != device && (e = cudaSetDevice(device)) != cudaSuccess) return noalloc_fail(err, err_len, RTCE_ERR_CUDA, "cudaSetDevice", e);
    struct Restore {
        int prev, dev;
        ~Restore() { if (prev != dev) cudaSetDevice(prev); }
    } restore{prev, device};

    NoAllocScratch* sc = nullptr;   // the DEVICE address of g_noalloc_scratch on `device`
    if ((e = noalloc_scratch_of(&sc)) != cudaSuccess) return noalloc_fail(err, err_len, RTCE_ERR_CUDA, "cudaGetSymbolAddress(scratch)", e);
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);

    NoAllocSetup setup{};
    for (int f = 0; f < frames; f++) {
        memcpy(setup.cams[f].rot, pcs[f].rot, sizeof(float) * 4);
        memcpy(setup.cams[f].pos, pcs[f].pos, sizeof(float) * 3);
        setup.cams[f].pad = std::numeric_limits<float>::infinity();   // no far-miss shortcut
    }
    setup.frames = frames;
    k_noalloc_setup<<<1, 256, 0, st>>>(setup);
    if ((e = cudaGetLastError()) != cudaSuccess) return noalloc_fail(err, err_len, RTCE_ERR_CUDA, "setup launch", e);

    float4 sph[kSmallSceneMax] = {};
    for (int i = 0; i < n_spheres; i++) sph[i] = make_float4(spheres[4 * i], spheres[4 * i + 1], spheres[4 * i + 2], spheres[4 * i + 3]);

    for (int i = 0; i < levels; i++) {   // main.rs:300-316
        float isz[2];
        host_image_size(levels, i, cb->view[0], cb->view[1], isz);
        StripParams S{};
        S.out = d_pyramid + loff[i];
        S.parent = i > 0 ? d_pyramid + loff[i - 1] : nullptr;
        S.slot_stride = frame_stride;
        S.counters = nullptr;
        S.tickets = sc->tickets[i];
        S.oq = sc->oq;
        S.oq_count = reinterpret_cast<unsigned*>(sc->tickets[i] + 1);
        S.oq_cap = kNoAllocOqCap;
        S.w = lw[i];
        S.pw = i > 0 ? lw[i - 1] : 0;
        S.x0 = 0; S.y0 = 0; S.x1 = lw[i]; S.y1 = lh[i];
        S.ntx = (lw[i] + 31) / 32;
        int rows = 16;
        while (rows > 2 && (long long)S.ntx * ((lh[i] + rows - 1) / rows) * frames < 2LL * 148 * 32) rows >>= 1;
        S.rows = rows;
        S.nty = (lh[i] + rows - 1) / rows;
        S.cta_share = 1;
        S.ctas_per_sm = 0;
        S.two_pass = 1;
        S.frames = frames;
        S.n_spheres = n_spheres;
        S.isx = isz[0]; S.isy = isz[1];
        S.ratx = cb->ratio[0]; S.raty = cb->ratio[1];
        S.threshold = host_threshold(isz[0]);
        S.render_dist = render_dist;
        memcpy(S.sph, sph, sizeof(S.sph));
        S.cams = sc->cams;
        if ((e = launch_trace_strip(S, false, st)) != cudaSuccess) return noalloc_fail(err, err_len, RTCE_ERR_CUDA, "strip launch", e);
        if (n_spheres > 0) {
            OverflowParams O{};
            O.out = S.out;
            O.oq = S.oq;
            O.oq_count = S.oq_count;
            O.oq_head = reinterpret_cast<unsigned*>(sc->tickets[i] + 2);
            O.oq_cap = S.oq_cap;
            O.counters = nullptr;
            O.n_spheres = n_spheres;
            O.threshold = S.threshold;
            O.render_dist = render_dist;
            memcpy(O.sph, sph, sizeof(O.sph));
            if ((e = launch_march_overflow(O, false, st)) != cudaSuccess) return noalloc_fail(err, err_len, RTCE_ERR_CUDA, "overflow launch", e);
        }
    }
    return RTCE_OK;
}
