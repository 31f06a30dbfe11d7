// rtce_api.cu — the C ABI declared in include/rtce.h, first part: context lifetime, options, scene (binding 1)
// and view (binding 2 + the pyramid of binding 0).  The pass loop lives in rtce_api_trace.cu.
// No CPU fallback exists anywhere in the library: every compute entry point launches CUDA kernels.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <new>
#include <limits>
#include <string>
#include <vector>

#include "rtce_ctx.hpp"

using namespace rtce;

namespace rtce {

int fail(rtce_ctx* c, int code, const char* what, cudaError_t e) {
    if (c) {
        c->err = what;
        if (e != cudaSuccess) {
            c->err += ": ";
            c->err += cudaGetErrorString(e);
        }
    }
    return code;
}

// Allocate what lane `l` needs (idempotent).
int lane_init(rtce_ctx* c, int l) {
    Lane& L = c->lanes[l];
    if (L.ready) return RTCE_OK;
    bool ok = cudaStreamCreateWithFlags(&L.st, cudaStreamNonBlocking) == cudaSuccess &&
              cudaEventCreateWithFlags(&L.ev_fork, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&L.ev_done, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&L.ev_free, cudaEventDisableTiming) == cudaSuccess;
    for (int g = 0; ok && g < kMaxGroups; g++)
        ok = cudaStreamCreateWithFlags(&L.grp_stream[g], cudaStreamNonBlocking) == cudaSuccess &&
             cudaEventCreateWithFlags(&L.ev_join[g], cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; ok && i < kCamRing; i++) ok = cudaEventCreateWithFlags(&L.ev_cam[i], cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaMalloc(&L.d_tickets, sizeof(int) * kMaxGroups * kMaxLevels * kTicketWords) == cudaSuccess &&
         cudaMalloc(&L.d_cams, sizeof(Cam) * kMaxBatch) == cudaSuccess &&
         cudaMallocHost(&L.h_cams, sizeof(Cam) * kMaxBatch * kCamRing) == cudaSuccess &&
         cudaMalloc(&L.d_oq, sizeof(float) * kMaxGroups * (size_t)kOverflowCap * (10 + kSmallSceneMax)) == cudaSuccess;
    if (!ok) {
        cudaGetLastError();
        return fail(c, RTCE_ERR_NOMEM, "lane resources");
    }
    L.ready = true;
    return RTCE_OK;
}

void lane_free(Lane& L) {
    if (L.st) cudaStreamSynchronize(L.st);
    for (int g = 0; g < kMaxGroups; g++) {
        if (L.grp_stream[g]) { cudaStreamSynchronize(L.grp_stream[g]); cudaStreamDestroy(L.grp_stream[g]); }
        if (L.ev_join[g]) cudaEventDestroy(L.ev_join[g]);
    }
    for (auto& g : L.graphs) cudaGraphExecDestroy(g.exec);
    L.graphs.clear();
    for (int i = 0; i < kCamRing; i++)
        if (L.ev_cam[i]) cudaEventDestroy(L.ev_cam[i]);
    if (L.ev_fork) cudaEventDestroy(L.ev_fork);
    if (L.ev_done) cudaEventDestroy(L.ev_done);
    if (L.ev_free) cudaEventDestroy(L.ev_free);
    if (L.d_cams) cudaFree(L.d_cams);
    if (L.h_cams) cudaFreeHost(L.h_cams);
    if (L.d_tickets) cudaFree(L.d_tickets);
    if (L.d_oq) cudaFree(L.d_oq);
    if (L.st) cudaStreamDestroy(L.st);
    L = Lane();
}

// Make the context's stream wait for every lane that has traced since it last did (stream-level, the host
// does not block): after this, work enqueued on the context's stream sees every image traced so far.
int flush_lanes(rtce_ctx* c) {
    int r = comm_wait_all(c, c->stream);   // frames / bands still travelling between ranks
    if (r) return r;
    for (int l = 0; l < kMaxLanes; l++) {
        Lane& L = c->lanes[l];
        if (L.ready && L.pending) {
            RTCE_CUDA(c, cudaStreamWaitEvent(c->stream, L.ev_done, 0));
            L.pending = false;
        }
    }
    return RTCE_OK;
}

// Everything issued so far has finished (host blocks).
int quiesce(rtce_ctx* c) {
    int r = flush_lanes(c);
    if (r) return r;
    if (c->host_copies_pending) {   // frames of rtce_render_to_host_async still on their way to host memory
        RTCE_CUDA(c, cudaStreamSynchronize(c->copy_stream));
        c->host_copies_pending = false;
    }
    RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
    return RTCE_OK;
}

// The pyramid allocation holds `windows` windows of c->slots frame slots (window w belongs to lane w).
int ensure_pyramid(rtce_ctx* c, int windows) {
    windows = std::max(windows, std::max(1, c->windows_alloc));   // never shrink: a lane's graphs point into its window
    const size_t need = c->slot_stride * (size_t)c->slots * (size_t)windows;
    if (c->d_pyr && c->d_pyr_floats >= need && c->slots_alloc == c->slots && c->windows_alloc >= windows) return RTCE_OK;
    int r = quiesce(c);
    if (r) return r;
    if (c->d_pyr) {
        RTCE_CUDA(c, cudaFree(c->d_pyr));
        c->d_pyr = nullptr;
        c->d_pyr_floats = 0;
    }
    cudaError_t e = cudaMalloc(&c->d_pyr, need * sizeof(float));
    if (e != cudaSuccess) {
        c->windows_alloc = 0;
        c->slots_alloc = 0;
        return fail(c, RTCE_ERR_NOMEM, "cudaMalloc(pyramid)", e);
    }
    c->plan_version++;
    c->d_pyr_floats = need;
    c->slots_alloc = c->slots;
    c->windows_alloc = windows;
    // the reference clears every level image before its dispatch (main.rs:312); texels outside a
    // set rect are never written by the pass, so clear once here to give them the same value
    RTCE_CUDA(c, cudaMemsetAsync(c->d_pyr, 0, need * sizeof(float), c->stream));
    RTCE_CUDA(c, cudaStreamSynchronize(c->stream));   // lanes on other streams start from the cleared allocation
    return RTCE_OK;
}

}  // namespace rtce

extern "C" {

int rtce_abi_version(void) { return RTCE_ABI_VERSION; }

const char* rtce_status_string(int s) {
    switch (s) {
        case RTCE_OK: return "ok";
        case RTCE_ERR_INVALID: return "invalid argument";
        case RTCE_ERR_CUDA: return "CUDA error";
        case RTCE_ERR_STATE: return "scene or view not set";
        case RTCE_ERR_LIMIT: return "limit exceeded";
        case RTCE_ERR_NOMEM: return "out of device memory";
        default: return "unknown status";
    }
}

const char* rtce_last_error(const rtce_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int rtce_create(int device, float render_dist, rtce_ctx** out) {
    if (!out) return RTCE_ERR_INVALID;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0) return RTCE_ERR_CUDA;   // no GPU: there is no fallback
    if (device < 0 || device >= ndev) return RTCE_ERR_INVALID;
    rtce_ctx* c = new (std::nothrow) rtce_ctx();
    if (!c) return RTCE_ERR_NOMEM;
    c->device = device;
    c->render_dist = render_dist;
    DeviceGuard g(device);
    if (cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_user, cudaEventDisableTiming) != cudaSuccess ||
        cudaMalloc(&c->d_counters, 4 * kMaxLevels * sizeof(unsigned long long)) != cudaSuccess) {
        rtce_destroy(c);
        return RTCE_ERR_CUDA;
    }
    // every lane the default options can use exists from the start: no stream / event creation and no
    // cudaMalloc happens inside a later rtce_trace* call
    for (int l = 0; l < std::max(c->in_flight, kAccumLanes); l++) {
        if (lane_init(c, l) != RTCE_OK) {
            rtce_destroy(c);
            return RTCE_ERR_CUDA;
        }
    }
    for (int i = 0; i < 2; i++) {
        cudaEventCreateWithFlags(&c->ev_traced[i], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&c->ev_copied[i], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&c->ev_host_done[i], cudaEventDisableTiming);
    }
    c->stream = c->own_stream;
    *out = c;
    return RTCE_OK;
}

void rtce_destroy(rtce_ctx* c) {
    if (!c) return;
    DeviceGuard g(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->copy_stream) cudaStreamSynchronize(c->copy_stream);
    comm_destroy(c);
    for (int l = 0; l < kMaxLanes; l++) lane_free(c->lanes[l]);
    if (c->large_state) cudaFree(c->large_state);
    if (c->ev_user) cudaEventDestroy(c->ev_user);
    for (int i = 0; i < 2; i++) {
        if (c->ev_traced[i]) cudaEventDestroy(c->ev_traced[i]);
        if (c->ev_copied[i]) cudaEventDestroy(c->ev_copied[i]);
        if (c->ev_host_done[i]) cudaEventDestroy(c->ev_host_done[i]);
    }
    if (c->d_pyr) cudaFree(c->d_pyr);
    if (c->d_accum) cudaFree(c->d_accum);
    if (c->d_spheres) cudaFree(c->d_spheres);
    if (c->d_counters) cudaFree(c->d_counters);
    if (c->d_dbg) cudaFree(c->d_dbg);
    if (c->d_color) cudaFree(c->d_color);
    if (c->d_rgba8) cudaFree(c->d_rgba8);
    if (c->d_shade_cams) cudaFree(c->d_shade_cams);
    if (c->h_shade_cams) cudaFreeHost(c->h_shade_cams);
    if (c->ev_shade_cams) cudaEventDestroy(c->ev_shade_cams);
    if (c->d_shade_counters) cudaFree(c->d_shade_counters);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    delete c;
}

int rtce_set_option(rtce_ctx* c, int option, int64_t value) {
    if (!c) return RTCE_ERR_INVALID;
    {
        DeviceGuard g(c->device);
        int r = flush_lanes(c);   // batches in flight were issued under the old options; order them before what follows
        if (r) return r;
    }
    c->plan_version++;
    switch (option) {
        case RTCE_OPT_FRAMES_IN_FLIGHT:
            if (value < 1 || value > kMaxLanes) return fail(c, RTCE_ERR_INVALID, "frames in flight out of range");
            c->in_flight = (int)value;
            c->next_lane = 0;
            {
                DeviceGuard g(c->device);
                for (int l = 0; l < std::max(c->in_flight, kAccumLanes); l++) {
                    int r = lane_init(c, l);
                    if (r) return r;
                }
                if (c->have_view) {   // the windows of the new lanes, now rather than inside the next trace call
                    int r = ensure_pyramid(c, c->in_flight);
                    if (r && r != RTCE_ERR_NOMEM) return r;
                }
            }
            return RTCE_OK;
        case RTCE_OPT_GRAPH: c->use_graph = value != 0; return RTCE_OK;
        case RTCE_OPT_COUNTERS: c->count = value != 0; return RTCE_OK;
        case RTCE_OPT_FRAME_SLOTS:
            if (value < 1 || value > 4096) return fail(c, RTCE_ERR_INVALID, "frame slots out of range");
            c->slots = (int)value;
            if (c->have_view) {
                DeviceGuard g(c->device);
                return ensure_pyramid(c);
            }
            return RTCE_OK;
        case RTCE_OPT_KERNEL:
            if (value != 0 && value != 1 && value != 3) return fail(c, RTCE_ERR_INVALID, "kernel plan: 0 auto, 1 per-pixel, 3 strip");
            c->kernel_mode = (int)value;
            return RTCE_OK;
        case RTCE_OPT_STRIP_ROWS:
            if (value < 1 || value > kStripMaxRows) return fail(c, RTCE_ERR_INVALID, "strip rows out of range");
            c->strip_rows = (int)value;
            return RTCE_OK;
        case RTCE_OPT_FAR_MISS: c->far_cull = value != 0; return RTCE_OK;
        case RTCE_OPT_LARGE_KERNEL:
            if (value != 0 && value != 2 && value != 3) return fail(c, RTCE_ERR_INVALID, "large-scene kernel: 0 auto, 2 register, 3 event-driven");
            c->large_mode = (int)value;
            return RTCE_OK;
        case RTCE_OPT_STRIP_CTAS_PER_SM:
            if (value < 0 || value > 8) return fail(c, RTCE_ERR_INVALID, "strip CTAs per SM out of range");
            c->strip_ctas_per_sm = (int)value;
            return RTCE_OK;
        case RTCE_OPT_MARCHER:
            if (value < 1 || value > 3) return fail(c, RTCE_ERR_INVALID, "marcher: 1, 2 or 3");
            c->marcher = (int)value;
            return RTCE_OK;
        case RTCE_OPT_GROUPS:
            if (value < 0 || value > kMaxGroups) return fail(c, RTCE_ERR_INVALID, "groups out of range");
            c->groups = (int)value;
            return RTCE_OK;
        case RTCE_OPT_STRIP_MIN_RAYS:
            if (value < 0) return fail(c, RTCE_ERR_INVALID, "strip threshold out of range");
            c->strip_min_rays = (long long)value;
            return RTCE_OK;
        default: return fail(c, RTCE_ERR_INVALID, "unknown option");
    }
}

int rtce_set_stream(rtce_ctx* c, void* cuda_stream) {
    if (!c) return RTCE_ERR_INVALID;
    DeviceGuard g(c->device);
    int r = quiesce(c);
    if (r) return r;
    c->plan_version++;
    c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
    return RTCE_OK;
}

int rtce_set_spheres(rtce_ctx* c, const rtce_sphere* spheres, uint32_t n) {
    if (!c || (n && !spheres)) return fail(c, RTCE_ERR_INVALID, "rtce_set_spheres: null");
    if (n > RTCE_MAX_SPHERES) return fail(c, RTCE_ERR_LIMIT, "rtce_set_spheres: n > RTCE_MAX_SPHERES");
    DeviceGuard g(c->device);
    int r = flush_lanes(c);
    if (r) return r;
    if (n > (uint32_t)kSmallSceneMax && (r = quiesce(c))) return r;   // the sphere buffer is about to be rewritten
    c->plan_version++;
    c->have_materials = false;
    c->n_spheres = (int)n;
    memset(c->small, 0, sizeof(c->small));
    if (n <= (uint32_t)kSmallSceneMax) {
        for (uint32_t i = 0; i < n; i++)
            c->small[i] = make_float4(spheres[i].pos[0], spheres[i].pos[1], spheres[i].pos[2], spheres[i].size);
    } else {
        if (c->d_spheres_cap < n) {
            if (c->d_spheres) {
                RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
                RTCE_CUDA(c, cudaFree(c->d_spheres));
                c->d_spheres = nullptr;
            }
            cudaError_t e = cudaMalloc(&c->d_spheres, sizeof(float4) * (size_t)n);
            if (e != cudaSuccess) return fail(c, RTCE_ERR_NOMEM, "cudaMalloc(spheres)", e);
            c->d_spheres_cap = n;
        }
        // rtce_sphere is {x,y,z,size} = the float4 layout; pageable source -> staged copy completes before return
        RTCE_CUDA(c, cudaMemcpyAsync(c->d_spheres, spheres, sizeof(float4) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
        RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    c->have_scene = true;
    return RTCE_OK;
}

int rtce_set_scene(rtce_ctx* c, const void* md, size_t nbytes) {
    if (!c || !md) return fail(c, RTCE_ERR_INVALID, "rtce_set_scene: null");
    if (nbytes != RTCE_MUTABLE_DATA_SIZE) return fail(c, RTCE_ERR_INVALID, "rtce_set_scene: MutableData is 656 bytes (std140)");
    const unsigned char* b = (const unsigned char*)md;
    uint32_t n = 0;
    memcpy(&n, b + RTCE_MUTABLE_DATA_OBJCOUNT_OFFSET, 4);
    // the shader indexes objs[i] for i < objCount with objs[MAX_OBJECTS]; more than 8 is out of bounds there
    if (n > RTCE_REF_MAX_OBJECTS) return fail(c, RTCE_ERR_LIMIT, "rtce_set_scene: objCount > MAX_OBJECTS (8)");
    uint32_t nl = 0;
    memcpy(&nl, b + 8, 4);
    if (nl > RTCE_REF_MAX_OBJECTS) return fail(c, RTCE_ERR_LIMIT, "rtce_set_scene: lightCount > MAX_LIGHTS (8)");
    rtce_sphere s[RTCE_REF_MAX_OBJECTS];
    memcpy(s, b + RTCE_MUTABLE_DATA_OBJS_OFFSET, sizeof(s));
    int r = rtce_set_spheres(c, s, n);
    if (r) return r;
    memcpy(c->md, b, RTCE_MUTABLE_DATA_SIZE);   // materials and lights: the shading pass reads them (fragment.glsl:18-25)
    c->have_materials = true;
    return RTCE_OK;
}

int rtce_host_level_count(float view_w, int resize_rule) { return host_level_count(view_w, resize_rule != 0); }

int rtce_host_level_dims(int levels, float vw, float vh, int level, int* w, int* h) {
    if (!w || !h || levels < 1 || levels > RTCE_MAX_LEVELS || level < 0 || level >= levels) return RTCE_ERR_INVALID;
    host_level_dims(levels, vw, vh, level, w, h);
    return RTCE_OK;
}

int rtce_host_image_size(int levels, int level, float vw, float vh, float out[2]) {
    if (!out || levels < 1 || levels > RTCE_MAX_LEVELS || level < 0 || level >= levels) return RTCE_ERR_INVALID;
    host_image_size(levels, level, vw, vh, out);
    return RTCE_OK;
}

int rtce_set_view(rtce_ctx* c, const rtce_constant_buffer* cb, int levels) {
    if (!c || !cb) return fail(c, RTCE_ERR_INVALID, "rtce_set_view: null");
    if (!(cb->view[0] >= 1.0f) || !(cb->view[1] >= 1.0f)) return fail(c, RTCE_ERR_INVALID, "rtce_set_view: view < 1");
    if (levels == 0) levels = host_level_count(cb->view[0], false);
    else if (levels == -1) levels = host_level_count(cb->view[0], true);
    if (levels < 1) return fail(c, RTCE_ERR_INVALID, "rtce_set_view: levels");
    if (levels > RTCE_MAX_LEVELS) return fail(c, RTCE_ERR_LIMIT, "rtce_set_view: levels > RTCE_MAX_LEVELS");
    DeviceGuard g(c->device);
    size_t off = 0;
    int lw[kMaxLevels], lh[kMaxLevels];
    size_t loff[kMaxLevels];
    for (int i = 0; i < levels; i++) {
        host_level_dims(levels, cb->view[0], cb->view[1], i, &lw[i], &lh[i]);
        if (lw[i] <= 0 || lh[i] <= 0) return fail(c, RTCE_ERR_INVALID, "rtce_set_view: empty level");
        loff[i] = off;
        off += ((size_t)lw[i] * (size_t)lh[i] + 63) & ~(size_t)63;   // 256-byte aligned levels
    }
    c->plan_version++;
    c->cb = *cb;
    c->levels = levels;
    memcpy(c->lw, lw, sizeof(lw));
    memcpy(c->lh, lh, sizeof(lh));
    memcpy(c->loff, loff, sizeof(loff));
    const bool same = c->have_view && c->slot_stride == off;
    int r = quiesce(c);   // nothing in flight reads the old geometry past this point
    if (r) return r;
    c->slot_stride = off;
    c->rect[0] = c->rect[1] = c->rect[2] = c->rect[3] = 0;
    c->have_view = true;
    if (c->d_accum) {   // the accumulator belongs to the old geometry
        RTCE_CUDA(c, cudaFree(c->d_accum));
        c->d_accum = nullptr;
        c->d_accum_floats = 0;
        c->accum_rows[0] = c->accum_rows[1] = -1;
    }
    if (!same) {
        c->d_pyr_floats = 0;   // force re-allocation (and the clear)
        if (c->d_pyr) {
            RTCE_CUDA(c, cudaFree(c->d_pyr));
            c->d_pyr = nullptr;
        }
    }
    // the slot windows of every lane in flight are allocated here, not inside the first trace call
    r = ensure_pyramid(c, c->in_flight);
    if (r == RTCE_ERR_NOMEM) r = ensure_pyramid(c, 1);
    return r;
}

int rtce_set_rect(rtce_ctx* c, int x0, int y0, int x1, int y1) {
    if (!c) return RTCE_ERR_INVALID;
    if (!c->have_view) return fail(c, RTCE_ERR_STATE, "rtce_set_rect: view not set");
    const int W = c->lw[c->levels - 1], H = c->lh[c->levels - 1];
    {
        DeviceGuard g(c->device);
        int r = flush_lanes(c);
        if (r) return r;
    }
    c->plan_version++;
    if (x0 == 0 && y0 == 0 && x1 == 0 && y1 == 0) {
        c->rect[0] = c->rect[1] = c->rect[2] = c->rect[3] = 0;
        return RTCE_OK;
    }
    if (x0 < 0 || y0 < 0 || x1 > W || y1 > H || x0 >= x1 || y0 >= y1) return fail(c, RTCE_ERR_INVALID, "rtce_set_rect: out of range");
    c->rect[0] = x0; c->rect[1] = y0; c->rect[2] = x1; c->rect[3] = y1;
    return RTCE_OK;
}

int rtce_level_count(const rtce_ctx* c) { return (c && c->have_view) ? c->levels : RTCE_ERR_STATE; }

int rtce_level_info(const rtce_ctx* c, int level, int slot, int* w, int* h, float** dptr) {
    if (!c) return RTCE_ERR_INVALID;
    if (!c->have_view) return RTCE_ERR_STATE;
    if (level < 0) level += c->levels;   // -1 = final level
    if (level < 0 || level >= c->levels || slot < 0 || slot >= c->slots) return RTCE_ERR_INVALID;
    if (w) *w = c->lw[level];
    if (h) *h = c->lh[level];
    if (dptr) *dptr = window_slot(c, c->last_lane, slot) + c->loff[level];   // the window of the latest batch
    return RTCE_OK;
}

}  // extern "C"

