// rtce_comm.cu — the ONE exchange of the multi-GPU path: gather of the accumulated image's row
// bands to one rank, behind the C ABI so that a Rust / C host reaches several GPUs without Python.
//
// The reference is single-device (one queue, src/main.rs:450-460); this is new, not ported.  Every rank traces its
// band of final-level rows (rtce_set_rect) into its own accumulator; a row band of a row-major image is contiguous,
// so rank r sends exactly its rows and the destination receives every band IN PLACE at its row offset: grouped
// ncclSend / ncclRecv over NVLink / NVSwitch, no padding, no reassembly.  The exchange runs on a communication stream
// of its own out of a staging copy of the band (two buffers), so the gather of step k overlaps the tracing of step
// k + 1, which overwrites the accumulator.
//
// NCCL is loaded on first use with dlopen("libnccl.so.2") — a single-GPU caller needs no NCCL at all, and inside a
// process that already holds an NCCL (e.g. PyTorch's bundled one) the loader hands back that same library.
#include <dlfcn.h>
#include <nccl.h>   // types and prototypes only; the library is not linked

#include <algorithm>
#include <cstring>
#include <vector>

#include "rtce_ctx.hpp"

namespace rtce {

struct NcclApi {
    void* lib = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclSend) Send = nullptr;
    decltype(&ncclRecv) Recv = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
};

static NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.lib ? &api : nullptr;
    tried = true;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return nullptr;
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(dlsym(h, "ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(dlsym(h, "ncclGroupEnd"));
    api.Send = reinterpret_cast<decltype(api.Send)>(dlsym(h, "ncclSend"));
    api.Recv = reinterpret_cast<decltype(api.Recv)>(dlsym(h, "ncclRecv"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
    if (!api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.GroupStart || !api.GroupEnd || !api.Send || !api.Recv) {
        dlclose(h);
        return nullptr;
    }
    api.lib = h;
    return &api;
}

struct CommState {
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    cudaStream_t st = nullptr;               // communication stream
    float* d_band[2] = {nullptr, nullptr};   // staging copies of this rank's band
    size_t band_floats[2] = {0, 0};
    float* d_full[2] = {nullptr, nullptr};   // destination rank: the gathered image
    size_t full_floats[2] = {0, 0};
    cudaEvent_t ev_ready[2] = {}, ev_done[2] = {}, ev_t0[2] = {}, ev_t1[2] = {};
    bool used[2] = {false, false};
    unsigned k = 0;
    int last = -1;                           // buffer of the most recent gather
    int last_dst = -1;
    // rtce_gather_frames: one "exchange finished" event per lane (the lane's window may be overwritten after it)
    cudaEvent_t ev_frames_ready = nullptr;
    cudaEvent_t ev_lane_done[kMaxLanes] = {};
    bool lane_used[kMaxLanes] = {};
    cudaEvent_t ev_f0 = nullptr, ev_f1 = nullptr;
};

static int comm_state(rtce_ctx* c, CommState** out) {
    if (!c->comm) {
        CommState* s = new (std::nothrow) CommState();
        if (!s) return fail(c, RTCE_ERR_NOMEM, "communication state");
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);   // hi = greatest priority: the exchange should not queue behind a persistent wave
        bool ok = cudaStreamCreateWithPriority(&s->st, cudaStreamNonBlocking, hi) == cudaSuccess;
        for (int i = 0; ok && i < 2; i++)
            ok = cudaEventCreateWithFlags(&s->ev_ready[i], cudaEventDisableTiming) == cudaSuccess &&
                 cudaEventCreateWithFlags(&s->ev_done[i], cudaEventDisableTiming) == cudaSuccess &&
                 cudaEventCreate(&s->ev_t0[i]) == cudaSuccess && cudaEventCreate(&s->ev_t1[i]) == cudaSuccess;
        ok = ok && cudaEventCreateWithFlags(&s->ev_frames_ready, cudaEventDisableTiming) == cudaSuccess;
        for (int l = 0; ok && l < kMaxLanes; l++) ok = cudaEventCreateWithFlags(&s->ev_lane_done[l], cudaEventDisableTiming) == cudaSuccess;
        c->comm = s;
        if (!ok) {
            cudaGetLastError();
            comm_destroy(c);
            return fail(c, RTCE_ERR_CUDA, "communication stream / events");
        }
    }
    *out = c->comm;
    return RTCE_OK;
}

void comm_destroy(rtce_ctx* c) {
    CommState* s = c->comm;
    if (!s) return;
    if (s->st) cudaStreamSynchronize(s->st);
    if (s->comm) {
        NcclApi* api = nccl_api();
        if (api) api->CommDestroy(s->comm);
    }
    for (int i = 0; i < 2; i++) {
        if (s->d_band[i]) cudaFree(s->d_band[i]);
        if (s->d_full[i]) cudaFree(s->d_full[i]);
        if (s->ev_ready[i]) cudaEventDestroy(s->ev_ready[i]);
        if (s->ev_done[i]) cudaEventDestroy(s->ev_done[i]);
        if (s->ev_t0[i]) cudaEventDestroy(s->ev_t0[i]);
        if (s->ev_t1[i]) cudaEventDestroy(s->ev_t1[i]);
    }
    if (s->ev_frames_ready) cudaEventDestroy(s->ev_frames_ready);
    for (int l = 0; l < kMaxLanes; l++)
        if (s->ev_lane_done[l]) cudaEventDestroy(s->ev_lane_done[l]);
    if (s->st) cudaStreamDestroy(s->st);
    delete s;
    c->comm = nullptr;
}

// `st` waits (device side) until the exchange that last read / wrote lane `lane`'s window has finished.
int comm_wait_lane(rtce_ctx* c, int lane, cudaStream_t st) {
    CommState* s = c->comm;
    if (s && s->lane_used[lane]) RTCE_CUDA(c, cudaStreamWaitEvent(st, s->ev_lane_done[lane], 0));
    return RTCE_OK;
}

// `st` waits for every exchange in flight.
int comm_wait_all(rtce_ctx* c, cudaStream_t st) {
    CommState* s = c->comm;
    if (!s) return RTCE_OK;
    for (int k = 0; k < 2; k++)
        if (s->used[k]) RTCE_CUDA(c, cudaStreamWaitEvent(st, s->ev_done[k], 0));
    for (int l = 0; l < kMaxLanes; l++)
        if (s->lane_used[l]) RTCE_CUDA(c, cudaStreamWaitEvent(st, s->ev_lane_done[l], 0));
    return RTCE_OK;
}

static int nccl_fail(rtce_ctx* c, NcclApi* api, const char* what, ncclResult_t r) {
    fail(c, RTCE_ERR_CUDA, what);
    if (c && api && api->GetErrorString) {
        c->err += ": ";
        c->err += api->GetErrorString(r);
    }
    return RTCE_ERR_CUDA;
}

static int grow(rtce_ctx* c, float** p, size_t* have, size_t need, cudaStream_t user) {
    if (*have >= need) return RTCE_OK;
    if (*p) {
        RTCE_CUDA(c, cudaStreamSynchronize(user));
        RTCE_CUDA(c, cudaStreamSynchronize(c->comm->st));
        RTCE_CUDA(c, cudaFree(*p));
        *p = nullptr;
        *have = 0;
    }
    cudaError_t e = cudaMalloc(p, need * sizeof(float));
    if (e != cudaSuccess) return fail(c, RTCE_ERR_NOMEM, "cudaMalloc(gather buffer)", e);
    *have = need;
    return RTCE_OK;
}

}  // namespace rtce

using namespace rtce;

extern "C" {

int rtce_host_band_rows(int height, int world, int rank, int* y0, int* y1) {
    if (!y0 || !y1 || height < 1 || world < 1 || rank < 0 || rank >= world) return RTCE_ERR_INVALID;
    const int align = 8;   // whole workgroup rows of the reference dispatch (compute.glsl:5, main.rs:314)
    const int units = (height + align - 1) / align;
    const int base = units / world, extra = units % world;
    const int u0 = rank * base + std::min(rank, extra);
    const int u1 = u0 + base + (rank < extra ? 1 : 0);
    *y0 = std::min(u0 * align, height);
    *y1 = std::min(u1 * align, height);
    return RTCE_OK;
}

int rtce_comm_unique_id(void* id_out) {
    if (!id_out) return RTCE_ERR_INVALID;
    NcclApi* api = nccl_api();
    if (!api) return RTCE_ERR_STATE;   // no libnccl.so.2 on this machine
    ncclUniqueId id;
    if (api->GetUniqueId(&id) != ncclSuccess) return RTCE_ERR_CUDA;
    static_assert(sizeof(id) == RTCE_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
    memcpy(id_out, &id, sizeof(id));
    return RTCE_OK;
}

int rtce_comm_init(rtce_ctx* c, const void* id_128, int rank, int world) {
    if (!c || !id_128 || world < 1 || rank < 0 || rank >= world) return fail(c, RTCE_ERR_INVALID, "rtce_comm_init: arguments");
    NcclApi* api = nccl_api();
    if (!api) return fail(c, RTCE_ERR_STATE, "rtce_comm_init: libnccl.so.2 not found");
    DeviceGuard g(c->device);
    CommState* s = nullptr;
    int r = comm_state(c, &s);
    if (r) return r;
    if (s->comm) {
        RTCE_CUDA(c, cudaStreamSynchronize(s->st));
        api->CommDestroy(s->comm);
        s->comm = nullptr;
    }
    ncclUniqueId id;
    memcpy(&id, id_128, sizeof(id));
    ncclResult_t nr = api->CommInitRank(&s->comm, world, id, rank);
    if (nr != ncclSuccess) {
        s->comm = nullptr;
        return nccl_fail(c, api, "ncclCommInitRank", nr);
    }
    s->rank = rank;
    s->world = world;
    return RTCE_OK;
}

int rtce_gather_accum(rtce_ctx* c, int dst, const int* bands) {
    if (!c) return RTCE_ERR_INVALID;
    if (!c->have_view || !c->d_accum) return fail(c, RTCE_ERR_STATE, "rtce_gather_accum: nothing accumulated");
    DeviceGuard g(c->device);
    CommState* s = nullptr;
    int r = comm_state(c, &s);
    if (r) return r;
    const int world = s->comm ? s->world : 1, rank = s->comm ? s->rank : 0;
    if (dst < 0 || dst >= world) return fail(c, RTCE_ERR_INVALID, "rtce_gather_accum: dst");
    const int W = c->lw[c->levels - 1], H = c->lh[c->levels - 1];
    std::vector<int> tab(2 * (size_t)world);
    for (int q = 0; q < world; q++) {
        if (bands) {
            tab[2 * q] = bands[2 * q];
            tab[2 * q + 1] = bands[2 * q + 1];
        } else {
            rtce_host_band_rows(H, world, q, &tab[2 * q], &tab[2 * q + 1]);
        }
        if (tab[2 * q] < 0 || tab[2 * q + 1] > H || tab[2 * q] > tab[2 * q + 1]) return fail(c, RTCE_ERR_INVALID, "rtce_gather_accum: band table");
    }
    const int y0 = tab[2 * rank], y1 = tab[2 * rank + 1];
    const size_t n_mine = (size_t)(y1 - y0) * (size_t)W;
    const int k = (int)(s->k++ % 2);
    // the gather that last used buffer pair k has finished with it
    if (s->used[k]) RTCE_CUDA(c, cudaStreamWaitEvent(c->stream, s->ev_done[k], 0));
    const float* src = c->d_accum + (size_t)y0 * W;
    if (rank == dst) {
        if ((r = grow(c, &s->d_full[k], &s->full_floats[k], (size_t)W * H, c->stream))) return r;
        if (n_mine)   // the destination's own band goes straight to its place
            RTCE_CUDA(c, cudaMemcpyAsync(s->d_full[k] + (size_t)y0 * W, src, n_mine * sizeof(float), cudaMemcpyDeviceToDevice, c->stream));
    } else if (n_mine) {
        if ((r = grow(c, &s->d_band[k], &s->band_floats[k], n_mine, c->stream))) return r;
        RTCE_CUDA(c, cudaMemcpyAsync(s->d_band[k], src, n_mine * sizeof(float), cudaMemcpyDeviceToDevice, c->stream));
    }
    RTCE_CUDA(c, cudaEventRecord(s->ev_ready[k], c->stream));
    RTCE_CUDA(c, cudaStreamWaitEvent(s->st, s->ev_ready[k], 0));
    RTCE_CUDA(c, cudaEventRecord(s->ev_t0[k], s->st));
    if (world > 1) {
        NcclApi* api = nccl_api();
        ncclResult_t nr = api->GroupStart();
        if (nr != ncclSuccess) return nccl_fail(c, api, "ncclGroupStart", nr);
        if (rank == dst) {
            for (int q = 0; q < world && nr == ncclSuccess; q++) {
                const size_t n = (size_t)(tab[2 * q + 1] - tab[2 * q]) * (size_t)W;
                if (q != dst && n) nr = api->Recv(s->d_full[k] + (size_t)tab[2 * q] * W, n, ncclFloat, q, s->comm, s->st);
            }
        } else if (n_mine) {
            nr = api->Send(s->d_band[k], n_mine, ncclFloat, dst, s->comm, s->st);
        }
        const ncclResult_t ne = api->GroupEnd();
        if (nr != ncclSuccess) return nccl_fail(c, api, "ncclSend / ncclRecv", nr);
        if (ne != ncclSuccess) return nccl_fail(c, api, "ncclGroupEnd", ne);
    }
    RTCE_CUDA(c, cudaEventRecord(s->ev_t1[k], s->st));
    RTCE_CUDA(c, cudaEventRecord(s->ev_done[k], s->st));
    s->used[k] = true;
    s->last = k;
    s->last_dst = dst;
    c->launches++;
    return RTCE_OK;
}

int rtce_gather_wait(rtce_ctx* c) {
    if (!c) return RTCE_ERR_INVALID;
    DeviceGuard g(c->device);
    return comm_wait_all(c, c->stream);
}

int rtce_gather_frames(rtce_ctx* c, int dst, const int* bands, int frames) {
    if (!c) return RTCE_ERR_INVALID;
    if (!c->have_view || !c->d_pyr) return fail(c, RTCE_ERR_STATE, "rtce_gather_frames: nothing traced");
    if (frames < 1 || frames > c->slots) return fail(c, RTCE_ERR_INVALID, "rtce_gather_frames: frames");
    DeviceGuard g(c->device);
    CommState* s = nullptr;
    int r = comm_state(c, &s);
    if (r) return r;
    const int world = s->comm ? s->world : 1, rank = s->comm ? s->rank : 0;
    if (dst < 0 || dst >= world) return fail(c, RTCE_ERR_INVALID, "rtce_gather_frames: dst");
    const int L = c->levels - 1;
    const int W = c->lw[L], H = c->lh[L];
    std::vector<int> tab(2 * (size_t)world);
    for (int q = 0; q < world; q++) {
        if (bands) {
            tab[2 * q] = bands[2 * q];
            tab[2 * q + 1] = bands[2 * q + 1];
        } else {
            rtce_host_band_rows(H, world, q, &tab[2 * q], &tab[2 * q + 1]);
        }
        if (tab[2 * q] < 0 || tab[2 * q + 1] > H || tab[2 * q] > tab[2 * q + 1]) return fail(c, RTCE_ERR_INVALID, "rtce_gather_frames: band table");
    }
    const int lane = c->last_lane;
    Lane& Ln = c->lanes[lane];
    // after the batch: its chain ran on the lane's stream when lanes rotate, else on the context's stream
    cudaStream_t src = c->lanes_active > 1 ? Ln.st : c->stream;
    RTCE_CUDA(c, cudaEventRecord(s->ev_frames_ready, src));
    RTCE_CUDA(c, cudaStreamWaitEvent(s->st, s->ev_frames_ready, 0));
    if (world > 1) {
        NcclApi* api = nccl_api();
        ncclResult_t nr = api->GroupStart();
        if (nr != ncclSuccess) return nccl_fail(c, api, "ncclGroupStart", nr);
        for (int k = 0; k < frames && nr == ncclSuccess; k++) {
            float* img = window_slot(c, lane, k) + c->loff[L];
            if (rank == dst) {
                for (int q = 0; q < world && nr == ncclSuccess; q++) {
                    const size_t n = (size_t)(tab[2 * q + 1] - tab[2 * q]) * (size_t)W;
                    if (q != dst && n) nr = api->Recv(img + (size_t)tab[2 * q] * W, n, ncclFloat, q, s->comm, s->st);
                }
            } else {
                const size_t n = (size_t)(tab[2 * rank + 1] - tab[2 * rank]) * (size_t)W;
                if (n) nr = api->Send(img + (size_t)tab[2 * rank] * W, n, ncclFloat, dst, s->comm, s->st);
            }
        }
        const ncclResult_t ne = api->GroupEnd();
        if (nr != ncclSuccess) return nccl_fail(c, api, "ncclSend / ncclRecv", nr);
        if (ne != ncclSuccess) return nccl_fail(c, api, "ncclGroupEnd", ne);
    }
    RTCE_CUDA(c, cudaEventRecord(s->ev_lane_done[lane], s->st));
    s->lane_used[lane] = true;
    c->launches++;
    return RTCE_OK;
}

int rtce_gather_info(const rtce_ctx* c, int* w, int* h, float** dptr) {
    if (!c) return RTCE_ERR_INVALID;
    const CommState* s = c->comm;
    if (!c->have_view || !s || s->last < 0 || !s->d_full[s->last]) return RTCE_ERR_STATE;   // not the destination, or nothing gathered
    if (w) *w = c->lw[c->levels - 1];
    if (h) *h = c->lh[c->levels - 1];
    if (dptr) *dptr = s->d_full[s->last];
    return RTCE_OK;
}

int rtce_readback_gathered(rtce_ctx* c, float* host, size_t pitch_bytes) {
    if (!c || !host) return fail(c, RTCE_ERR_INVALID, "rtce_readback_gathered: null");
    int w = 0, h = 0;
    float* d = nullptr;
    int r = rtce_gather_info(c, &w, &h, &d);
    if (r) return fail(c, r, "rtce_readback_gathered: nothing gathered on this rank");
    DeviceGuard g(c->device);
    if ((r = rtce_gather_wait(c))) return r;
    const size_t row = (size_t)w * sizeof(float);
    if (pitch_bytes == 0) pitch_bytes = row;
    if (pitch_bytes < row) return fail(c, RTCE_ERR_INVALID, "rtce_readback_gathered: pitch < row");
    RTCE_CUDA(c, cudaMemcpy2DAsync(host, pitch_bytes, d, row, row, (size_t)h, cudaMemcpyDeviceToHost, c->stream));
    RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
    return RTCE_OK;
}

int rtce_gather_last_ms(rtce_ctx* c, float* ms) {
    if (!c || !ms) return RTCE_ERR_INVALID;
    CommState* s = c->comm;
    if (!s || s->last < 0) return fail(c, RTCE_ERR_STATE, "rtce_gather_last_ms: nothing gathered");
    DeviceGuard g(c->device);
    RTCE_CUDA(c, cudaStreamSynchronize(s->st));
    RTCE_CUDA(c, cudaEventElapsedTime(ms, s->ev_t0[s->last], s->ev_t1[s->last]));
    return RTCE_OK;
}

}  // extern "C"
