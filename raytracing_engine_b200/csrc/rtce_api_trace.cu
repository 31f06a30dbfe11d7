// rtce_api_trace.cu — the C ABI declared in include/rtce.h, second part: the pass loop (src/main.rs:291-316)
// expressed as kernel launches / replayed launch graphs, and the consumers of binding 0 (headless readback,
// the presentation shading pass).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>

#include "rtce_ctx.hpp"

using namespace rtce;

namespace {

// Diagnostics: bracket a launch with a timed event pair on its stream.
struct MarkScope {
    rtce_ctx* c;
    cudaStream_t st;
    cudaEvent_t e1 = nullptr;
    MarkScope(rtce_ctx* c_, cudaStream_t st_, int kind, int level, int group) : c(c_), st(st_) {
        if (!c->mark_launches) return;
        rtce_ctx::LaunchMark m{nullptr, nullptr, kind, level, group};
        cudaEventCreate(&m.e0);
        cudaEventCreate(&m.e1);
        cudaEventRecord(m.e0, st);
        e1 = m.e1;
        c->marks.push_back(m);
    }
    ~MarkScope() {
        if (e1) cudaEventRecord(e1, st);
    }
};

// Far-miss distance of a camera (Cam::pad): a ray whose parent texel is at least this far starts beyond every
// sphere by more than RENDER_DIST, so every distances[i] of compute.glsl:37-39 is >= RENDER_DIST, the first march
// trip takes dist = RENDER_DIST (:49, :56), is no hit (:62, RENDER_DIST > threshold) and ends the loop (:44):
// the texel is start + RENDER_DIST without evaluating a single SDF.  distances[i] = |pos + dir start - c_i| - r_i
// >= start |dir| - (|pos - c_i| + r_i); the bound below keeps a 0.1 % + 1 margin over the ~1e-6 relative rounding
// of the computed value (|dir| = 1 +- 3e-7), plus a term for the rounding of origin = pos + dir * start itself, which
// scales with the COORDINATES (half an ulp of |pos| + start per component), not with the distances: a camera at 1e8
// with a sphere 10 units away has k ~ 10 but an origin known only to +- 8.  The bound is a fixed point of
// t = base + u (|pos|_inf + t) with u = 1e-6 (> sqrt(3) * 2^-23 * 1.5, three components, generous).  +inf disables
// the shortcut.
float far_miss_distance(const rtce_ctx* c, const float pos[3]) {
    const float inf = std::numeric_limits<float>::infinity();
    if (!(c->render_dist > 0.0f) || c->n_spheres > kSmallSceneMax) return inf;
    for (int i = 0; i < c->levels; i++) {
        float isz[2];
        host_image_size(c->levels, i, c->cb.view[0], c->cb.view[1], isz);
        if (!(host_threshold(isz[0]) < c->render_dist)) return inf;
    }
    double k = 0.0;
    for (int i = 0; i < c->n_spheres; i++) {
        const double dx = (double)pos[0] - c->small[i].x, dy = (double)pos[1] - c->small[i].y, dz = (double)pos[2] - c->small[i].z;
        const double d = std::sqrt(dx * dx + dy * dy + dz * dz) + std::fabs((double)c->small[i].w);
        if (!(d <= 1e30)) return inf;
        k = std::max(k, d);
    }
    const double pinf = std::max(std::fabs((double)pos[0]), std::max(std::fabs((double)pos[1]), std::fabs((double)pos[2])));
    const double u = 1.0e-6;
    const double t = (1.001 * ((double)c->render_dist + k) + 1.0 + u * pinf) / (1.0 - u);
    if (!(t < 1e30) || !(pinf < 1e30)) return inf;
    return std::nextafter((float)t, inf);
}

void fill_cams(const rtce_ctx* c, Cam* cams, const rtce_push_constants* pcs, int frames) {
    for (int f = 0; f < frames; f++) {
        memcpy(cams[f].rot, pcs[f].rot, sizeof(float) * 4);
        memcpy(cams[f].pos, pcs[f].pos, sizeof(float) * 3);
        cams[f].pad = c->far_cull ? far_miss_distance(c, cams[f].pos) : std::numeric_limits<float>::infinity();
    }
}

// One level (main.rs:301-314: push constants {iter, imageSize} + dispatch) with the kernel `kind`:
// 0 = per-pixel kernel, 1 = strip kernel, 2 = large-scene kernel, 3 = latency-oriented per-pixel kernel.
// `g` selects the group's ticket block / overflow queue (strip kernel only), `st` the stream.
// `cams` are the frames' cameras on the host, `dcams` the same in device memory.
// `crowd` = how many lanes trace side by side (0 or 1: this chain has the device to itself).
int launch_level(rtce_ctx* c, Lane& L, const Cam* cams, const Cam* dcams, int frames, float* base, int i, int kind, cudaStream_t st,
                 int g, int share = 1, int crowd = 0) {
    int* const tickets = L.d_tickets + (size_t)g * kMaxLevels * kTicketWords;
    float isz[2];
    host_image_size(c->levels, i, c->cb.view[0], c->cb.view[1], isz);   // main.rs:303-305
    int rc[4];
    host_level_rect(c->levels, i, c->rect, c->lw[i], c->lh[i], rc);
    cudaError_t e;
    if (kind == 1) {
        StripParams S{};
        S.out = base + c->loff[i];
        S.parent = i > 0 ? base + c->loff[i - 1] : nullptr;
        S.slot_stride = c->slot_stride;
        S.counters = c->d_counters + 4 * i;
        S.tickets = tickets + kTicketWords * i;
        S.oq = L.d_oq + (size_t)g * kOverflowCap * (10 + kSmallSceneMax);
        S.oq_count = reinterpret_cast<unsigned*>(tickets + kTicketWords * i + 1);
        S.oq_cap = kOverflowCap;
        S.dbg = (c->d_dbg && c->dbg_level == i && g == 0) ? c->d_dbg : nullptr;
        S.w = c->lw[i];
        S.pw = i > 0 ? c->lw[i - 1] : 0;
        S.x0 = rc[0]; S.y0 = rc[1]; S.x1 = rc[2]; S.y1 = rc[3];
        // rows per task: 16, fewer while the level has too few tasks to fill the persistent warps
        int rows = c->strip_rows;
        S.ntx = (rc[2] - rc[0] + 31) / 32;
        while (rows > 2 && (long long)S.ntx * ((rc[3] - rc[1] + rows - 1) / rows) * frames < 2LL * 148 * 32) rows >>= 1;
        S.rows = rows;
        S.cta_share = share;
        // auto: with several lanes in flight a strip launch takes one CTA per SM (two with 2-3 lanes), so that the launches of the
        // other lanes (strips included) find room beside it; alone it takes every resident slot
        S.ctas_per_sm = c->strip_ctas_per_sm > 0 ? c->strip_ctas_per_sm : (crowd > 1 ? (crowd >= 4 ? 1 : 2) : 0);
        S.nty = (rc[3] - rc[1] + rows - 1) / rows;
        S.two_pass = crowd > 1 ? 0 : 1;
        S.frames = frames;
        {
            const long long tiles = (long long)S.ntx * S.nty;
            if (tiles > 0 && tiles < (1ll << 16) && tiles * frames < (1ll << 24)) {
                S.magic_tiles = (1ull << 40) / (unsigned long long)tiles + 1ull;
                S.magic_ntx = (1ull << 40) / (unsigned long long)S.ntx + 1ull;
            }
        }
        S.n_spheres = c->n_spheres;
        S.isx = isz[0]; S.isy = isz[1];
        S.ratx = c->cb.ratio[0]; S.raty = c->cb.ratio[1];
        S.threshold = host_threshold(isz[0]);
        S.render_dist = c->render_dist;
        memcpy(S.sph, c->small, sizeof(S.sph));
        S.cams = dcams;
        {
            MarkScope mark(c, st, 1, i, g);
            e = launch_trace_strip(S, c->count, st);
        }
        if (e == cudaSuccess && c->n_spheres > 0) {   // second half of the dispatch: the long rays
            OverflowParams O{};
            O.out = S.out;
            O.oq = S.oq;
            O.oq_count = S.oq_count;
            O.oq_head = reinterpret_cast<unsigned*>(tickets + kTicketWords * i + 2);
            O.oq_cap = S.oq_cap;
            O.counters = c->d_counters + 4 * i;
            O.n_spheres = c->n_spheres;
            O.threshold = S.threshold;
            O.render_dist = S.render_dist;
            memcpy(O.sph, c->small, sizeof(O.sph));
            c->launches++;
            MarkScope mark(c, st, 2, i, g);
            e = launch_march_overflow(O, c->count, st);
        }
    } else if (kind == 0 || kind == 3) {
        LevelParams P{};
        P.out = base + c->loff[i];
        P.parent = i > 0 ? base + c->loff[i - 1] : nullptr;
        P.slot_stride = c->slot_stride;
        P.counters = c->d_counters + 4 * i;
        P.w = c->lw[i];
        P.pw = i > 0 ? c->lw[i - 1] : 0;
        P.x0 = rc[0]; P.y0 = rc[1]; P.x1 = rc[2]; P.y1 = rc[3];
        P.iter = i;   // main.rs:302
        P.n_spheres = c->n_spheres;
        P.isx = isz[0]; P.isy = isz[1];
        P.ratx = c->cb.ratio[0]; P.raty = c->cb.ratio[1];
        P.threshold = host_threshold(isz[0]);
        P.render_dist = c->render_dist;
        memcpy(P.sph, c->small, sizeof(P.sph));
        P.cams = dcams;
        if (kind == 0) memcpy(P.cam, cams, sizeof(Cam) * (size_t)frames);
        MarkScope mark(c, st, kind == 3 ? 3 : 0, i, g);
        e = kind == 3 ? launch_trace_level_pix(P, frames, c->count, st)
            : c->marcher != 3 ? launch_trace_level_alt(P, frames, c->count, c->marcher, st)
                              : launch_trace_level_small(P, frames, c->count, st);   // main.rs:314 dispatch
    } else {
        LevelParamsLarge G{};
        G.out = base + c->loff[i];
        G.parent = i > 0 ? base + c->loff[i - 1] : nullptr;
        G.slot_stride = c->slot_stride;
        G.counters = c->d_counters + 4 * i;
        G.spheres = c->d_spheres;
        G.w = c->lw[i];
        G.pw = i > 0 ? c->lw[i - 1] : 0;
        G.x0 = rc[0]; G.y0 = rc[1]; G.x1 = rc[2]; G.y1 = rc[3];
        G.iter = i;
        G.n_spheres = c->n_spheres;
        G.isx = isz[0]; G.isy = isz[1];
        G.ratx = c->cb.ratio[0]; G.raty = c->cb.ratio[1];
        G.threshold = host_threshold(isz[0]);
        G.render_dist = c->render_dist;
        memcpy(G.cam, cams, sizeof(Cam) * (size_t)frames);
        if (c->n_spheres > large_reg_capacity()) return fail(c, RTCE_ERR_LIMIT, "more spheres than RTCE_MAX_SPHERES");
        if (c->large_mode != 3) {   // auto = the register kernel (the event-driven form is slower on BASELINE's scene: DESIGN.md)
            e = launch_trace_level_large_reg(G, frames, c->count, st);
        } else {
            // event-driven kernel (+ the register kernel behind it for the rays it hands back); its scratch is sized
            // for the largest launch of the view and allocated on first use (large scenes run in strict stream order)
            const long long max_rays = (long long)c->lw[c->levels - 1] * c->lh[c->levels - 1] * std::min(c->slots, kMaxBatch);
            const size_t need = large_ev_scratch_bytes(c->n_spheres, max_rays, nullptr);
            if (c->large_state_bytes < need) {
                RTCE_CUDA(c, cudaStreamSynchronize(st));
                if (c->large_state) RTCE_CUDA(c, cudaFree(c->large_state));
                c->large_state = nullptr;
                c->large_state_bytes = 0;
                cudaError_t em = cudaMalloc(&c->large_state, need);
                if (em != cudaSuccess) return fail(c, RTCE_ERR_NOMEM, "cudaMalloc(event-driven kernel scratch)", em);
                c->large_state_bytes = need;
            }
            const long long cap_rays = (long long)((c->large_state_bytes - large_ev_scratch_bytes(c->n_spheres, 0, nullptr)) / sizeof(int));
            e = launch_trace_level_large_ev(G, frames, c->count, c->large_state, std::min(max_rays, cap_rays), st);
            c->launches++;
        }
    }
    if (e != cudaSuccess) return fail(c, RTCE_ERR_CUDA, "kernel launch", e);
    c->launches++;
    return RTCE_OK;
}

// Trace `frames` (<= min(slots, kMaxBatch)) frames into slots [slot0, slot0 + frames): the level
// loop of main.rs:300-316.  Kernel plan (RTCE_OPT_KERNEL):
//   strip (auto)  levels with >= strip_min_rays rays per launch -> one strip-kernel launch (+ its
//                 overflow launch) each; the coarse levels below them -> one latency-oriented per-pixel
//                 launch each.  The frames are split in RTCE_OPT_GROUPS groups that run the same launch
//                 chain on their own streams (forked from / joined to the chain's stream), and a strip
//                 launch takes only a share of the CTA slots: the coarse levels, the overflow
//                 launches and the ragged end of every launch are latency-bound (a grazing cone
//                 marches ~100 dependent trips), so the machine is kept busy by whichever groups
//                 are in an issue-bound launch at the moment.
//   per-level     one per-pixel kernel launch per level (RTCE_OPT_KERNEL = 1, RENDER_DIST <= 0, marchers 1 / 2)
//   objCount > 8  one large-scene kernel launch per level
bool strip_plan(const rtce_ctx* c) {   // the strip kernel assumes RENDER_DIST > 0
    return c->n_spheres <= kSmallSceneMax && (c->kernel_mode == 0 || c->kernel_mode == 3) && c->render_dist > 0.0f &&
           c->marcher == 3;
}

// The launch chain of one chunk; `cams` (host) go into the parameters of the kernels that take them
// there, the lane's d_cams (already uploaded) is read by the strip-plan kernels.
int emit_chunk(rtce_ctx* c, int lane, cudaStream_t ls, const Cam* cams, int frames, int slot0, int crowd) {
    Lane& L = c->lanes[lane];
    const Cam* dcams = L.d_cams;
    float* base = window_slot(c, lane, slot0);
    int r;
    if (c->n_spheres > kSmallSceneMax) {
        if (c->marcher != 3) return fail(c, RTCE_ERR_INVALID, "RTCE_OPT_MARCHER 1 / 2 need objCount <= 8");
        for (int i = 0; i < c->levels; i++)
            if ((r = launch_level(c, L, cams, dcams, frames, base, i, 2, ls, 0))) return r;
        return RTCE_OK;
    }
    if (strip_plan(c)) {
        // frame groups: contiguous frame ranges, sizes differ by at most one
        // groups: auto = 2 when this chain has the device to itself, 1 when several lanes trace side by side (they already overlap)
        const int want = c->groups > 0 ? c->groups : (crowd > 1 ? 1 : 2);
        const int ng = std::max(1, std::min(want, frames));
        int nf[kMaxGroups], f0[kMaxGroups];
        for (int g = 0, f = 0; g < ng; g++) {
            nf[g] = frames / ng + (g < frames % ng ? 1 : 0);
            f0[g] = f;
            f += nf[g];
        }
        // first level that goes to the strip kernel (level sizes grow ~4x per level), by the largest group
        int k = c->levels;
        while (k > 0) {
            int rc[4];
            host_level_rect(c->levels, k - 1, c->rect, c->lw[k - 1], c->lh[k - 1], rc);
            if ((long long)(rc[2] - rc[0]) * (rc[3] - rc[1]) * nf[0] < c->strip_min_rays) break;
            if ((unsigned long long)nf[0] * c->slot_stride >= (1ull << 32)) break;   // strip queue offsets are 32-bit
            k--;
        }
        // With one group the launches run on the chain's stream; with several, every group has
        // its own stream, forked from and joined to the chain's stream.
        cudaStream_t st[kMaxGroups];
        for (int g = 0; g < ng; g++) st[g] = ng == 1 ? ls : L.grp_stream[g];
        if (ng > 1) {
            RTCE_CUDA(c, cudaEventRecord(L.ev_fork, ls));
            for (int g = 0; g < ng; g++) RTCE_CUDA(c, cudaStreamWaitEvent(st[g], L.ev_fork, 0));
        }
        // a strip launch takes a share of the resident CTA slots, so that the strip launches of up to
        // three groups (and the small launches of the others) are resident side by side
        const int share = std::min(ng, 3);
        for (int g = 0; g < ng; g++) {
            if (k < c->levels)   // task counters / overflow cursors of this group's strip launches
                RTCE_CUDA(c, cudaMemsetAsync(L.d_tickets + (size_t)g * kMaxLevels * kTicketWords, 0, sizeof(int) * kMaxLevels * kTicketWords, st[g]));
        }
        for (int i = 0; i < k; i++)   // the coarse levels
            for (int g = 0; g < ng; g++)
                if ((r = launch_level(c, L, cams + f0[g], dcams + f0[g], nf[g], base + (size_t)f0[g] * c->slot_stride, i, 3, st[g], g, share)))
                    return r;
        for (int i = k; i < c->levels; i++)
            for (int g = 0; g < ng; g++)
                if ((r = launch_level(c, L, cams + f0[g], dcams + f0[g], nf[g], base + (size_t)f0[g] * c->slot_stride, i, 1, st[g], g, share,
                                      crowd)))
                    return r;
        if (ng > 1) {
            for (int g = 0; g < ng; g++) {
                RTCE_CUDA(c, cudaEventRecord(L.ev_join[g], st[g]));
                RTCE_CUDA(c, cudaStreamWaitEvent(ls, L.ev_join[g], 0));
            }
        }
        return RTCE_OK;
    }
    for (int i = 0; i < c->levels; i++)
        if ((r = launch_level(c, L, cams, dcams, frames, base, i, 0, ls, 0))) return r;
    return RTCE_OK;
}

bool graphable(const rtce_ctx* c) {
    return c->use_graph && strip_plan(c) && !c->count && !c->mark_launches && c->dbg_level < 0;
}

void drop_stale_graphs(rtce_ctx* c, Lane& L) {
    for (auto it = L.graphs.begin(); it != L.graphs.end();) {
        if (it->version != c->plan_version) {
            cudaGraphExecDestroy(it->exec);
            it = L.graphs.erase(it);
        } else {
            ++it;
        }
    }
}

const ChunkGraph* find_graph(const Lane& L, int frames, int slot0, int crowd) {
    for (const auto& g : L.graphs)
        if (g.frames == frames && g.slot0 == slot0 && g.lanes == crowd) return &g;
    return nullptr;
}

// Capture, instantiate and upload the launch graph of chunk shape (frames, slot0, crowd) on lane `lane`, whose
// chain runs on `ls` (idempotent).  The camera values are not part of the graph: they travel through the lane's
// device buffer, so the capture needs no cameras (the kernels that take them in their parameters are not
// graphable plans).
int capture_graph(rtce_ctx* c, int lane, cudaStream_t ls, int frames, int slot0, int crowd) {
    Lane& L = c->lanes[lane];
    drop_stale_graphs(c, L);
    if (find_graph(L, frames, slot0, crowd)) return RTCE_OK;
    const uint64_t l0 = c->launches;
    Cam dummy[kMaxBatch] = {};
    RTCE_CUDA(c, cudaStreamBeginCapture(ls, cudaStreamCaptureModeRelaxed));
    const int r = emit_chunk(c, lane, ls, dummy, frames, slot0, crowd);
    cudaGraph_t graph = nullptr;
    const cudaError_t ec = cudaStreamEndCapture(ls, &graph);
    const uint64_t n_launch = c->launches - l0;
    c->launches = l0;   // a capture launches nothing
    if (r || ec != cudaSuccess) {
        if (graph) cudaGraphDestroy(graph);
        cudaGetLastError();
        return r ? r : fail(c, RTCE_ERR_CUDA, "cudaStreamEndCapture", ec);
    }
    ChunkGraph cg{frames, slot0, crowd, c->plan_version, nullptr, n_launch};
    const cudaError_t ei = cudaGraphInstantiate(&cg.exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ei != cudaSuccess) return fail(c, RTCE_ERR_CUDA, "cudaGraphInstantiate", ei);
    cudaGraphUpload(cg.exec, ls);   // stream-ordered; the first replay then costs what every replay costs
    if (L.graphs.size() >= 32) {   // a caller cycling through many chunk shapes: start over
        for (auto& g : L.graphs) cudaGraphExecDestroy(g.exec);
        L.graphs.clear();
    }
    L.graphs.push_back(cg);
    return RTCE_OK;
}

// The stream lane `l` chains on when `nl` lanes rotate: its own, or the context's in strict stream order.
inline cudaStream_t lane_stream(rtce_ctx* c, int l, int nl) { return nl > 1 ? c->lanes[l].st : c->stream; }

// Graphs of a chunk shape on EVERY lane of the rotation, both for a chain that has the device to itself and for
// chains side by side: the first call that meets a new shape pays for all lanes, no later call instantiates.
int capture_shape(rtce_ctx* c, int nl, int frames, int slot0) {
    for (int l = 0; l < nl; l++) {
        int r;
        if ((r = capture_graph(c, l, lane_stream(c, l, nl), frames, slot0, 0))) return r;
        if (nl > 1 && (r = capture_graph(c, l, lane_stream(c, l, nl), frames, slot0, nl))) return r;
    }
    return RTCE_OK;
}

// Trace one chunk on lane `lane`, whose chain runs on stream `ls`: stage its cameras (pinned ring ->
// the lane's device buffer, stream-ordered), then either issue the launch chain or — strip plan, no
// counters / diagnostics — replay the CUDA graph captured from it (RTCE_OPT_GRAPH): ~90 launches on
// up to 8 streams cost one graph launch on the host.
int trace_chunk(rtce_ctx* c, int lane, cudaStream_t ls, const rtce_push_constants* pcs, int frames, int slot0, int nl, int crowd) {
    Lane& L = c->lanes[lane];
    const int ring = (int)(L.cam_seq++ % kCamRing);
    RTCE_CUDA(c, cudaEventSynchronize(L.ev_cam[ring]));   // the copy that last used this entry is done
    Cam* cams = L.h_cams + (size_t)ring * kMaxBatch;
    fill_cams(c, cams, pcs, frames);
    RTCE_CUDA(c, cudaMemcpyAsync(L.d_cams, cams, sizeof(Cam) * (size_t)frames, cudaMemcpyHostToDevice, ls));
    RTCE_CUDA(c, cudaEventRecord(L.ev_cam[ring], ls));
    if (!graphable(c)) return emit_chunk(c, lane, ls, cams, frames, slot0, crowd);
    drop_stale_graphs(c, L);
    const ChunkGraph* g = find_graph(L, frames, slot0, crowd);
    if (!g) {
        int r = capture_shape(c, nl, frames, slot0);
        if (r) return r;
        g = find_graph(L, frames, slot0, crowd);
        if (!g) return fail(c, RTCE_ERR_STATE, "launch graph missing after capture");
    }
    RTCE_CUDA(c, cudaGraphLaunch(g->exec, ls));
    c->launches += g->launches;
    return RTCE_OK;
}

// `keep_host_copies`: the caller is rtce_render_to_host(_async) itself, which orders its own reuse of window 0 behind
// the copies half by half; every other tracer first makes the context's stream wait for copies that may still read it.
int check_ready(rtce_ctx* c, const void* p, int frames, bool keep_host_copies = false) {
    if (!c) return RTCE_ERR_INVALID;
    if (!p || frames < 1) return fail(c, RTCE_ERR_INVALID, "null push constants or frames < 1");
    if (!c->have_scene || !c->have_view) return fail(c, RTCE_ERR_STATE, "set scene and view before tracing");
    if (c->marcher != 3 && (c->n_spheres < 1 || c->n_spheres > kSmallSceneMax))
        return fail(c, RTCE_ERR_INVALID, "RTCE_OPT_MARCHER 1 / 2 need 1 <= objCount <= 8 (algorithm 1 reads objs[0])");
    if (c->host_copies_pending && !keep_host_copies) {
        DeviceGuard g(c->device);
        for (int i = 0; i < 2; i++)
            if (c->half_used[i]) RTCE_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_copied[i], 0));
        c->host_copies_pending = false;
    }
    return RTCE_OK;
}

int begin_counters(rtce_ctx* c) {
    c->launches = 0;
    if (c->count) RTCE_CUDA(c, cudaMemsetAsync(c->d_counters, 0, 4 * kMaxLevels * sizeof(unsigned long long), c->stream));
    return RTCE_OK;
}

int chunk_size(const rtce_ctx* c) { return std::min(c->slots, kMaxBatch); }

// Lanes a call may use right now: counters and the debug timelines keep everything on the context's stream.
int active_lanes(const rtce_ctx* c, int want) {
    // large scenes: a launch is hundreds of milliseconds of a full device and the kernels share one scratch area
    return (c->count || c->mark_launches || c->dbg_level >= 0 || c->n_spheres > kSmallSceneMax) ? 1 : std::min(want, kMaxLanes);
}

// Windows for `nl` lanes, falling back to strict stream order when the device has no room for them.
int lanes_with_windows(rtce_ctx* c, int nl) {
    int r = ensure_pyramid(c, nl);
    if (r == RTCE_ERR_NOMEM && nl > 1) {
        nl = 1;
        r = ensure_pyramid(c, 1);
    }
    return r ? r : nl;
}

// How many lanes other than `lane` still have work on the device.
int busy_lanes(rtce_ctx* c, int nl, int lane) {
    int n = 0;
    for (int l = 0; l < nl; l++) {
        Lane& L = c->lanes[l];
        if (l != lane && L.ready && L.pending && cudaEventQuery(L.ev_done) == cudaErrorNotReady) n++;
    }
    cudaGetLastError();
    return n;
}

// Slot of the window in which chunk [f0, f0 + n) of a batch starts: frame k lands in slot k % slots.
inline int chunk_slot0(const rtce_ctx* c, int f0, int n) { return f0 % c->slots + n <= c->slots ? f0 % c->slots : 0; }

}  // namespace

extern "C" {

int rtce_trace_batch(rtce_ctx* c, const rtce_push_constants* pcs, int frames) {
    int r = check_ready(c, pcs, frames);
    if (r) return r;
    DeviceGuard g(c->device);
    int nl = lanes_with_windows(c, active_lanes(c, c->in_flight));
    if (nl < 0) return nl;
    c->lanes_active = nl;
    if ((r = begin_counters(c))) return r;
    int lane = 0, crowd = 0;
    cudaStream_t ls = c->stream;
    if (nl > 1) {
        // frames in flight: this batch goes to the next lane (its own window of slots and its own streams) and
        // starts after everything the caller has enqueued so far — not after the batches on the other lanes
        lane = c->next_lane % nl;
        c->next_lane = (lane + 1) % nl;
        ls = c->lanes[lane].st;
        RTCE_CUDA(c, cudaEventRecord(c->ev_user, c->stream));
        RTCE_CUDA(c, cudaStreamWaitEvent(ls, c->ev_user, 0));
        if ((r = comm_wait_lane(c, lane, ls))) return r;   // rtce_gather_frames still reading / filling this window
        // a caller that waits for every batch gets the launch shape of a chain that owns the device (two frame
        // groups, every CTA slot); batches issued back to back get the shape that leaves room for each other
        crowd = busy_lanes(c, nl, lane) > 0 ? nl : 0;
    } else if ((r = flush_lanes(c))) {
        return r;
    }
    const int cs = chunk_size(c);
    for (int f0 = 0; f0 < frames; f0 += cs) {
        const int n = std::min(cs, frames - f0);
        if ((r = trace_chunk(c, lane, ls, pcs + f0, n, chunk_slot0(c, f0, n), nl, crowd))) return r;
    }
    c->last_lane = lane;
    if (nl > 1) {
        Lane& L = c->lanes[lane];
        RTCE_CUDA(c, cudaEventRecord(L.ev_done, ls));
        L.pending = true;
        // the context's stream catches up with the batch whose window the NEXT call overwrites
        Lane& O = c->lanes[c->next_lane];
        if (O.ready && O.pending) {
            RTCE_CUDA(c, cudaStreamWaitEvent(c->stream, O.ev_done, 0));
            O.pending = false;
        }
    }
    return RTCE_OK;
}

int rtce_trace(rtce_ctx* c, const rtce_push_constants* pc) { return rtce_trace_batch(c, pc, 1); }

int rtce_flush(rtce_ctx* c) {
    if (!c) return RTCE_ERR_INVALID;
    DeviceGuard g(c->device);
    return flush_lanes(c);
}

// Lanes rtce_trace_accumulate rotates `frames` frames over.
static int accum_lanes(rtce_ctx* c, int frames) {
    const int cs = chunk_size(c);
    int nl = active_lanes(c, std::max(c->in_flight, kAccumLanes));
    return std::max(1, std::min(nl, (frames + cs - 1) / cs));
}

int rtce_prepare(rtce_ctx* c, int frames, int accumulate) {
    if (!c || frames < 1) return fail(c, RTCE_ERR_INVALID, "rtce_prepare: frames < 1");
    if (!c->have_scene || !c->have_view) return fail(c, RTCE_ERR_STATE, "set scene and view before rtce_prepare");
    DeviceGuard g(c->device);
    int nl = lanes_with_windows(c, accumulate ? accum_lanes(c, frames) : active_lanes(c, c->in_flight));
    if (nl < 0) return nl;
    if (!graphable(c)) return RTCE_OK;
    const int cs = chunk_size(c);
    const uint64_t keep = c->launches;
    for (int f0 = 0; f0 < frames; f0 += cs) {
        const int n = std::min(cs, frames - f0);
        int r = capture_shape(c, nl, n, accumulate ? 0 : chunk_slot0(c, f0, n));   // idempotent: repeated shapes cost a lookup
        if (r) return r;
    }
    c->launches = keep;
    return RTCE_OK;
}

int rtce_trace_accumulate(rtce_ctx* c, const rtce_push_constants* pcs, int frames) {
    int r = check_ready(c, pcs, frames);
    if (r) return r;
    DeviceGuard g(c->device);
    const int cs = chunk_size(c);
    int nl = lanes_with_windows(c, accum_lanes(c, frames));
    if (nl < 0) return nl;
    c->lanes_active = nl;
    if ((r = flush_lanes(c))) return r;
    if ((r = begin_counters(c))) return r;
    const int L = c->levels - 1;
    const size_t npx = (size_t)c->lw[L] * (size_t)c->lh[L];
    if (c->d_accum_floats != npx) {
        if (c->d_accum) {
            RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
            RTCE_CUDA(c, cudaFree(c->d_accum));
            c->d_accum = nullptr;
            c->d_accum_floats = 0;
        }
        cudaError_t e = cudaMalloc(&c->d_accum, npx * sizeof(float));
        if (e != cudaSuccess) return fail(c, RTCE_ERR_NOMEM, "cudaMalloc(accum)", e);
        c->d_accum_floats = npx;
        c->accum_rows[0] = 0;
        c->accum_rows[1] = c->lh[L];   // everything is stale: cleared below
    }
    // only the rows of the region of interest are accumulated (whole rows: the image is row-major); the rows
    // outside it are defined to be 0: rows an earlier call wrote under another rect are cleared here
    int rc[4];
    host_level_rect(c->levels, L, c->rect, c->lw[L], c->lh[L], rc);
    if (c->accum_rows[0] < rc[1]) {
        const int y1 = std::min(c->accum_rows[1], rc[1]);
        if (y1 > c->accum_rows[0])
            RTCE_CUDA(c, cudaMemsetAsync(c->d_accum + (size_t)c->accum_rows[0] * c->lw[L], 0, (size_t)(y1 - c->accum_rows[0]) * c->lw[L] * sizeof(float), c->stream));
    }
    if (c->accum_rows[1] > rc[3]) {
        const int y0 = std::max(c->accum_rows[0], rc[3]);
        if (c->accum_rows[1] > y0)
            RTCE_CUDA(c, cudaMemsetAsync(c->d_accum + (size_t)y0 * c->lw[L], 0, (size_t)(c->accum_rows[1] - y0) * c->lw[L] * sizeof(float), c->stream));
    }
    c->accum_rows[0] = rc[1];
    c->accum_rows[1] = rc[3];
    const size_t roi_off = (size_t)rc[1] * (size_t)c->lw[L];
    const size_t roi_n = (size_t)(rc[3] - rc[1]) * (size_t)c->lw[L];
    // The chunks rotate over `nl` lanes (own streams, own window of slots): up to nl chunks are traced side by
    // side — one lane's latency-bound coarse levels under another's final level — while the accumulation runs
    // on the context's stream, chunk after chunk, in frame order.
    if (nl > 1) RTCE_CUDA(c, cudaEventRecord(c->ev_user, c->stream));
    bool first = true;
    int j = 0, lane = 0;
    for (int f0 = 0; f0 < frames; f0 += cs, j++) {
        const int n = std::min(cs, frames - f0);
        lane = nl > 1 ? j % nl : 0;
        Lane& W = c->lanes[lane];
        cudaStream_t ls = nl > 1 ? W.st : c->stream;
        if (nl > 1) RTCE_CUDA(c, cudaStreamWaitEvent(ls, j < nl ? c->ev_user : W.ev_free, 0));   // window free again
        if ((r = trace_chunk(c, lane, ls, pcs + f0, n, 0, nl, nl > 1 ? nl : 0))) return r;
        if (nl > 1) {
            RTCE_CUDA(c, cudaEventRecord(W.ev_done, ls));
            RTCE_CUDA(c, cudaStreamWaitEvent(c->stream, W.ev_done, 0));
            W.pending = false;
        }
        // one pass over the chunk's final levels, frames added in order: the sum is order-defined
        const float* img = window_slot(c, lane, 0) + c->loff[L] + roi_off;
        cudaError_t e = launch_accumulate(c->d_accum + roi_off, img, c->slot_stride, n, roi_n, first, c->stream);
        if (e != cudaSuccess) return fail(c, RTCE_ERR_CUDA, "accumulate launch", e);
        c->launches++;
        if (nl > 1) RTCE_CUDA(c, cudaEventRecord(W.ev_free, c->stream));
        first = false;
    }
    c->last_lane = lane;
    cudaError_t e = launch_scale(c->d_accum + roi_off, roi_n, (float)frames, c->stream);
    if (e != cudaSuccess) return fail(c, RTCE_ERR_CUDA, "scale launch", e);
    c->launches++;
    return RTCE_OK;
}

int rtce_accum_info(const rtce_ctx* c, int* w, int* h, float** dptr) {
    if (!c) return RTCE_ERR_INVALID;
    if (!c->have_view || !c->d_accum) return RTCE_ERR_STATE;
    if (w) *w = c->lw[c->levels - 1];
    if (h) *h = c->lh[c->levels - 1];
    if (dptr) *dptr = c->d_accum;
    return RTCE_OK;
}

int rtce_sync(rtce_ctx* c) {
    if (!c) return RTCE_ERR_INVALID;
    DeviceGuard g(c->device);
    return quiesce(c);
}

int rtce_readback(rtce_ctx* c, int level, int slot, float* host, size_t pitch_bytes) {
    if (!c || !host) return fail(c, RTCE_ERR_INVALID, "rtce_readback: null");
    int w, h;
    float* d;
    int r = rtce_level_info(c, level, slot, &w, &h, &d);
    if (r) return fail(c, r, "rtce_readback: bad level/slot or view not set");
    DeviceGuard g(c->device);
    if ((r = flush_lanes(c))) return r;
    const size_t row = (size_t)w * sizeof(float);
    if (pitch_bytes == 0) pitch_bytes = row;
    if (pitch_bytes < row) return fail(c, RTCE_ERR_INVALID, "rtce_readback: pitch < row");
    RTCE_CUDA(c, cudaMemcpy2DAsync(host, pitch_bytes, d, row, row, (size_t)h, cudaMemcpyDeviceToHost, c->stream));
    RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
    return RTCE_OK;
}

int rtce_readback_accum(rtce_ctx* c, float* host, size_t pitch_bytes) {
    if (!c || !host) return fail(c, RTCE_ERR_INVALID, "rtce_readback_accum: null");
    if (!c->have_view || !c->d_accum) return fail(c, RTCE_ERR_STATE, "rtce_readback_accum: nothing accumulated");
    DeviceGuard g(c->device);
    const int w = c->lw[c->levels - 1], h = c->lh[c->levels - 1];
    const size_t row = (size_t)w * sizeof(float);
    if (pitch_bytes == 0) pitch_bytes = row;
    if (pitch_bytes < row) return fail(c, RTCE_ERR_INVALID, "rtce_readback_accum: pitch < row");
    RTCE_CUDA(c, cudaMemcpy2DAsync(host, pitch_bytes, c->d_accum, row, row, (size_t)h, cudaMemcpyDeviceToHost, c->stream));
    RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
    return RTCE_OK;
}

// ---- presentation shading pass (fragment.glsl; the draw of main.rs:318-338) --------------------
// The fragments are the WINDOW's pixels (cs.view); the final level image may be larger (main.rs:209-213).
static inline int win_w(const rtce_ctx* c) { return std::min((int)c->cb.view[0], c->lw[c->levels - 1]); }
static inline int win_h(const rtce_ctx* c) { return std::min((int)c->cb.view[1], c->lh[c->levels - 1]); }

static int ensure_color(rtce_ctx* c) {
    const size_t px = (size_t)win_w(c) * (size_t)win_h(c);
    if (c->d_color && c->d_color_px == px && c->color_slots >= c->slots) return RTCE_OK;
    RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
    if (c->d_color) cudaFree(c->d_color);
    if (c->d_rgba8) cudaFree(c->d_rgba8);
    c->d_color = nullptr;
    c->d_rgba8 = nullptr;
    if (cudaMalloc(&c->d_color, px * (size_t)c->slots * sizeof(float4)) != cudaSuccess ||
        cudaMalloc(&c->d_rgba8, px * (size_t)c->slots * sizeof(uchar4)) != cudaSuccess) {
        cudaGetLastError();
        return fail(c, RTCE_ERR_NOMEM, "cudaMalloc(colour slots)");
    }
    c->d_color_px = px;
    c->color_slots = c->slots;
    RTCE_CUDA(c, cudaMemsetAsync(c->d_color, 0, px * (size_t)c->slots * sizeof(float4), c->stream));
    if (!c->d_shade_cams) {
        if (cudaMalloc(&c->d_shade_cams, sizeof(Cam) * kMaxBatch) != cudaSuccess ||
            cudaMallocHost(&c->h_shade_cams, sizeof(Cam) * kMaxBatch) != cudaSuccess ||
            cudaEventCreateWithFlags(&c->ev_shade_cams, cudaEventDisableTiming) != cudaSuccess ||
            cudaMalloc(&c->d_shade_counters, 4 * sizeof(unsigned long long)) != cudaSuccess)
            return fail(c, RTCE_ERR_NOMEM, "shading pass staging");
    }
    return RTCE_OK;
}

// Shade the final level of slots [slot0, slot0 + frames) of window c->last_lane into the same colour slots.
static int shade_slots(rtce_ctx* c, const rtce_push_constants* pcs, int frames, int slot0) {
    const int L = c->levels - 1;
    ShadeParams S{};
    S.depth_stride = c->slot_stride;
    S.out_stride = c->d_color_px;
    S.counters = c->count ? c->d_shade_counters : nullptr;
    S.cams = c->d_shade_cams;
    S.w = c->lw[L];
    S.h = c->lh[L];
    S.out_w = win_w(c);
    int rc[4];
    host_level_rect(c->levels, L, c->rect, S.w, S.h, rc);
    S.x0 = rc[0]; S.y0 = rc[1]; S.x1 = std::min(rc[2], win_w(c)); S.y1 = std::min(rc[3], win_h(c));
    S.vw = c->cb.view[0]; S.vh = c->cb.view[1];
    S.ratx = c->cb.ratio[0]; S.raty = c->cb.ratio[1];
    S.render_dist = c->render_dist;
    // std140 MutableData (fragment.glsl:18-25): counts at 0 / 4 / 8, mats[8] x 32 B at 16, objs[8] x 16 B at 272,
    // lights[8] x 32 B at 400
    uint32_t nobj = 0, nlight = 0;
    memcpy(&nobj, c->md + 4, 4);
    memcpy(&nlight, c->md + 8, 4);
    if (nobj < 1) return fail(c, RTCE_ERR_STATE, "rtce_shade_batch: the shader reads objs[0]; objCount is 0");
    S.n_obj = (int)nobj;
    S.n_light = (int)nlight;
    for (int i = 0; i < kSmallSceneMax; i++) {
        float m[7], o[4], l[7];
        memcpy(m, c->md + 16 + 32 * i, sizeof(m));
        memcpy(o, c->md + 272 + 16 * i, sizeof(o));
        memcpy(l, c->md + 400 + 32 * i, sizeof(l));
        S.matc[i] = make_float4(m[0], m[1], m[2], m[5]);   // color, shine
        S.amb[i] = m[6];
        S.obj[i] = make_float4(o[0], o[1], o[2], o[3]);
        S.lpos[i] = make_float4(l[0], l[1], l[2], 0.0f);
        S.lcol[i] = make_float4(l[4], l[5], l[6], 0.0f);
    }
    for (int f0 = 0; f0 < frames; f0 += kMaxBatch) {
        const int n = std::min(kMaxBatch, frames - f0);
        RTCE_CUDA(c, cudaEventSynchronize(c->ev_shade_cams));   // the previous upload has left the staging buffer
        fill_cams(c, c->h_shade_cams, pcs + f0, n);
        RTCE_CUDA(c, cudaMemcpyAsync(c->d_shade_cams, c->h_shade_cams, sizeof(Cam) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
        RTCE_CUDA(c, cudaEventRecord(c->ev_shade_cams, c->stream));
        S.depth = window_slot(c, c->last_lane, slot0 + f0) + c->loff[L];
        S.out = c->d_color + (size_t)(slot0 + f0) * c->d_color_px;
        cudaError_t e = launch_shade(S, n, c->count, c->stream);
        if (e != cudaSuccess) return fail(c, RTCE_ERR_CUDA, "shading kernel launch", e);
        c->launches++;
    }
    return RTCE_OK;
}

int rtce_shade_batch(rtce_ctx* c, const rtce_push_constants* pcs, int frames) {
    int r = check_ready(c, pcs, frames);
    if (r) return r;
    if (!c->have_materials) return fail(c, RTCE_ERR_STATE, "rtce_shade_batch: materials and lights come with rtce_set_scene");
    if (frames > c->slots) return fail(c, RTCE_ERR_LIMIT, "rtce_shade_batch: more frames than frame slots");
    if (!c->d_pyr) return fail(c, RTCE_ERR_STATE, "rtce_shade_batch: nothing traced yet");
    DeviceGuard g(c->device);
    if ((r = flush_lanes(c))) return r;
    if ((r = ensure_color(c))) return r;
    if (c->count) RTCE_CUDA(c, cudaMemsetAsync(c->d_shade_counters, 0, 4 * sizeof(unsigned long long), c->stream));
    return shade_slots(c, pcs, frames, 0);
}

int rtce_color_info(const rtce_ctx* c, int slot, int* w, int* h, float** dptr) {
    if (!c) return RTCE_ERR_INVALID;
    if (!c->have_view || !c->d_color) return RTCE_ERR_STATE;
    if (slot < 0 || slot >= c->color_slots) return RTCE_ERR_INVALID;
    if (w) *w = win_w(c);
    if (h) *h = win_h(c);
    if (dptr) *dptr = reinterpret_cast<float*>(c->d_color + (size_t)slot * c->d_color_px);
    return RTCE_OK;
}

int rtce_readback_color(rtce_ctx* c, int slot, float* host_rgba) {
    if (!c || !host_rgba) return fail(c, RTCE_ERR_INVALID, "rtce_readback_color: null");
    float* d = nullptr;
    int r = rtce_color_info(c, slot, nullptr, nullptr, &d);
    if (r) return fail(c, r, "rtce_readback_color: nothing shaded or bad slot");
    DeviceGuard g(c->device);
    RTCE_CUDA(c, cudaMemcpyAsync(host_rgba, d, c->d_color_px * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
    RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
    return RTCE_OK;
}

int rtce_readback_color_rgba8(rtce_ctx* c, int slot, unsigned char* host_rgba8) {
    if (!c || !host_rgba8) return fail(c, RTCE_ERR_INVALID, "rtce_readback_color_rgba8: null");
    float* d = nullptr;
    int r = rtce_color_info(c, slot, nullptr, nullptr, &d);
    if (r) return fail(c, r, "rtce_readback_color_rgba8: nothing shaded or bad slot");
    DeviceGuard g(c->device);
    uchar4* d8 = c->d_rgba8 + (size_t)slot * c->d_color_px;
    cudaError_t e = launch_rgba8(reinterpret_cast<const float4*>(d), d8, c->d_color_px, c->stream);
    if (e != cudaSuccess) return fail(c, RTCE_ERR_CUDA, "rgba8 kernel launch", e);
    c->launches++;
    RTCE_CUDA(c, cudaMemcpyAsync(host_rgba8, d8, c->d_color_px * sizeof(uchar4), cudaMemcpyDeviceToHost, c->stream));
    RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
    return RTCE_OK;
}

int rtce_get_shade_counters(rtce_ctx* c, uint64_t out[4]) {
    if (!c || !out) return RTCE_ERR_INVALID;
    out[0] = out[1] = out[2] = out[3] = 0;
    if (!c->d_shade_counters) return RTCE_OK;
    DeviceGuard g(c->device);
    unsigned long long h[4] = {0, 0, 0, 0};
    if (c->count) RTCE_CUDA(c, cudaMemcpyAsync(h, c->d_shade_counters, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
    for (int i = 0; i < 4; i++) out[i] = h[i];
    return RTCE_OK;
}

int rtce_host_alloc(void** out, size_t nbytes) {
    if (!out) return RTCE_ERR_INVALID;
    return cudaMallocHost(out, nbytes) == cudaSuccess ? RTCE_OK : RTCE_ERR_NOMEM;
}

int rtce_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? RTCE_OK : RTCE_ERR_CUDA; }

// End of a rtce_render_*_to_host(_async) call: block for everything (synchronous form), or record this call's
// completion on the copy stream and block for the PREVIOUS asynchronous call only.
static int finish_host_call(rtce_ctx* c, bool wait) {
    if (wait) {
        RTCE_CUDA(c, cudaStreamSynchronize(c->copy_stream));
        RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
        c->host_copies_pending = false;
        c->host_calls = 0;
        return RTCE_OK;
    }
    const unsigned k = c->host_calls++;
    RTCE_CUDA(c, cudaEventRecord(c->ev_host_done[k & 1], c->copy_stream));
    if (k > 0) RTCE_CUDA(c, cudaEventSynchronize(c->ev_host_done[(k - 1) & 1]));
    return RTCE_OK;
}

// `wait`: synchronise before returning (rtce_render_to_host).  Otherwise (rtce_render_to_host_async) everything of this
// call is enqueued, then the host waits for the PREVIOUS asynchronous call only: when call k returns, the frames of call
// k - 1 are in their host buffer while the device works on call k — a caller with two host buffers never drains the
// pipeline between batches (the synchronous form idles the copy engine while the first half of every batch is traced,
// and the SMs while the last half is copied).
static int render_to_host_impl(rtce_ctx* c, const rtce_push_constants* pcs, int frames, float* host_out, bool wait) {
    int r = check_ready(c, pcs, frames, true);
    if (r) return r;
    if (!host_out) return fail(c, RTCE_ERR_INVALID, "rtce_render_to_host: null output");
    DeviceGuard g(c->device);
    if ((r = ensure_pyramid(c))) return r;
    if ((r = flush_lanes(c))) return r;
    if ((r = begin_counters(c))) return r;
    c->last_lane = 0;
    c->lanes_active = 1;
    const int L = c->levels - 1;
    const int w = c->lw[L], h = c->lh[L];
    int rc[4];
    host_level_rect(c->levels, L, c->rect, w, h, rc);
    const size_t npx = (size_t)w * (size_t)h;
    // ping-pong two halves of the slot window so the D2H copy of one half overlaps the trace of the other
    const int half = std::max(1, std::min(c->slots / 2, kMaxBatch));
    const bool pingpong = c->slots >= 2;
    int phase = 0;
    for (int f0 = 0; f0 < frames; f0 += half, phase ^= 1) {
        const int n = std::min(half, frames - f0);
        const int slot0 = pingpong ? phase * half : 0;
        if (c->half_used[phase]) RTCE_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_copied[phase], 0));   // half is free again
        if (!pingpong && c->half_used[phase ^ 1]) RTCE_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_copied[phase ^ 1], 0));   // one slot: both
        if ((r = trace_chunk(c, 0, c->stream, pcs + f0, n, slot0, 1, 0))) return r;
        RTCE_CUDA(c, cudaEventRecord(c->ev_traced[phase], c->stream));
        RTCE_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_traced[phase], 0));
        for (int k = 0; k < n; k++) {
            const float* src = window_slot(c, 0, slot0 + k) + c->loff[L] + (size_t)rc[1] * w;
            float* dst = host_out + (size_t)(f0 + k) * npx + (size_t)rc[1] * w;
            RTCE_CUDA(c, cudaMemcpyAsync(dst, src, (size_t)(rc[3] - rc[1]) * w * sizeof(float), cudaMemcpyDeviceToHost,
                                         c->copy_stream));
        }
        RTCE_CUDA(c, cudaEventRecord(c->ev_copied[phase], c->copy_stream));
        c->half_used[phase] = true;
        c->host_copies_pending = true;
    }
    return finish_host_call(c, wait);
}

int rtce_render_to_host(rtce_ctx* c, const rtce_push_constants* pcs, int frames, float* host_out) {
    return render_to_host_impl(c, pcs, frames, host_out, true);
}

int rtce_render_to_host_async(rtce_ctx* c, const rtce_push_constants* pcs, int frames, float* host_out) {
    return render_to_host_impl(c, pcs, frames, host_out, false);
}

// main.rs:886-914 headless: compute half (trace), render-pass half (shade), "present" (RGBA8 UNORM to host memory).
// Chunks of min(slots, kMaxBatch) frames on the context's stream; the D2H copy of a chunk's RGBA8 images runs on the
// copy stream under the tracing of the next chunk (the RGBA8 staging holds one image per slot: two chunks of
// half the window alternate, as in rtce_render_to_host).
static int render_rgba8_impl(rtce_ctx* c, const rtce_push_constants* pcs, int frames, unsigned char* host_rgba8, bool wait) {
    int r = check_ready(c, pcs, frames, true);
    if (r) return r;
    if (!host_rgba8) return fail(c, RTCE_ERR_INVALID, "rtce_render_rgba8_to_host: null output");
    if (!c->have_materials) return fail(c, RTCE_ERR_STATE, "rtce_render_rgba8_to_host: materials and lights come with rtce_set_scene");
    DeviceGuard g(c->device);
    if ((r = ensure_pyramid(c))) return r;
    if ((r = flush_lanes(c))) return r;
    if ((r = ensure_color(c))) return r;
    const int half = std::max(1, std::min(c->slots / 2, kMaxBatch));
    const bool pingpong = c->slots >= 2;
    const size_t img_bytes = c->d_color_px * sizeof(uchar4);
    int phase = 0;
    for (int f0 = 0; f0 < frames; f0 += half, phase ^= 1) {
        const int n = std::min(half, frames - f0);
        const int slot0 = pingpong ? phase * half : 0;
        if (c->half_used[phase]) RTCE_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_copied[phase], 0));   // half is free again
        if (!pingpong && c->half_used[phase ^ 1]) RTCE_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_copied[phase ^ 1], 0));   // one slot: both
        c->last_lane = 0;
        c->lanes_active = 1;
        if ((r = trace_chunk(c, 0, c->stream, pcs + f0, n, slot0, 1, 0))) return r;
        if ((r = shade_slots(c, pcs + f0, n, slot0))) return r;
        cudaError_t e = launch_rgba8(c->d_color + (size_t)slot0 * c->d_color_px, c->d_rgba8 + (size_t)slot0 * c->d_color_px,
                                     c->d_color_px * (size_t)n, c->stream);
        if (e != cudaSuccess) return fail(c, RTCE_ERR_CUDA, "rgba8 kernel launch", e);
        c->launches++;
        RTCE_CUDA(c, cudaEventRecord(c->ev_traced[phase], c->stream));
        RTCE_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_traced[phase], 0));
        RTCE_CUDA(c, cudaMemcpyAsync(host_rgba8 + (size_t)f0 * img_bytes, c->d_rgba8 + (size_t)slot0 * c->d_color_px, img_bytes * (size_t)n,
                                     cudaMemcpyDeviceToHost, c->copy_stream));
        RTCE_CUDA(c, cudaEventRecord(c->ev_copied[phase], c->copy_stream));
        c->half_used[phase] = true;
        c->host_copies_pending = true;
    }
    return finish_host_call(c, wait);
}

int rtce_render_rgba8_to_host(rtce_ctx* c, const rtce_push_constants* pcs, int frames, unsigned char* host_rgba8) {
    return render_rgba8_impl(c, pcs, frames, host_rgba8, true);
}

int rtce_render_rgba8_to_host_async(rtce_ctx* c, const rtce_push_constants* pcs, int frames, unsigned char* host_rgba8) {
    return render_rgba8_impl(c, pcs, frames, host_rgba8, false);
}

static int fetch_counters(rtce_ctx* c, unsigned long long* h /* [kMaxLevels][4] */) {
    memset(h, 0, sizeof(unsigned long long) * 4 * kMaxLevels);
    int r = flush_lanes(c);
    if (r) return r;
    if (c->count) {
        RTCE_CUDA(c, cudaMemcpyAsync(h, c->d_counters, sizeof(unsigned long long) * 4 * kMaxLevels, cudaMemcpyDeviceToHost, c->stream));
    }
    RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
    return RTCE_OK;
}

int rtce_get_counters(rtce_ctx* c, rtce_counters* out) {
    if (!c || !out) return RTCE_ERR_INVALID;
    DeviceGuard g(c->device);
    unsigned long long h[4 * kMaxLevels];
    int r = fetch_counters(c, h);
    if (r) return r;
    out->rays = out->sdf_evals = out->march_steps = 0;
    for (int i = 0; i < kMaxLevels; i++) {
        out->rays += h[4 * i + 0];
        out->sdf_evals += h[4 * i + 1];
        out->march_steps += h[4 * i + 2];
    }
    out->kernel_launches = c->launches;
    return RTCE_OK;
}

int rtce_get_level_counters(rtce_ctx* c, int level, rtce_counters* out) {
    if (!c || !out) return RTCE_ERR_INVALID;
    if (!c->have_view) return fail(c, RTCE_ERR_STATE, "rtce_get_level_counters: view not set");
    if (level < 0) level += c->levels;
    if (level < 0 || level >= c->levels) return fail(c, RTCE_ERR_INVALID, "rtce_get_level_counters: level");
    DeviceGuard g(c->device);
    unsigned long long h[4 * kMaxLevels];
    int r = fetch_counters(c, h);
    if (r) return r;
    out->rays = h[4 * level + 0];
    out->sdf_evals = h[4 * level + 1];
    out->march_steps = h[4 * level + 2];
    out->kernel_launches = 0;
    return RTCE_OK;
}

int rtce_get_far_miss_rays(rtce_ctx* c, int level, uint64_t* out) {
    if (!c || !out) return RTCE_ERR_INVALID;
    if (!c->have_view) return fail(c, RTCE_ERR_STATE, "rtce_get_far_miss_rays: view not set");
    if (level >= c->levels) return fail(c, RTCE_ERR_INVALID, "rtce_get_far_miss_rays: level");
    DeviceGuard g(c->device);
    unsigned long long h[4 * kMaxLevels];
    int r = fetch_counters(c, h);
    if (r) return r;
    *out = 0;
    if (level >= 0) *out = h[4 * level + 3];
    else for (int i = 0; i < c->levels; i++) *out += h[4 * i + 3];   // negative: every level
    return RTCE_OK;
}

// Diagnostics (not part of the reference boundary): record, for the strip-kernel launch of `level`
// (negative = counted from the last level), four words per warp: globaltimer ns at start, when the
// task tickets ran out, at exit, and the number of tasks taken.  max_warps = 0 switches it off.
#define RTCE_DBG_MAX_WARPS 16384
int rtce_debug_strip_timeline(rtce_ctx* c, int level, uint64_t* out, int max_warps) {
    if (!c) return RTCE_ERR_INVALID;
    if (!c->have_view) return fail(c, RTCE_ERR_STATE, "rtce_debug_strip_timeline: view not set");
    DeviceGuard g(c->device);
    if (max_warps <= 0) { c->dbg_level = -1; return RTCE_OK; }
    if (max_warps > RTCE_DBG_MAX_WARPS) max_warps = RTCE_DBG_MAX_WARPS;
    if (!c->d_dbg) {
        cudaError_t e = cudaMalloc(&c->d_dbg, sizeof(uint64_t) * 4 * RTCE_DBG_MAX_WARPS);
        if (e != cudaSuccess) return fail(c, RTCE_ERR_NOMEM, "cudaMalloc(dbg)", e);
        RTCE_CUDA(c, cudaMemset(c->d_dbg, 0, sizeof(uint64_t) * 4 * RTCE_DBG_MAX_WARPS));
    }
    RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
    if (out && c->dbg_level >= 0)
        RTCE_CUDA(c, cudaMemcpy(out, c->d_dbg, sizeof(uint64_t) * 4 * (size_t)max_warps, cudaMemcpyDeviceToHost));
    if (level < 0) level += c->levels;
    if (level < 0 || level >= c->levels) return fail(c, RTCE_ERR_INVALID, "rtce_debug_strip_timeline: level");
    c->dbg_level = level;
    return RTCE_OK;
}

// Diagnostics: arm (out == NULL) or fetch the launch timeline of the calls since arming: 6 doubles per
// launch {start ms, end ms (relative to the first launch), kind (1 strip, 2 overflow, 10 fused), level,
// group, 0}.  Returns the number of launches written, or a negative status.  Synchronises.
int rtce_debug_launch_timeline(rtce_ctx* c, double* out, int max_launches) {
    if (!c) return RTCE_ERR_INVALID;
    DeviceGuard g(c->device);
    RTCE_CUDA(c, cudaStreamSynchronize(c->stream));
    for (int gi = 0; gi < kMaxGroups; gi++) RTCE_CUDA(c, cudaStreamSynchronize(c->lanes[0].grp_stream[gi]));
    int n = 0;
    if (out && !c->marks.empty()) {
        for (const auto& m : c->marks) {
            if (n >= max_launches) break;
            float a = 0.0f, b = 0.0f;
            cudaEventElapsedTime(&a, c->marks[0].e0, m.e0);
            cudaEventElapsedTime(&b, c->marks[0].e0, m.e1);
            double* o = out + 6 * (size_t)n++;
            o[0] = a; o[1] = b; o[2] = m.kind; o[3] = m.level; o[4] = m.group; o[5] = 0.0;
        }
    }
    for (auto& m : c->marks) {
        cudaEventDestroy(m.e0);
        cudaEventDestroy(m.e1);
    }
    c->marks.clear();
    c->mark_launches = out == nullptr || max_launches > 0;
    if (out && max_launches <= 0) c->mark_launches = false;
    return n;
}

}  // extern "C"
