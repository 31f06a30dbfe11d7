// rtce_kernels_strip.cu — ONE pyramid level of a frame batch per launch, objCount <= 8, traced by
// persistent barrier-free warps (DESIGN.md "Strip kernel").  This is the dispatch of
// src/main.rs:314 for the levels that hold almost all rays.
//
// Work decomposition
//   grid = one resident wave of CTAs; warps draw tasks (frame, tile) from one atomic ticket counter.
//   A task is a 32-wide x `rows`-high tile of the level: lane = image column for the whole
//   task, so the column terms of the ray direction live in registers; the row terms of the task's
//   rows are computed once (lane r computes row r) into a per-warp shared-memory table and read back
//   as one broadcast LDS.128 per row.  Stores are full 128-byte rows.
//
// Divergence (measured on the default scene, all levels >= 2: 77 % of rays end in the first march
// trip, 20 % in the second, 3 % form a long tail of up to ~30 trips; hit pixels are spatially
// coherent, the tail is not):
//   trip 1  is straight-line code for all 32 lanes (no SDF evaluation: see first_trip);
//   trip 2  runs in place, under one warp-uniform branch (taken by the ~23 % of warps that cover
//           hit pixels, where nearly all lanes take it);
//   trips 3.. are rare per warp, so those rays are compacted: ballot + prefix-popcount push into a
//           per-warp shared-memory queue that PERSISTS across the warp's tasks, and are marched 32
//           at a time with lane refill (a finished lane pops the next queued ray as soon as
//           kRefillMin lanes are free), so the tail runs at >= 50 % lane occupancy instead of the
//           ~20 % of lock-step marching.  Nothing waits on a parked ray: the level's children are
//           traced by the next launch.
#include "rtce_march.cuh"

namespace rtce {

constexpr int kQCap = 64;          // queue entries per warp: < 32 parked + <= 32 pushed by one row
constexpr int kRefillMin = 8;      // refill when this many lanes are idle
constexpr int kPinWords = 40;       // per-warp scratch for pinning uniforms (>= 2 + 4 * 8, multiple of 4)
#ifndef RTCE_MAX_AGE
#define RTCE_MAX_AGE 3
#endif
#ifndef RTCE_TAIL_BUDGET
#define RTCE_TAIL_BUDGET (1 << 30)
#endif
#ifndef RTCE_TICKET_PROBE
#define RTCE_TICKET_PROBE 3
#endif
constexpr int kTicketProbe = RTCE_TICKET_PROBE;   // counters a warp draws from before it leaves (its own + the next ones)
constexpr int kFarPairs = 8;        // parent rows of a task the far-miss fast path handles (16 rows per task)
constexpr int kMaxAge = RTCE_MAX_AGE;   // tail trips after which a ray is evicted to the overflow queue
constexpr int kTailBudget = RTCE_TAIL_BUDGET;   // tail trips a warp spends on ONE task before it evicts every tail ray it
                                                // holds: a tile whose rays all need ~6 trips would otherwise keep one
                                                // warp busy for 30+ us after the others have run out of tasks
constexpr int kEvictMin = 8;       // evict in groups of at least this many rays (or when the local queue is empty)
constexpr int kLeaveBelow = 16;    // with the queue empty, park the stragglers below this many active lanes
#define RTCE_SQ_WORDS(N) (10 + (N))   // o(3) s(3) len last start offset d[N]
#define RTCE_FULL 0xffffffffu

template <int N>
RTCE_DEV void sq_push(float* q, int idx, const Ray<N>& r, float len, float last, float start, unsigned off) {
    q[0 * kQCap + idx] = r.ox;
    q[1 * kQCap + idx] = r.oy;
    q[2 * kQCap + idx] = r.oz;
    q[3 * kQCap + idx] = r.sx;
    q[4 * kQCap + idx] = r.sy;
    q[5 * kQCap + idx] = r.sz;
    q[6 * kQCap + idx] = len;
    q[7 * kQCap + idx] = last;
    q[8 * kQCap + idx] = start;
    q[9 * kQCap + idx] = __uint_as_float(off);
#pragma unroll
    for (int i = 0; i < N; i++) q[(10 + i) * kQCap + idx] = r.d[i];
}

template <int N>
RTCE_DEV void sq_pop(const float* q, int idx, Ray<N>& r, float& len, float& last, float& start, unsigned& off) {
    r.ox = q[0 * kQCap + idx];
    r.oy = q[1 * kQCap + idx];
    r.oz = q[2 * kQCap + idx];
    r.sx = q[3 * kQCap + idx];
    r.sy = q[4 * kQCap + idx];
    r.sz = q[5 * kQCap + idx];
    len = q[6 * kQCap + idx];
    last = q[7 * kQCap + idx];
    start = q[8 * kQCap + idx];
    off = __float_as_uint(q[9 * kQCap + idx]);
#pragma unroll
    for (int i = 0; i < N; i++) r.d[i] = q[(10 + i) * kQCap + idx];
}

// Overflow queue in global memory (structure of arrays, word w of ray k at oq[w * cap + k]).
struct OverflowQueue {
    float* buf;
    unsigned* count;
    unsigned cap;
};

template <int N>
RTCE_DEV void oq_store(const OverflowQueue& oq, unsigned k, const Ray<N>& r, float len, float last, float start, unsigned off) {
    float* b = oq.buf + k;
    const size_t c = oq.cap;
    b[0 * c] = r.ox; b[1 * c] = r.oy; b[2 * c] = r.oz;
    b[3 * c] = r.sx; b[4 * c] = r.sy; b[5 * c] = r.sz;
    b[6 * c] = len; b[7 * c] = last; b[8 * c] = start;
    b[9 * c] = __uint_as_float(off);
#pragma unroll
    for (int i = 0; i < N; i++) b[(10 + i) * c] = r.d[i];
}

template <int N>
RTCE_DEV void oq_load(const float* buf, unsigned cap, unsigned k, Ray<N>& r, float& len, float& last, float& start, unsigned& off) {
    const float* b = buf + k;
    const size_t c = cap;
    r.ox = __ldcg(b + 0 * c); r.oy = __ldcg(b + 1 * c); r.oz = __ldcg(b + 2 * c);
    r.sx = __ldcg(b + 3 * c); r.sy = __ldcg(b + 4 * c); r.sz = __ldcg(b + 5 * c);
    len = __ldcg(b + 6 * c); last = __ldcg(b + 7 * c); start = __ldcg(b + 8 * c);
    off = __float_as_uint(__ldcg(b + 9 * c));
#pragma unroll
    for (int i = 0; i < N; i++) r.d[i] = __ldcg(b + (10 + i) * c);
}

// March queued rays (trips 3.. of compute.glsl:44-66) with lane refill.  drain = false: called with
// a full warp's worth queued; returns once the queue is empty and fewer than kLeaveBelow lanes are
// still marching, parking those back in the queue.  drain = true: runs until every ray is finished
// or evicted.  A ray that has marched kMaxAge trips in here is a long one (a cone grazing a sphere:
// 10 .. 100+ trips): it is evicted to the global overflow queue, which the whole device marches
// after this kernel (k_march_overflow), instead of serialising in this warp.
template <int N, bool COUNT>
RTCE_DEV int tail_march(float* q, int qn, const bool drain, float* __restrict__ img, const float thr, const float rd,
                        const float4* __restrict__ sph, const OverflowQueue& oq, bool& oq_open, int& budget, unsigned& cE,
                        unsigned& cS, const int lane) {
    const unsigned lt = (1u << lane) - 1u;
    bool active = false;
    Ray<N> r;
    float len = 0.0f, last = 0.0f, start = 0.0f;
    unsigned off = 0;
    int age = 0;
    for (;;) {
        unsigned am = __ballot_sync(RTCE_FULL, active);
        if (oq_open) {
            const bool spent = budget <= 0;
            const bool old = active && (age >= kMaxAge || spent);
            const unsigned om = __ballot_sync(RTCE_FULL, old);
            if (om != 0u && (__popc(om) >= kEvictMin || qn == 0 || spent)) {
                const int n = __popc(om);
                unsigned base = 0;
                if (lane == __ffs(om) - 1) base = atomicAdd(oq.count, (unsigned)n);
                base = __shfl_sync(RTCE_FULL, base, __ffs(om) - 1);
                const unsigned k = base + (unsigned)__popc(om & lt);
                if (old && k < oq.cap) {   // every slot below min(count, cap) is written: the consumer reads exactly those
                    oq_store<N>(oq, k, r, len, last, start, off);
                    active = false;
                }
                if (base + (unsigned)n > oq.cap) oq_open = false;   // full: the rest is marched here
                am = __ballot_sync(RTCE_FULL, active);
            }
        }
        const int nact = __popc(am);
        if (qn > 0 && 32 - nact >= kRefillMin) {
            const int rank = __popc(~am & lt);
            if (!active && rank < qn) {
                sq_pop<N>(q, qn - 1 - rank, r, len, last, start, off);
                active = true;
                age = 0;
            }
            qn -= min(32 - nact, qn);
            __syncwarp();
        } else if (am == 0u) {
            break;
        } else if (!drain && qn == 0 && nact < kLeaveBelow) {
            if (active) sq_push<N>(q, __popc(am & lt), r, len, last, start, off);
            qn = nact;
            __syncwarp();
            break;
        }
        if (active) {
            unsigned e = 0, s = 0;
            const bool go = march_trip<N, COUNT, false>(r, len, last, thr, rd, sph, e, s);
            if (COUNT) { cE += e; cS += s; }
            age++;
            if (!go) {
                img[off] = glsl_max(fadd(start, len), 0.0f);   // compute.glsl:84, :86
                active = false;
            }
        }
        budget--;
    }
    return qn;
}

// The overflow queue of one strip launch, marched to the end by every warp of the device.
#ifndef RTCE_OQ_PER_WARP
#define RTCE_OQ_PER_WARP 32
#endif
#ifndef RTCE_OQ_BRANCHY
#define RTCE_OQ_BRANCHY 0
#endif
#ifndef RTCE_OQ_REFILL
#define RTCE_OQ_REFILL kRefillMin
#endif
template <int N, bool COUNT>
__global__ void __launch_bounds__(128)
    k_march_overflow(const __grid_constant__ OverflowParams P) {
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned n = min(*P.oq_count, P.oq_cap);
    // Warps in use grow with the queue (~48 rays each, refilled as rays finish): a short queue is
    // bound by the dependent chain of its longest ray and runs fastest with one warp per
    // scheduler; a long one is issue-bound and wants them all.
    if ((blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * (unsigned)RTCE_OQ_PER_WARP >= n) return;
    float4 sph[N > 0 ? N : 1];
#pragma unroll
    for (int i = 0; i < N; i++) sph[i] = P.sph[i];
    const float thr = P.threshold, rd = P.render_dist;
    float* const img = P.out;
    unsigned cE = 0, cS = 0;
    bool active = false, exhausted = false;
    Ray<N> r;
    float len = 0.0f, last = 0.0f, start = 0.0f;
    unsigned off = 0;
    for (;;) {
        const unsigned am = __ballot_sync(RTCE_FULL, active);
        const int nfree = 32 - __popc(am);
        if (!exhausted && nfree >= RTCE_OQ_REFILL) {
            unsigned base = 0;
            if (lane == 0) base = atomicAdd(P.oq_head, (unsigned)nfree);
            base = __shfl_sync(RTCE_FULL, base, 0);
            const unsigned k = base + (unsigned)__popc(~am & lt);
            if (!active && k < n) {
                oq_load<N>(P.oq, P.oq_cap, k, r, len, last, start, off);
                active = true;
            }
            if (base + (unsigned)nfree >= n) exhausted = true;
        } else if (am == 0u) {
            if (exhausted) break;
            continue;
        }
        if (active) {
            unsigned e = 0, s = 0;
#if RTCE_OQ_BRANCHY
            const bool go = march_trip<N, COUNT, false>(r, len, last, thr, rd, sph, e, s);
#else
            const bool go = march_block<N, COUNT, RTCE_PRED_BLOCK>(r, len, last, thr, rd, sph, e, s);
#endif
            if (COUNT) { cE += e; cS += s; }
            if (!go) {
                img[off] = glsl_max(fadd(start, len), 0.0f);   // compute.glsl:84, :86
                active = false;
            }
        }
    }
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) {
            cE += __shfl_down_sync(RTCE_FULL, cE, o);
            cS += __shfl_down_sync(RTCE_FULL, cS, o);
        }
        if (lane == 0) {
            atomicAdd(P.counters + 1, (unsigned long long)cE);
            atomicAdd(P.counters + 2, (unsigned long long)cS);
        }
    }
}

RTCE_DEV unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

#ifndef RTCE_STRIP_MIN_BLOCKS
#define RTCE_STRIP_MIN_BLOCKS 3
#endif
// Compile-time switches of the per-ticket path (all on; 0 restores the round-1 form for A/B runs).  ncu of the final
// level (profiles/r02_v9_strip_final_level_hot_blocks.txt): fetching and decoding a ticket was 120 instructions x
// 130 560 tickets = 14.5 % of the launch's instructions and a third of its stall samples, the whole-task far-miss path
// another 7.5 %.
#ifndef RTCE_STRIP_MAGIC
#define RTCE_STRIP_MAGIC 1      // ticket -> (frame, tile row, tile column) by two 64-bit multiplies instead of two divisions
#endif
#ifndef RTCE_STRIP_FARFAST
#define RTCE_STRIP_FARFAST 1    // whole-task far miss of a full-height task: loads and stores at constant row offsets
#endif

// n / d for n * d < 2^40 (tasks < 2^24, divisors < 2^16: every view up to 8K x 32 frames) by one 64-bit multiply with
// m = floor(2^40 / d) + 1 from the host; exact in that range (error term n (m d - 2^40) < 2^40).  m = 0: divide.
RTCE_DEV unsigned div_magic(unsigned n, unsigned long long m) { return (unsigned)(((unsigned long long)n * m) >> 40); }

template <int N, bool COUNT, bool HAS_PARENT>
__global__ void __launch_bounds__(kStripThreads, (N <= 4 ? RTCE_STRIP_MIN_BLOCKS : 2))
    k_trace_strip(const __grid_constant__ StripParams P) {
    extern __shared__ __align__(16) float s_mem[];
    constexpr int kWarpWords = RTCE_SQ_WORDS(N) * kQCap + 4 * kStripMaxRows + kPinWords;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* const q = s_mem + warp * kWarpWords;
    float4* const rowtab = reinterpret_cast<float4*>(q + RTCE_SQ_WORDS(N) * kQCap);

    // Loop-invariant uniforms are PINNED in registers: ptxas re-materialises plain parameter loads
    // inside the row loop (13 LDC/LDCU per row = 10 % of its issue slots; it also sees through a
    // shuffle of a uniform value).  A value read back through a volatile shared-memory load cannot
    // be re-materialised.  Scene and level constants once per warp, the camera once per task.
    volatile float* const pin = reinterpret_cast<volatile float*>(rowtab + kStripMaxRows);   // kPinWords floats
    float4 sph[N > 0 ? N : 1];
    float thr, rd;
    {
        pin[0] = P.threshold;
        pin[1] = P.render_dist;
#pragma unroll
        for (int i = 0; i < N; i++) {
            pin[2 + 4 * i + 0] = P.sph[i].x;
            pin[2 + 4 * i + 1] = P.sph[i].y;
            pin[2 + 4 * i + 2] = P.sph[i].z;
            pin[2 + 4 * i + 3] = P.sph[i].w;
        }
        __syncwarp();
        thr = pin[0];
        rd = pin[1];
#pragma unroll
        for (int i = 0; i < N; i++) sph[i] = make_float4(pin[2 + 4 * i + 0], pin[2 + 4 * i + 1], pin[2 + 4 * i + 2], pin[2 + 4 * i + 3]);
        __syncwarp();
    }
    float* const img = P.out;   // frame slot 0; queue entries carry offsets from here
    const OverflowQueue oq{P.oq, P.oq_count, P.oq_cap};
    bool oq_open = P.oq != nullptr;   // warp-uniform
    const int tiles = P.ntx * P.nty;
    const int tasks = tiles * P.frames;
    const int nwarps = gridDim.x * (kStripThreads / 32);
    const unsigned lt = (1u << lane) - 1u;
    unsigned cR = 0, cE = 0, cS = 0, cF = 0;   // cF: rays resolved by the far-miss shortcut (a subset of cR)
    int qn = 0;   // warp-uniform

    // Tasks (frame-major) are handed out dynamically, one atomic per task fetched one task ahead: march cost differs
    // by > 5x between tiles that cover spheres and tiles that see only background.  With a parent level and the
    // device to itself (P.two_pass) the ticket range is walked TWICE: the first walk takes only the tiles whose
    // parent texels show a hit ("hot": second trips, tail marches), the second walk the rest, whose cost is small and
    // uniform, so the launch ends without a ragged tail of expensive tiles (longest-processing-time-first scheduling
    // with a two-class estimate).  With other lanes' launches running beside it the ragged end is filled by them and
    // the second walk only costs (every ticket fetched, decoded and classified twice): one walk.
    // Measured and rejected (profiles/r02_strip_variants.md): tickets issued 2 tasks ahead, the classifying parent
    // texels fetched one task ahead, 2 or 4 tickets per atomic — each faster on some stall counter, none faster in
    // batches per second with six lanes in flight.
    const bool two_pass = HAS_PARENT && P.two_pass != 0;
    const int tickets_total = two_pass ? 2 * tasks : tasks;
    // Tickets: the first round is static (warp index); the rest come from kTicketCounters striped counters
    // (rtce_kernels.cuh), fetched one task ahead from the warp's current counter.
    const int wglobal = blockIdx.x * (kStripThreads / 32) + warp;
    int* const tcount = P.tickets + 32;
    // every stripe is drained by its home warps, so each counter needs some: a small grid uses one counter
    const int kc = nwarps >= 8 * kTicketCounters ? kTicketCounters : 1;
    const int probe = min(kTicketProbe, kc);
    int cur = wglobal & (kc - 1);   // warp-uniform; moves on when the counter's stripe is exhausted
    int tried = 1;                               // counters this warp has drawn from
    int next = 0;
    unsigned long long t0 = 0, t1 = 0;
    int ndone = 0;
    if (P.dbg) t0 = globaltimer_ns();
    // `continue` below runs the increment: the prefetched ticket, or — once the current counter is past the end of its
    // stripe — a ticket from the next counter; a warp leaves after it has seen kTicketProbe counters exhausted (the
    // stripes interleave the tasks and 222 warps share each, so they run out within a few tasks of each other: probing
    // all 16 would be 15 dependent atomics of pure latency at the end of every warp)
    auto draw = [&](int from) { return nwarps + atomicAdd(tcount + from * kTicketStride, 1) * kc + from; };
    auto advance = [&]() {
        int t = __shfl_sync(RTCE_FULL, next, 0);
        while (t >= tickets_total && tried < probe) {
            tried++;
            cur = (cur + 1) & (kc - 1);
            if (lane == 0) next = draw(cur);
            t = __shfl_sync(RTCE_FULL, next, 0);
        }
        return t;
    };
    for (int ticket = wglobal; ticket < tickets_total; ticket = advance()) {
        if (lane == 0) next = draw(cur);
        const int second = ticket >= tasks ? 1 : 0;
        const unsigned task = (unsigned)(ticket - second * tasks);
        unsigned uf, tile, uty;
        if (RTCE_STRIP_MAGIC && P.magic_tiles != 0ull) {
            uf = div_magic(task, P.magic_tiles);
            tile = task - uf * (unsigned)tiles;
            uty = div_magic(tile, P.magic_ntx);
        } else {
            uf = task / (unsigned)tiles;
            tile = task - uf * (unsigned)tiles;
            uty = tile / (unsigned)P.ntx;
        }
        const int f = (int)uf, ty = (int)uty, tx = (int)(tile - uty * (unsigned)P.ntx);
        const int x = P.x0 + tx * 32 + lane;
        const bool xok = x < P.x1;
        const int xc = min(x, P.x1 - 1);   // lanes past the region repeat its last column and store nothing
        const int yb = P.y0 + ty * P.rows;
        const int rows = min(P.rows, P.y1 - yb);
        // :80-82  rows 2k and 2k+1 share the parent texel; it is fetched one row pair ahead
        const float* pp = nullptr;
        float start = 1.0f, ahead = 0.0f;   // compute.glsl:79
        int pairs_left = 0;                 // parent rows still to fetch after `ahead`
        if (HAS_PARENT) {
            pp = P.parent + ((size_t)f * P.slot_stride + (size_t)(yb >> 1) * P.pw + (xc >> 1));
            pairs_left = ((yb + rows - 1) >> 1) - (yb >> 1);
            // Classification and the whole-task far miss read the SAME parent texels, so they are fetched in ONE round trip:
            // the task's (up to) 8 parent rows of this lane's column and the camera's far-miss distance.  (Fetched as
            // first / last row for the class and then all rows for the far miss, a background task was two dependent L2
            // round trips around ~150 instructions: with 6 warps per scheduler that is 0.6 instructions per cycle, which
            // is what ncu measured for the whole launch.)
            //   hot  = a parent ray of the task stopped on a sphere (second trips, tail marches: the expensive class);
            //   far  = every parent texel is at least the camera's far-miss distance away, so every ray of the task is
            //          start + RENDER_DIST (see the row loop): 71 % of the final level's rays.
            const bool farshape = !(yb & 1) && pairs_left < kFarPairs;   // warp-uniform: the task's rows are whole pairs, <= 8 of them
            int hot;
            if (farshape) {
                const float tf = __ldg(reinterpret_cast<const float*>(P.cams + f) + 7);
                const bool full = RTCE_STRIP_FARFAST && rows == 2 * kFarPairs;   // a full-height task (warp-uniform)
                float pv[kFarPairs];
                if (full) {
#pragma unroll
                    for (int k = 0; k < kFarPairs; k++) pv[k] = __ldg(pp + (size_t)(k * P.pw));
                } else {
#pragma unroll
                    for (int k = 0; k < kFarPairs; k++) pv[k] = k <= pairs_left ? __ldg(pp + (size_t)k * P.pw) : pv[0];
                }
                start = pv[0];
                float nearest = pv[0];
#pragma unroll
                for (int k = 1; k < kFarPairs; k++) nearest = fminf(nearest, pv[k]);
                hot = __any_sync(RTCE_FULL, nearest < rd) ? 1 : 0;
                if (two_pass && hot == second) {   // not this walk's class
                    continue;
                }
                if (!hot && __all_sync(RTCE_FULL, !xok || nearest >= tf)) {
                    if (xok) {
                        float* o = img + ((size_t)f * P.slot_stride + (size_t)yb * P.w + xc);
                        if (full) {
#pragma unroll
                            for (int rr = 0; rr < 2 * kFarPairs; rr++) o[(size_t)(rr * P.w)] = fadd(pv[rr >> 1], rd);   // :84, :86 (positive)
                        } else {
#pragma unroll
                            for (int rr = 0; rr < 2 * kFarPairs; rr++)
                                if (rr < rows) o[(size_t)rr * P.w] = fadd(pv[rr >> 1], rd);
                        }
                        if (COUNT) { cR += (unsigned)rows; cE += (unsigned)(N * rows); cS += (unsigned)rows; cF += (unsigned)rows; }
                    }
                    ndone++;
                    continue;
                }
            } else {   // odd first row or a taller task (RTCE_OPT_STRIP_ROWS > 16): class from the first and last parent row
                start = __ldg(pp);
                const float bottom = __ldg(pp + (size_t)pairs_left * P.pw);
                hot = __any_sync(RTCE_FULL, start < rd || bottom < rd) ? 1 : 0;
                if (two_pass && hot == second) {
                    continue;
                }
            }
            if (pairs_left > 0) { pp += P.pw; ahead = __ldg(pp); pairs_left--; }
        }
        ndone++;
        int budget = kTailBudget;   // of this task
        Cam cam;
        {
            if (lane < 8) pin[lane] = __ldg(reinterpret_cast<const float*>(P.cams + f) + lane);
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 4; i++) cam.rot[i] = pin[i];
#pragma unroll
            for (int i = 0; i < 3; i++) cam.pos[i] = pin[4 + i];
            cam.pad = pin[7];   // far-miss distance of this camera (host: far_miss_distance)
        }
        const float tfar = cam.pad;
        const float4 col = col_terms(cam, norm_coord(xc, P.isx, P.ratx));
        __syncwarp();   // the previous task's table reads are done
        if (lane < rows) rowtab[lane] = row_terms(cam, norm_coord(yb + lane, P.isy, P.raty));
        __syncwarp();
        float* op = img + ((size_t)f * P.slot_stride + (size_t)yb * P.w + xc);   // this lane's texel, row by row
        for (int rr = 0; rr < rows; rr++, op += P.w) {
            if (HAS_PARENT && __all_sync(RTCE_FULL, !xok || start >= tfar)) {
                // every ray of the row starts beyond every sphere by more than RENDER_DIST: all distances[i] >=
                // RENDER_DIST, so the first trip takes dist = RENDER_DIST, is no hit and ends the march (:44-66)
                if (xok) *op = fadd(start, rd);   // :84, :86 (positive)
                if (COUNT && xok) { cR++; cE += (unsigned)N; cS++; cF++; }
                if ((yb + rr) & 1) {   // next row starts a new pair
                    start = ahead;
                    if (pairs_left > 0) { pp += P.pw; ahead = __ldg(pp); pairs_left--; }
                }
                continue;
            }
            const float4 row = rowtab[rr];
            Ray<N> ray;
            ray_setup<N>(ray, col, row, start, cam, sph);
            float len;
            unsigned e = 0, s = 0;
            bool cont = first_trip<N, COUNT>(ray, len, thr, rd, e, s);
            if (COUNT && xok) { cR++; cE += e; cS += s; }
            if (xok && !cont) *op = glsl_max(fadd(start, len), 0.0f);   // :84, :86
            cont = cont && xok;
            if (__any_sync(RTCE_FULL, cont)) {   // some lane needs a second trip
                bool cont2 = false;
                float last = len;   // len = 0 + last after the first trip
                if (cont) {
                    e = 0, s = 0;
                    cont2 = march_trip<N, COUNT, false>(ray, len, last, thr, rd, sph, e, s);
                    if (COUNT) { cE += e; cS += s; }
                    if (!cont2) *op = glsl_max(fadd(start, len), 0.0f);
                }
                const unsigned m2 = __ballot_sync(RTCE_FULL, cont2);
                if (m2) {   // park the rays that need a third trip
                    if (cont2) sq_push<N>(q, qn + __popc(m2 & lt), ray, len, last, start, (unsigned)(op - img));
                    qn += __popc(m2);
                    __syncwarp();
                    if (qn >= 32) qn = tail_march<N, COUNT>(q, qn, false, img, thr, rd, sph, oq, oq_open, budget, cE, cS, lane);
                }
            }
            if (HAS_PARENT) {
                if ((yb + rr) & 1) {   // next row starts a new pair
                    start = ahead;
                    if (pairs_left > 0) { pp += P.pw; ahead = __ldg(pp); pairs_left--; }
                }
            }
        }
    }
    if (P.dbg) t1 = globaltimer_ns();
    int budget = kTailBudget;
    if (qn > 0) qn = tail_march<N, COUNT>(q, qn, true, img, thr, rd, sph, oq, oq_open, budget, cE, cS, lane);
    if (P.dbg && lane == 0) {
        unsigned long long* d = P.dbg + 4 * (size_t)(blockIdx.x * (kStripThreads / 32) + warp);
        d[0] = t0; d[1] = t1; d[2] = globaltimer_ns(); d[3] = (unsigned long long)ndone;
    }

    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) {
            cR += __shfl_down_sync(RTCE_FULL, cR, o);
            cE += __shfl_down_sync(RTCE_FULL, cE, o);
            cS += __shfl_down_sync(RTCE_FULL, cS, o);
            cF += __shfl_down_sync(RTCE_FULL, cF, o);
        }
        if (lane == 0) {
            atomicAdd(P.counters + 0, (unsigned long long)cR);
            atomicAdd(P.counters + 1, (unsigned long long)cE);
            atomicAdd(P.counters + 2, (unsigned long long)cS);
            if (cF) atomicAdd(P.counters + 3, (unsigned long long)cF);
        }
    }
}

template <int N, bool COUNT, bool HAS_PARENT>
static cudaError_t launch_strip_ncp(const StripParams& p, cudaStream_t st) {
    const int smem = (kStripThreads / 32) * (RTCE_SQ_WORDS(N) * kQCap + 4 * kStripMaxRows + kPinWords) * (int)sizeof(float);
    // per instantiation AND per device: function attributes and occupancy belong to the device's copy of the kernel
    static int resident_of[kMaxDevices] = {};   // CTAs the device holds at once; 0 = not configured yet
    const int dslot = current_device_slot();
    const int n_sms = device_sm_count();
    if (!resident_of[dslot]) {
        cudaError_t e = cudaFuncSetAttribute(k_trace_strip<N, COUNT, HAS_PARENT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(k_trace_strip<N, COUNT, HAS_PARENT>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        if (e != cudaSuccess) return e;
        int per_sm = 3;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_trace_strip<N, COUNT, HAS_PARENT>, kStripThreads, smem);
        if (e != cudaSuccess) return e;
        resident_of[dslot] = n_sms * (per_sm > 0 ? per_sm : 1);
    }
    const int resident = resident_of[dslot];
    // persistent CTAs: one resident wave for the whole batch
    const long long tasks = (long long)p.ntx * p.nty * p.frames;
    long long ctas = resident / (p.cta_share > 1 ? p.cta_share : 1);
    if (p.ctas_per_sm > 0 && ctas > (long long)n_sms * p.ctas_per_sm) ctas = (long long)n_sms * p.ctas_per_sm;
    if (ctas < 1) ctas = 1;
    const long long need = (tasks + kStripThreads / 32 - 1) / (kStripThreads / 32);
    if (ctas > need) ctas = need;
    dim3 grid((unsigned)ctas, 1, 1);
    k_trace_strip<N, COUNT, HAS_PARENT><<<grid, kStripThreads, smem, st>>>(p);
    return cudaGetLastError();
}

template <int N>
static cudaError_t launch_strip_n(const StripParams& p, bool count, cudaStream_t st) {
    if (p.parent) return count ? launch_strip_ncp<N, true, true>(p, st) : launch_strip_ncp<N, false, true>(p, st);
    return count ? launch_strip_ncp<N, true, false>(p, st) : launch_strip_ncp<N, false, false>(p, st);
}

cudaError_t launch_trace_strip(const StripParams& p, bool count, cudaStream_t st) {
    if (p.frames <= 0 || p.ntx <= 0 || p.nty <= 0) return cudaSuccess;
    if (p.rows < 1 || p.rows > kStripMaxRows || !(p.render_dist > 0.0f)) return cudaErrorInvalidValue;
#ifdef RTCE_EXPERIMENT   // compile-time experiments: one instantiation only
    return launch_strip_ncp<4, false, true>(p, st);
#else
    switch (p.n_spheres) {
        case 0: return launch_strip_n<0>(p, count, st);
        case 1: return launch_strip_n<1>(p, count, st);
        case 2: return launch_strip_n<2>(p, count, st);
        case 3: return launch_strip_n<3>(p, count, st);
        case 4: return launch_strip_n<4>(p, count, st);
        case 5: return launch_strip_n<5>(p, count, st);
        case 6: return launch_strip_n<6>(p, count, st);
        case 7: return launch_strip_n<7>(p, count, st);
        case 8: return launch_strip_n<8>(p, count, st);
        default: return cudaErrorInvalidValue;
    }
#endif
}


template <int N>
static cudaError_t launch_overflow_n(const OverflowParams& p, bool count, cudaStream_t st) {
    const int ctas = device_sm_count() * 2;   // up to two warps per scheduler (see the kernel); a 128-thread CTA fits
                                              // beside three resident strip CTAs
    if (count)
        k_march_overflow<N, true><<<ctas, 128, 0, st>>>(p);
    else
        k_march_overflow<N, false><<<ctas, 128, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_march_overflow(const OverflowParams& p, bool count, cudaStream_t st) {
#ifdef RTCE_EXPERIMENT
    return launch_overflow_n<4>(p, count, st);
#else
    switch (p.n_spheres) {
        case 0: return cudaSuccess;   // no objects: every ray ends in its first trip
        case 1: return launch_overflow_n<1>(p, count, st);
        case 2: return launch_overflow_n<2>(p, count, st);
        case 3: return launch_overflow_n<3>(p, count, st);
        case 4: return launch_overflow_n<4>(p, count, st);
        case 5: return launch_overflow_n<5>(p, count, st);
        case 6: return launch_overflow_n<6>(p, count, st);
        case 7: return launch_overflow_n<7>(p, count, st);
        case 8: return launch_overflow_n<8>(p, count, st);
        default: return cudaErrorInvalidValue;
    }
#endif
}

}  // namespace rtce
