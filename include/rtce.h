/* rtce.h — C ABI of the B200-native cone-march compute pass.
 *
 * Drop-in boundary for ONE path of IvoteSligte/raytracing_engine: the per-pixel
 * compute pass (shaders/compute.glsl + shaders/utilities.glsl) and the host loop
 * that drives it (src/main.rs:291-316).  The reference has no FFI of its own;
 * the pass sits behind a Vulkan bind-and-dispatch sequence, so every entry
 * point below names the Vulkan-side code it replaces (file:line under the
 * reference tree).  Byte layouts are the reference's own (std140 / push-constant
 * layouts generated from the shaders, src/main.rs:47-66).
 *
 * Conventions: every function returns RTCE_OK (0) or a negative rtce_status and
 * never unwinds across the ABI (the reference's convention is unwrap()-panic —
 * src/main.rs:91,166,315 — a Rust wrapper maps codes to Result).  Host buffers
 * are owned by the caller and copied during the call; device memory is owned by
 * the context.  A context is bound to one CUDA device and one stream, is NOT
 * thread-safe (the reference is one host thread and one queue, main.rs:460,721),
 * and work is asynchronous until rtce_sync / rtce_readback*.  There is no CPU
 * fallback: without a CUDA device rtce_create fails with RTCE_ERR_CUDA.
 */
#ifndef RTCE_H
#define RTCE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTCE_ABI_VERSION 2

/* limits of the reference shader (utilities.glsl:1-3, compute.glsl:14-15, main.rs:359) */
#define RTCE_REF_MAX_OBJECTS 8
#define RTCE_REF_MAX_LEVELS 9
/* limits of this build (SURVEY.md finding 4: both caps are lifted by parameterising the same algorithm) */
#define RTCE_MAX_LEVELS 16
#define RTCE_MAX_SPHERES 16384

typedef enum rtce_status {
    RTCE_OK = 0,
    RTCE_ERR_INVALID = -1,   /* bad argument (null pointer, size, range) */
    RTCE_ERR_CUDA = -2,      /* CUDA runtime error; see rtce_last_error */
    RTCE_ERR_STATE = -3,     /* call order: scene / view not set yet */
    RTCE_ERR_LIMIT = -4,     /* exceeds RTCE_MAX_* */
    RTCE_ERR_NOMEM = -5
} rtce_status;

/* compute.glsl:7-12 `PushConstantData` — 40 bytes, filled at main.rs:301-307, 624-629, 770-773.
 * rot = glam Quat::to_array() = [x, y, z, w].  `iter` and `imageSize` are per-level values the
 * host loop writes (main.rs:301-307); rtce_trace* ignore the caller's and fill them per level. */
typedef struct rtce_push_constants {
    float rot[4];        /* @0  */
    float pos[3];        /* @16 */
    uint32_t iter;       /* @28 */
    float imageSize[2];  /* @32 */
} rtce_push_constants;

/* compute.glsl:26-29 `ConstantBuffer` — 16 bytes, built at main.rs:608-611 / 800-803:
 * view = window size in pixels as f32, ratio = [FOV, FOV * H / W]. */
typedef struct rtce_constant_buffer {
    float view[2];   /* @0 */
    float ratio[2];  /* @8 */
} rtce_constant_buffer;

/* utilities.glsl:16-19 `Object` — one std140 array element of MutableData.objs (16 bytes). */
typedef struct rtce_sphere {
    float pos[3];
    float size;
} rtce_sphere;

/* compute.glsl:17-24 `MutableData` (std140, 656 bytes), built at main.rs:524-605.
 * matCount@0 objCount@4 lightCount@8 | mats[8]@16 stride 32 | objs[8]@272 stride 16 |
 * lights[8]@400 stride 32.  The compute pass reads objCount and objs only. */
#define RTCE_MUTABLE_DATA_SIZE 656
#define RTCE_MUTABLE_DATA_OBJCOUNT_OFFSET 4
#define RTCE_MUTABLE_DATA_OBJS_OFFSET 272

typedef struct rtce_ctx rtce_ctx;

/* Work counters of the last rtce_trace* call: R = rays (pixel-invocations over
 * all levels), E = sphereSDF evaluations, S = march-loop iterations; kernel_launches = CUDA
 * kernels this library launched.  Only filled when rtce_set_option(RTCE_OPT_COUNTERS, 1). */
typedef struct rtce_counters {
    uint64_t rays;
    uint64_t sdf_evals;
    uint64_t march_steps;
    uint64_t kernel_launches;
} rtce_counters;

typedef enum rtce_option {
    RTCE_OPT_COUNTERS = 1,      /* 0/1: count R/E/S on the device (slower kernels, one lane, no graph replay; off by default) */
    RTCE_OPT_FRAME_SLOTS = 2,   /* frames traced concurrently by rtce_trace_batch (pyramid replicas per window), default 16 */
    RTCE_OPT_KERNEL = 3,        /* 0 = auto (= 3), 1 = per-pixel kernel per level (the second opinion of the parity tests),
                                   3 = strip kernel per large level + one latency-oriented per-pixel launch per coarse level */
    RTCE_OPT_STRIP_MIN_RAYS = 4,/* kernel plan 3: a level is traced by the strip kernel when it has at least this
                                   many rays per launch (level pixels x frames of the launch), default 400000 */
    RTCE_OPT_GROUPS = 5,        /* kernel plan 3: number of frame groups (1..8) a batch is split into; every group runs its
                                   launch chain on its own stream so that latency-bound launches of one group overlap
                                   issue-bound launches of the others.  0 (default) = auto: 2 in strict stream order,
                                   1 when frames are in flight on several lanes (those already overlap) */
    RTCE_OPT_STRIP_ROWS = 7,    /* kernel plan 3: rows of a strip-kernel task (a task is 32 columns x this many rows), 1..32 */
    RTCE_OPT_GRAPH = 8,         /* kernel plan 3: 1 (default) = the launch chain of a frame batch (every level of every group,
                                   forks and joins included) is captured once into a CUDA graph and replayed by later
                                   calls with the same batch shape; the per-call cameras travel through device memory.
                                   0 = issue every launch.  Counters and the debug timelines always issue every launch. */
    RTCE_OPT_FRAMES_IN_FLIGHT = 9, /* 1..8, default 8 (measured on BASELINE.json configs[1]: 4 -> 196, 6 -> 214,
                                   8 -> 220 Grays/s).  The reference keeps frames in flight behind its swapchain
                                   (main.rs:859-884: acquire, submit, fence of the PREVIOUS frame).  With W > 1 successive
                                   rtce_trace_batch calls rotate over W windows of frame slots, each with its own streams:
                                   a call starts after everything enqueued on the context's stream so far, but NOT after
                                   the W-1 calls before it, so the latency-bound levels of one batch run under the
                                   issue-bound final level of another.  The context's stream is only ordered after the
                                   batch whose window the next call reuses; rtce_flush (stream-level) or rtce_sync /
                                   rtce_readback* / rtce_shade_batch (which flush) order it after everything: a caller that
                                   reads device pointers of rtce_level_info with its OWN kernels calls rtce_flush first.
                                   rtce_level_info and rtce_readback address the window of the latest call.
                                   W = 1 is strict stream order. */
    RTCE_OPT_STRIP_CTAS_PER_SM = 10, /* kernel plan 3: persistent CTAs per SM of a strip launch; 0 (default) = auto: every
                                   resident slot in strict stream order, 2 with two or three lanes in flight, 1 with more
                                   (leaves room for the launches of the other lanes). */
    RTCE_OPT_FAR_MISS = 11,     /* kernel plan 3: 1 (default) = a ray whose parent texel lies beyond every sphere by more than
                                   RENDER_DIST (host-side conservative bound per camera) is resolved as start + RENDER_DIST
                                   without evaluating an SDF: exactly what compute.glsl:37-66 computes for it (all
                                   distances[i] >= RENDER_DIST -> one trip, no hit).  0 = evaluate every ray in full. */
    RTCE_OPT_LARGE_KERNEL = 12, /* objCount > 8 (extension of utilities.glsl:3): 0 (default) = auto = the register kernel;
                                   2 = register kernel (per-ray cache of lower bounds as an unrolled register array, every
                                   bound touched on every trip);
                                   3 = event-driven kernel (a trip touches only the objects that can be due or be the
                                   minimum; exact, but slower on BASELINE.json configs[3] whose long rays dominate: DESIGN.md
                                   "Large scenes"; the second opinion of the parity tests).  Both hold RTCE_MAX_SPHERES. */
    RTCE_OPT_MARCHER = 13       /* 3 (default) = tracing algorithm 3, the one compute.glsl:44-66 ships; 1, 2 = the author's
                                   alternatives (shaders/tracing_algorithms.txt:2-13 and :16-37) dropped into the same
                                   traceCone frame.  The reference holds only pseudo-code for 1 and 2: their checker is a
                                   restatement, parity UNPINNED by the reference.  objCount 1..8, per-pixel kernels. */
} rtce_option;

/* -- lifetime ------------------------------------------------------------------------------
 * Replaces device/queue selection and compute-pipeline creation with its specialization
 * constant RENDER_DIST (main.rs:68-92, 154-167, 450-460, 631-637; compute.glsl:32). */
int rtce_create(int device, float render_dist, rtce_ctx** out);
void rtce_destroy(rtce_ctx* ctx);
int rtce_abi_version(void);
const char* rtce_last_error(const rtce_ctx* ctx);   /* valid until the next call on ctx */
const char* rtce_status_string(int status);
int rtce_set_option(rtce_ctx* ctx, int option, int64_t value);
/* Run on a caller-owned CUDA stream (cudaStream_t as void*; NULL = the context's own stream). */
int rtce_set_stream(rtce_ctx* ctx, void* cuda_stream);

/* -- scene (binding 1) ------------------------------------------------------------------------
 * Replaces DeviceLocalBuffer::from_data(mutable_data) (main.rs:592-605).  Takes the exact
 * 656-byte std140 image of `MutableData`. */
int rtce_set_scene(rtce_ctx* ctx, const void* mutable_data_std140, size_t nbytes);
/* Extension for objCount > 8 (BASELINE.json configs[3]): n spheres as {x,y,z,size}. */
int rtce_set_spheres(rtce_ctx* ctx, const rtce_sphere* spheres, uint32_t n);

/* -- view (binding 2 + binding 0) -----------------------------------------------------------
 * Replaces the ConstantBuffer upload (main.rs:608-615, 800-834), the level count
 * (main.rs:639; levels = 0 -> floor rule, levels = -1 -> the resize rule main.rs:845) and
 * get_compute_images (main.rs:203-234): allocates the r32f pyramid. */
int rtce_set_view(rtce_ctx* ctx, const rtce_constant_buffer* cb, int levels);
/* Multi-GPU / windowing extension: restrict work to final-level rows [y0, y1) and columns
 * [x0, x1) plus their ancestor chain on the coarser levels.  All zero = whole image. */
int rtce_set_rect(rtce_ctx* ctx, int x0, int y0, int x1, int y1);
int rtce_level_count(const rtce_ctx* ctx);
/* Level `level` of frame slot `slot`: size and device pointer (row-major f32, pitch = w). */
int rtce_level_info(const rtce_ctx* ctx, int level, int slot, int* w, int* h, float** device_ptr);
/* Host mirrors of the pyramid arithmetic, usable without a context or a GPU. */
int rtce_host_level_count(float view_w, int resize_rule);                               /* main.rs:639 / 845 */
int rtce_host_level_dims(int levels, float view_w, float view_h, int level, int* w, int* h); /* main.rs:209-213 */
int rtce_host_image_size(int levels, int level, float view_w, float view_h, float out[2]);   /* main.rs:303-305 */

/* -- the pass ---------------------------------------------------------------------------------
 * Replaces get_command_buffer's compute half + queue submit (main.rs:291-316, 909-914): for each
 * pyramid level, push constants {iter, imageSize}, dispatch over the level image.  One frame into
 * slot 0, or `frames` frames (frame k into slot k % slots).  Asynchronous. */
int rtce_trace(rtce_ctx* ctx, const rtce_push_constants* pc);
int rtce_trace_batch(rtce_ctx* ctx, const rtce_push_constants* pcs, int frames);
/* Extension ("progressive accumulation"): trace `frames` frames and keep only the
 * running FP32 sum of their final levels, divided by `frames` at the end, in an internal image.
 * Up to four chunks of frames are traced side by side (own streams, own windows of slots); the sum
 * is taken on the context's stream in frame order. */
int rtce_trace_accumulate(rtce_ctx* ctx, const rtce_push_constants* pcs, int frames);
int rtce_accum_info(const rtce_ctx* ctx, int* w, int* h, float** device_ptr);
/* Optional: do now what the first rtce_trace_batch (accumulate = 0) or rtce_trace_accumulate (accumulate = 1) of `frames`
 * frames would do once — allocate the slot windows of every lane and capture + instantiate + upload the launch graphs
 * of every chunk shape on every lane (command-buffer pre-recording; the reference re-records per frame, main.rs:284-289,
 * 886).  Without it the first call of a shape pays for all lanes; no later call allocates or instantiates anything. */
int rtce_prepare(rtce_ctx* ctx, int frames, int accumulate);

/* -- consumers of binding 0 (main.rs:650-655 -> fragment.glsl:15,135) ---------------------------
 * Headless readback; both synchronise.  pitch_bytes = 0 -> tightly packed rows. */
int rtce_readback(rtce_ctx* ctx, int level, int slot, float* host, size_t pitch_bytes);
int rtce_readback_accum(rtce_ctx* ctx, float* host, size_t pitch_bytes);
/* -- the presentation shading pass ---------------------------------------------
 * Replaces the draw of get_command_buffer's render-pass half (main.rs:318-338: fullscreen triangle of
 * vertex.glsl:3-10, one fragment per window pixel running fragment.glsl:123-187): shades the final pyramid
 * level of frame slots 0 .. frames-1 of the latest batch with the SAME push constants the batch was traced
 * with (the reference pushes one PushConstantData to both pipelines, main.rs:296-299, 327-331) into RGBA32F
 * colour slots owned by the context — the headless stand-in for the swapchain image.  Needs the materials and
 * lights of rtce_set_scene.  pow() is pinned (csrc/rtce_shade_math.cuh); fragColor.a, which the shader never
 * writes, is 1.  rtce_readback_color_rgba8 converts like a UNORM attachment (clamp, x255, round to nearest even).
 * Counters (RTCE_OPT_COUNTERS): {pixels, hit pixels, distance() calls, light evaluations}. */
int rtce_shade_batch(rtce_ctx* ctx, const rtce_push_constants* pcs, int frames);
int rtce_color_info(const rtce_ctx* ctx, int slot, int* w, int* h, float** device_rgba32f);
int rtce_readback_color(rtce_ctx* ctx, int slot, float* host_rgba32f);
int rtce_readback_color_rgba8(rtce_ctx* ctx, int slot, unsigned char* host_rgba8);
int rtce_get_shade_counters(rtce_ctx* ctx, uint64_t out[4]);

/* End-to-end convenience used by headless callers: trace `frames` frames from HOST push constants
 * and copy every frame's final level to `host_out` (frames * w * h floats), copies overlapped
 * with the tracing of later frames.  Synchronises before returning. */
int rtce_render_to_host(rtce_ctx* ctx, const rtce_push_constants* pcs, int frames, float* host_out);
/* The same for a caller that streams batches into two (or more) host buffers, the way the reference keeps frames in
 * flight behind its swapchain (main.rs:859-884: submit frame k, wait for the fence of frame k - 1): everything of this
 * call is enqueued, then the host waits for the PREVIOUS rtce_render_to_host_async call only.  When call k returns, the
 * frames of call k - 1 are complete in the buffer that call was given, and the device is working on call k.  rtce_sync
 * completes the last call.  `pcs` is consumed during the call; `host_out` must stay valid until the next call (or
 * rtce_sync) returns, and should be page-locked (rtce_host_alloc) or the copies serialise. */
int rtce_render_to_host_async(rtce_ctx* ctx, const rtce_push_constants* pcs, int frames, float* host_out);
/* The headless replacement of one reference frame loop body (main.rs:886-914: compute half, render-pass half, present):
 * trace `frames` frames, shade them (fragment.glsl) and copy every frame's RGBA8 UNORM image (window size, 4 bytes per
 * pixel) to `host_rgba8` (frames * view_w * view_h * 4 bytes).  Synchronises before returning. */
int rtce_render_rgba8_to_host(rtce_ctx* ctx, const rtce_push_constants* pcs, int frames, unsigned char* host_rgba8);
/* Asynchronous form, same contract as rtce_render_to_host_async (the two share the double buffering and may alternate). */
int rtce_render_rgba8_to_host_async(rtce_ctx* ctx, const rtce_push_constants* pcs, int frames, unsigned char* host_rgba8);
/* Page-locked host buffers for rtce_render_to_host / rtce_readback (pageable memory works, slower). */
int rtce_host_alloc(void** out, size_t nbytes);
int rtce_host_free(void* p);
int rtce_sync(rtce_ctx* ctx);                        /* fence wait, main.rs:882-884 */
/* Make the context's stream wait (on the device, the host does not block) for every batch traced so far:
 * after it, work enqueued on the context's stream sees all images.  A no-op with one frame in flight. */
int rtce_flush(rtce_ctx* ctx);
int rtce_get_counters(rtce_ctx* ctx, rtce_counters* out);
/* The same per pyramid level (negative = from the last).  A fused-pyramid launch books all the levels it
 * traces under its finest one. */
int rtce_get_level_counters(rtce_ctx* ctx, int level, rtce_counters* out);
/* Rays of the last counted call that the far-miss shortcut (RTCE_OPT_FAR_MISS) resolved without an SDF: of pyramid
 * level `level`, or of every level when level < 0.  They are included in rtce_counters with the work the reference
 * algorithm does for them (1 trip, objCount evaluations). */
int rtce_get_far_miss_rays(rtce_ctx* ctx, int level, uint64_t* out);

/* -- the step before the pass: camera -> push constants (main.rs:343-357, 402-414, 770-773) ------------
 * Host arithmetic in binary32, one rounding per operation, in glam 0.21's operation order:
 *   rtce_camera_rotation  Data::rotation (main.rs:402-404) = Quat::from_rotation_z(-yaw) * Quat::from_rotation_x(pitch),
 *                         as Quat::to_array() = [x, y, z, w]; sin / cos of the half angles are libm sinf / cosf (glam calls
 *                         f32::sin_cos), so the bits are those of THIS libm: parity with glam on the author's machine is
 *                         unpinned (nothing to compare with: the Rust host cannot be built, SURVEY.md finding 6).
 *   rtce_camera_pose      the per-frame update main.rs:770-773: pitch clamped to [-pi/2, pi/2], inout->rot = rotation,
 *                         inout->pos += local_delta.x * right + local_delta.y * forward + local_delta.z * up with the axes
 *                         (main.rs:351-357: right +X, forward +Y, up +Z) rotated by Quat::mul_vec3 (Data::position,
 *                         main.rs:406-414).  local_delta = NULL adds nothing.  iter / imageSize are left alone. */
int rtce_camera_rotation(float yaw, float pitch, float rot_xyzw[4]);
int rtce_camera_pose(float yaw, float pitch, const float local_delta[3], rtce_push_constants* inout);

/* -- several GPUs (no reference counterpart: one device, one queue, main.rs:450-460) ----------------------------------
 * One process (or thread) per GPU, one context each.  The pass itself never communicates (a pixel depends only on its
 * ancestor chain and uniforms, compute.glsl:81): every rank restricts itself to a row band with rtce_set_rect, and the
 * ONE exchange is the gather of the accumulated image's bands to rank `dst`: grouped ncclSend / ncclRecv, every band
 * received in place (a row band of a row-major image is contiguous).  NCCL is loaded on first use (dlopen of
 * libnccl.so.2: a single-GPU caller needs no NCCL); the caller distributes the 128-byte unique id by its own means.
 *   rtce_host_band_rows   equal split of `height` rows over `world` ranks in whole 8-row dispatch rows (compute.glsl:5)
 *   rtce_gather_accum     stream-ordered after rtce_trace_accumulate: copies this rank's band out of the accumulator into
 *                         one of two staging buffers and exchanges from there on the context's communication stream, so
 *                         the gather of one step overlaps the tracing of the next.  bands = world x {y0, y1} final-level
 *                         rows per rank (NULL = rtce_host_band_rows).  On `dst` the image lands in one of two gather
 *                         buffers (rtce_gather_info, valid after rtce_gather_wait / rtce_readback_gathered).
 *                         world = 1 (no rtce_comm_init needed): the band is copied, nothing is sent. */
#define RTCE_COMM_ID_BYTES 128
int rtce_comm_unique_id(void* id_out_128);
int rtce_comm_init(rtce_ctx* ctx, const void* id_128, int rank, int world);
int rtce_host_band_rows(int height, int world, int rank, int* y0, int* y1);
int rtce_gather_accum(rtce_ctx* ctx, int dst, const int* bands);
/* The frame-by-frame form: gathers the final-level row bands of frame slots 0 .. frames-1 of the LATEST batch to rank
 * `dst`, every band received in place in `dst`'s own frame slots (whose other rows its kernels never write), so that
 * rtce_level_info / rtce_readback on `dst` see whole frames.  Ordered after the batch; the window is not overwritten
 * by a later batch before the exchange has read it; rtce_flush / rtce_sync / rtce_readback* wait for it. */
int rtce_gather_frames(rtce_ctx* ctx, int dst, const int* bands, int frames);
int rtce_gather_wait(rtce_ctx* ctx);   /* the context's stream waits (device side) for the gathers in flight */
int rtce_gather_info(const rtce_ctx* ctx, int* w, int* h, float** device_ptr);
int rtce_readback_gathered(rtce_ctx* ctx, float* host, size_t pitch_bytes);
/* Device time of the most recent completed exchange (staging copy excluded), ms; synchronises the communication stream. */
int rtce_gather_last_ms(rtce_ctx* ctx, float* ms);

/* -- measurement support (no reference counterpart) ---------------------------------------------
 * Sustained FP32 CUDA-core issue rate of `device`: a register-only kernel of independent,
 * non-contracted FMUL/FADD chains, timed with CUDA events.  Writes lane-operations per second
 * (the denominator of the FP32-issue roofline) and the kernel time in ms. */
int rtce_measure_fp32_issue(int device, int iters, double* lane_ops_per_s, double* ms);
/* Exhaustive check (all 2^32 float bit patterns, on the GPU) that the short sqrt / reciprocal
 * sequences of the hot kernels return the bits of IEEE sqrt.rn / rcp.rn on their stated domains.
 * out4 = {sqrt mismatches, rcp mismatches, sqrt inputs checked, rcp inputs checked}. */
int rtce_selftest_math(int device, uint64_t out4[4]);
/* utilities.glsl:32-34 `repeat(p, r) = mod(p + 0.5 r, r) - 0.5 r` (unused by the reference's shaders, provided as a
 * device function for completeness) evaluated on the GPU, componentwise, for n (p, r) pairs of host floats. */
int rtce_selftest_repeat(int device, const float* p, const float* r, float* out, size_t n);

/* Diagnostics: warp timeline of the strip-kernel launch that traces `level` (negative = from the last
 * level).  First call arms it; a later call copies 4 words per warp {start ns, tickets exhausted ns,
 * exit ns, tasks taken} of the most recent launch into `out` and re-arms.  max_warps = 0 disarms. */
int rtce_debug_strip_timeline(rtce_ctx* ctx, int level, uint64_t* out, int max_warps);
/* Diagnostics: launch timeline.  out == NULL arms it (an event pair around every kernel launch from now
 * on); a later call writes 6 doubles per launch {start ms, end ms, kind (1 strip, 2 overflow, 10 fused),
 * level, frame group, 0} and returns how many; it stays armed while max_launches > 0.  Synchronises. */
int rtce_debug_launch_timeline(rtce_ctx* ctx, double* out, int max_launches);

/* ---- allocation-free variant ----------------------------------------------------------------------------------
 * rtce_trace_noalloc has NO counterpart in the reference: it was added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * It traces `frames` (1..RTCE_NOALLOC_MAX_FRAMES) frames of a scene of at most 8 spheres (host array, n x
 * {x, y, z, size}) through all `levels` pyramid levels (0 = the startup rule of rtce_set_view) with the strip kernel on
 * `cuda_stream` of `device`, asynchronously, and needs no context.  The caller owns inputs and outputs: d_pyramid is
 * DEVICE memory, frame f at d_pyramid + f * frame_stride floats, its levels packed back to back in level order (level i =
 * w_i * h_i floats, row pitch w_i, dims from rtce_host_level_dims; frame_stride >= their sum; frames * frame_stride
 * < 2^32).  The kernel's scratch (task tickets, overflow queue, cameras: about 2.4 MB) is a static device array inside
 * the library: every GPU that loads librtce.so holds it until the process exits; the call itself performs no
 * cudaMalloc / cudaFree / allocating library call.  All calls share that array: what must start at zero is cleared on
 * the stream at the head of every call, and calls that OVERLAP on one GPU race on it (keep them in stream order;
 * different GPUs are independent).  Sizes beyond the limits are rejected with RTCE_ERR_LIMIT.  On failure the
 * message is written to `err` (may be NULL).  Texels are bit-identical to rtce_trace_batch + rtce_readback. */
#define RTCE_NOALLOC_MAX_FRAMES 16
int rtce_trace_noalloc(int device, void* cuda_stream, const float* spheres, int n_spheres, const rtce_constant_buffer* cb,
                       int levels, float render_dist, const rtce_push_constants* pcs, int frames, float* d_pyramid,
                       size_t frame_stride, char* err, size_t err_len);

#ifdef __cplusplus
}
#endif
#endif /* RTCE_H */
