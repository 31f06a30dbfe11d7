//! rtce.rs — Rust binding of the C ABI in include/rtce.h (librtce.so), for the reference's host
//! (IvoteSligte/raytracing_engine, src/main.rs).  SOURCE ONLY: this image has no rustc/cargo and
//! the reference itself does not build outside its author's machine (Cargo.toml:9 is a local
//! Windows path dependency), so the ABI is exercised from Python/ctypes and C++ in tests/.
//!
//! The #[repr(C)] structs are byte-for-byte the types vulkano-shaders generates from the GLSL
//! (`shaders::ty::*`, src/main.rs:47-66), so the values src/main.rs already builds at
//! :524-629 can be passed through unchanged.
#![allow(non_camel_case_types, non_snake_case, dead_code)]

use std::ffi::CStr;
use std::os::raw::{c_char, c_int, c_void};

/// compute.glsl:7-12 PushConstantData, 40 bytes (src/main.rs:301-307, 624-629, 770-773).
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct PushConstantData {
    pub rot: [f32; 4],       // glam Quat::to_array() = [x, y, z, w]
    pub pos: [f32; 3],
    pub iter: u32,           // filled per level by the library (main.rs:302)
    pub imageSize: [f32; 2], // filled per level by the library (main.rs:303-305)
}

/// compute.glsl:26-29 ConstantBuffer, 16 bytes (src/main.rs:608-611).
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct ConstantBuffer {
    pub view: [f32; 2],
    pub ratio: [f32; 2],
}

/// utilities.glsl:16-19 Object = one element of MutableData.objs.
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct Sphere {
    pub pos: [f32; 3],
    pub size: f32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct Counters {
    pub rays: u64,
    pub sdf_evals: u64,
    pub march_steps: u64,
    pub kernel_launches: u64,
}

#[repr(C)]
pub struct rtce_ctx {
    _private: [u8; 0],
}

pub const RTCE_MUTABLE_DATA_SIZE: usize = 656;

#[link(name = "rtce")]
extern "C" {
    fn rtce_abi_version() -> c_int;
    fn rtce_create(device: c_int, render_dist: f32, out: *mut *mut rtce_ctx) -> c_int;
    fn rtce_destroy(ctx: *mut rtce_ctx);
    fn rtce_last_error(ctx: *const rtce_ctx) -> *const c_char;
    fn rtce_status_string(status: c_int) -> *const c_char;
    fn rtce_set_option(ctx: *mut rtce_ctx, option: c_int, value: i64) -> c_int;
    fn rtce_set_scene(ctx: *mut rtce_ctx, mutable_data_std140: *const c_void, nbytes: usize) -> c_int;
    fn rtce_set_spheres(ctx: *mut rtce_ctx, spheres: *const Sphere, n: u32) -> c_int;
    fn rtce_set_view(ctx: *mut rtce_ctx, cb: *const ConstantBuffer, levels: c_int) -> c_int;
    fn rtce_set_rect(ctx: *mut rtce_ctx, x0: c_int, y0: c_int, x1: c_int, y1: c_int) -> c_int;
    fn rtce_level_count(ctx: *const rtce_ctx) -> c_int;
    fn rtce_level_info(ctx: *const rtce_ctx, level: c_int, slot: c_int, w: *mut c_int, h: *mut c_int,
                       device_ptr: *mut *mut f32) -> c_int;
    fn rtce_trace(ctx: *mut rtce_ctx, pc: *const PushConstantData) -> c_int;
    fn rtce_trace_batch(ctx: *mut rtce_ctx, pcs: *const PushConstantData, frames: c_int) -> c_int;
    fn rtce_readback(ctx: *mut rtce_ctx, level: c_int, slot: c_int, host: *mut f32, pitch_bytes: usize) -> c_int;
    fn rtce_render_to_host(ctx: *mut rtce_ctx, pcs: *const PushConstantData, frames: c_int, host_out: *mut f32) -> c_int;
    fn rtce_render_to_host_async(ctx: *mut rtce_ctx, pcs: *const PushConstantData, frames: c_int, host_out: *mut f32) -> c_int;
    fn rtce_sync(ctx: *mut rtce_ctx) -> c_int;
    fn rtce_flush(ctx: *mut rtce_ctx) -> c_int;
    fn rtce_get_counters(ctx: *mut rtce_ctx, out: *mut Counters) -> c_int;
    fn rtce_set_stream(ctx: *mut rtce_ctx, cuda_stream: *mut c_void) -> c_int;
    fn rtce_trace_accumulate(ctx: *mut rtce_ctx, pcs: *const PushConstantData, frames: c_int) -> c_int;
    fn rtce_readback_accum(ctx: *mut rtce_ctx, host: *mut f32, pitch_bytes: usize) -> c_int;
    fn rtce_shade_batch(ctx: *mut rtce_ctx, pcs: *const PushConstantData, frames: c_int) -> c_int;
    fn rtce_color_info(ctx: *const rtce_ctx, slot: c_int, w: *mut c_int, h: *mut c_int, device_rgba32f: *mut *mut f32) -> c_int;
    fn rtce_readback_color(ctx: *mut rtce_ctx, slot: c_int, host_rgba32f: *mut f32) -> c_int;
    fn rtce_readback_color_rgba8(ctx: *mut rtce_ctx, slot: c_int, host_rgba8: *mut u8) -> c_int;
    fn rtce_host_alloc(out: *mut *mut c_void, nbytes: usize) -> c_int;
    fn rtce_host_free(p: *mut c_void) -> c_int;
    fn rtce_prepare(ctx: *mut rtce_ctx, frames: c_int, accumulate: c_int) -> c_int;
    fn rtce_render_rgba8_to_host(ctx: *mut rtce_ctx, pcs: *const PushConstantData, frames: c_int, host_rgba8: *mut u8) -> c_int;
    fn rtce_render_rgba8_to_host_async(ctx: *mut rtce_ctx, pcs: *const PushConstantData, frames: c_int, host_rgba8: *mut u8) -> c_int;
    fn rtce_camera_rotation(yaw: f32, pitch: f32, rot_xyzw: *mut f32) -> c_int;
    fn rtce_camera_pose(yaw: f32, pitch: f32, local_delta: *const f32, inout: *mut PushConstantData) -> c_int;
    fn rtce_comm_unique_id(id_out_128: *mut c_void) -> c_int;
    fn rtce_comm_init(ctx: *mut rtce_ctx, id_128: *const c_void, rank: c_int, world: c_int) -> c_int;
    fn rtce_host_band_rows(height: c_int, world: c_int, rank: c_int, y0: *mut c_int, y1: *mut c_int) -> c_int;
    fn rtce_gather_accum(ctx: *mut rtce_ctx, dst: c_int, bands: *const c_int) -> c_int;
    fn rtce_gather_frames(ctx: *mut rtce_ctx, dst: c_int, bands: *const c_int, frames: c_int) -> c_int;
    fn rtce_gather_wait(ctx: *mut rtce_ctx) -> c_int;
    fn rtce_readback_gathered(ctx: *mut rtce_ctx, host: *mut f32, pitch_bytes: usize) -> c_int;
}

pub const RTCE_COMM_ID_BYTES: usize = 128;

/// `Data::rotation()` (src/main.rs:402-404) without glam: `Quat::from_rotation_z(-yaw) * Quat::from_rotation_x(pitch)`
/// as `[x, y, z, w]`, evaluated in f32 by the library (rtce_camera_rotation).
pub fn camera_rotation(yaw: f32, pitch: f32) -> [f32; 4] {
    let mut q = [0f32; 4];
    unsafe { rtce_camera_rotation(yaw, pitch, q.as_mut_ptr()) };
    q
}

/// The per-frame camera update of src/main.rs:770-773: clamps the pitch, writes `pc.rot`, adds the rotated
/// `local_delta` (x right, y forward, z up; `Data::position()`, main.rs:406-414) to `pc.pos`.
pub fn camera_pose(yaw: f32, pitch: f32, local_delta: [f32; 3], pc: &mut PushConstantData) {
    unsafe { rtce_camera_pose(yaw, pitch, local_delta.as_ptr(), pc) };
}

/// NCCL unique id for `ConeMarcher::comm_init` (rank 0 creates it, the host distributes the 128 bytes).
pub fn comm_unique_id() -> Result<[u8; RTCE_COMM_ID_BYTES], Error> {
    let mut id = [0u8; RTCE_COMM_ID_BYTES];
    let st = unsafe { rtce_comm_unique_id(id.as_mut_ptr() as *mut c_void) };
    if st != 0 {
        return Err(Error { status: st, message: "rtce_comm_unique_id (libnccl.so.2 not found?)".to_string() });
    }
    Ok(id)
}

/// Rows `[y0, y1)` of a `height`-row final level owned by `rank` of `world` (whole 8-row dispatch rows, compute.glsl:5).
pub fn band_rows(height: i32, world: i32, rank: i32) -> (i32, i32) {
    let (mut y0, mut y1) = (0, 0);
    unsafe { rtce_host_band_rows(height, world, rank, &mut y0, &mut y1) };
    (y0, y1)
}

#[derive(Debug)]
pub struct Error {
    pub status: i32,
    pub message: String,
}

/// Owns one context = one GPU, one stream; replaces the compute pipeline, its descriptor set and
/// the pyramid images of src/main.rs:154-266, 631-662.
pub struct ConeMarcher {
    ctx: *mut rtce_ctx,
}

impl ConeMarcher {
    /// Device selection + compute pipeline with the RENDER_DIST specialization constant
    /// (src/main.rs:68-92, 450-460, 631-637).
    pub fn new(device: i32, render_dist: f32) -> Result<Self, Error> {
        let mut ctx: *mut rtce_ctx = std::ptr::null_mut();
        let st = unsafe { rtce_create(device, render_dist, &mut ctx) };
        if st != 0 {
            let msg = unsafe { CStr::from_ptr(rtce_status_string(st)) }.to_string_lossy().into_owned();
            return Err(Error { status: st, message: msg });
        }
        Ok(ConeMarcher { ctx })
    }

    fn check(&self, st: c_int) -> Result<(), Error> {
        if st == 0 {
            return Ok(());
        }
        let msg = unsafe { CStr::from_ptr(rtce_last_error(self.ctx)) }.to_string_lossy().into_owned();
        Err(Error { status: st, message: msg })
    }

    /// Binding 1: `DeviceLocalBuffer::from_data(mutable_data)` (src/main.rs:592-605).  Pass
    /// `bytemuck::bytes_of(&mutable_data)`.
    pub fn set_scene(&mut self, mutable_data: &[u8]) -> Result<(), Error> {
        let st = unsafe { rtce_set_scene(self.ctx, mutable_data.as_ptr() as *const c_void, mutable_data.len()) };
        self.check(st)
    }

    /// Binding 2 + the pyramid (src/main.rs:608-615, 639-648); `levels = 0` = the startup rule
    /// (main.rs:639), `-1` = the rule after a resize (main.rs:845).
    pub fn set_view(&mut self, cb: &ConstantBuffer, levels: i32) -> Result<(), Error> {
        let st = unsafe { rtce_set_view(self.ctx, cb, levels) };
        self.check(st)
    }

    /// The compute half of `get_command_buffer` + submit (src/main.rs:291-316, 909-914).
    pub fn trace(&mut self, pc: &PushConstantData) -> Result<(), Error> {
        let st = unsafe { rtce_trace(self.ctx, pc) };
        self.check(st)
    }

    pub fn trace_batch(&mut self, pcs: &[PushConstantData]) -> Result<(), Error> {
        let st = unsafe { rtce_trace_batch(self.ctx, pcs.as_ptr(), pcs.len() as c_int) };
        self.check(st)
    }

    /// What the fragment pass reads (src/main.rs:650-655 -> fragment.glsl:15,135): the last level.
    pub fn final_level(&mut self, slot: i32) -> Result<(usize, usize, Vec<f32>), Error> {
        let (mut w, mut h) = (0, 0);
        let mut p: *mut f32 = std::ptr::null_mut();
        let st = unsafe { rtce_level_info(self.ctx, -1, slot, &mut w, &mut h, &mut p) };
        self.check(st)?;
        let mut out = vec![0f32; (w * h) as usize];
        let st = unsafe { rtce_readback(self.ctx, -1, slot, out.as_mut_ptr(), 0) };
        self.check(st)?;
        Ok((w as usize, h as usize, out))
    }

    /// Fence wait (src/main.rs:882-884).
    pub fn sync(&mut self) -> Result<(), Error> {
        let st = unsafe { rtce_sync(self.ctx) };
        self.check(st)
    }

    /// Frames in flight (src/main.rs:859-884): `w` batches rotate over `w` windows of frame slots
    /// (RTCE_OPT_FRAMES_IN_FLIGHT = 9); default 8; 1 = strict stream order.
    pub fn set_frames_in_flight(&mut self, w: u32) -> Result<(), Error> {
        let st = unsafe { rtce_set_option(self.ctx, 9, w as i64) };
        self.check(st)
    }

    /// Kernel for scenes of more than 8 spheres (RTCE_OPT_LARGE_KERNEL = 12): 0 auto, 2 register version,
    /// 3 event-driven version.  No counterpart in the reference (utilities.glsl:3 caps MAX_OBJECTS at 8).
    pub fn set_large_scene_kernel(&mut self, k: u32) -> Result<(), Error> {
        let st = unsafe { rtce_set_option(self.ctx, 12, k as i64) };
        self.check(st)
    }

    /// Run the pass on a CUDA stream the caller owns (`cudaStream_t` as a pointer); null = the context's own stream.
    pub fn set_stream(&mut self, cuda_stream: *mut c_void) -> Result<(), Error> {
        let st = unsafe { rtce_set_stream(self.ctx, cuda_stream) };
        self.check(st)
    }

    /// The render-pass half of `get_command_buffer` (src/main.rs:318-338: fullscreen triangle, one fragment per
    /// window pixel running fragment.glsl:123-187) for the frames just traced with the same push constants.
    /// Needs the materials and lights of `set_scene`.
    pub fn shade_batch(&mut self, pcs: &[PushConstantData]) -> Result<(), Error> {
        let st = unsafe { rtce_shade_batch(self.ctx, pcs.as_ptr(), pcs.len() as c_int) };
        self.check(st)
    }

    /// The headless stand-in for presenting the swapchain image (src/main.rs:899-916): the shaded frame of `slot`
    /// as RGBA8 UNORM, row-major, window size.
    pub fn color_rgba8(&mut self, slot: i32) -> Result<(usize, usize, Vec<u8>), Error> {
        let (mut w, mut h) = (0, 0);
        let mut p: *mut f32 = std::ptr::null_mut();
        let st = unsafe { rtce_color_info(self.ctx, slot, &mut w, &mut h, &mut p) };
        self.check(st)?;
        let mut out = vec![0u8; (w * h * 4) as usize];
        let st = unsafe { rtce_readback_color_rgba8(self.ctx, slot, out.as_mut_ptr()) };
        self.check(st)?;
        Ok((w as usize, h as usize, out))
    }

    /// Labelled extension (no reference counterpart): FP32 running sum of the final level over `pcs` in frame order,
    /// divided by the frame count; returns the accumulated image (final-level size).
    pub fn trace_accumulate(&mut self, pcs: &[PushConstantData]) -> Result<(usize, usize, Vec<f32>), Error> {
        let st = unsafe { rtce_trace_accumulate(self.ctx, pcs.as_ptr(), pcs.len() as c_int) };
        self.check(st)?;
        let (mut w, mut h) = (0, 0);
        let mut p: *mut f32 = std::ptr::null_mut();
        let st = unsafe { rtce_level_info(self.ctx, -1, 0, &mut w, &mut h, &mut p) };
        self.check(st)?;
        let mut out = vec![0f32; (w * h) as usize];
        let st = unsafe { rtce_readback_accum(self.ctx, out.as_mut_ptr(), 0) };
        self.check(st)?;
        Ok((w as usize, h as usize, out))
    }

    /// Multi-GPU callers: restrict the pass to rows [y0, y1) x columns [x0, x1) of the final level and the ancestors
    /// they touch (one context per GPU; a band traced alone equals the same rows of the whole image bit for bit).
    pub fn set_rect(&mut self, x0: i32, y0: i32, x1: i32, y1: i32) -> Result<(), Error> {
        let st = unsafe { rtce_set_rect(self.ctx, x0, y0, x1, y1) };
        self.check(st)
    }

    /// Order the context's stream after every batch traced so far (no host wait).
    pub fn flush(&mut self) -> Result<(), Error> {
        let st = unsafe { rtce_flush(self.ctx) };
        self.check(st)
    }

    /// Pre-record: allocate every lane's frame slots and capture the launch graphs of a batch of `frames` frames now
    /// (the reference re-records its command buffer every frame, src/main.rs:284-289, 886).
    pub fn prepare(&mut self, frames: i32, accumulate: bool) -> Result<(), Error> {
        let st = unsafe { rtce_prepare(self.ctx, frames, accumulate as c_int) };
        self.check(st)
    }

    /// One loop body of src/main.rs:886-914 per frame, headless: compute pass, shading pass, RGBA8 frames to `out`
    /// (`pcs.len() * view_w * view_h * 4` bytes).
    pub fn render_rgba8(&mut self, pcs: &[PushConstantData], out: &mut [u8]) -> Result<(), Error> {
        let st = unsafe { rtce_render_rgba8_to_host(self.ctx, pcs.as_ptr(), pcs.len() as c_int, out.as_mut_ptr()) };
        self.check(st)
    }

    /// Several GPUs, one `ConeMarcher` per GPU (process or thread): join the communicator of `world` ranks.
    pub fn comm_init(&mut self, id: &[u8; RTCE_COMM_ID_BYTES], rank: i32, world: i32) -> Result<(), Error> {
        let st = unsafe { rtce_comm_init(self.ctx, id.as_ptr() as *const c_void, rank, world) };
        self.check(st)
    }

    /// The one exchange of the accumulating path: every rank's row band of the accumulated image to rank `dst`
    /// (grouped ncclSend / ncclRecv, in place), overlapping whatever is traced next.  `bands` = `[y0, y1]` per rank.
    pub fn gather_accum(&mut self, dst: i32, bands: Option<&[[i32; 2]]>) -> Result<(), Error> {
        let p = bands.map_or(std::ptr::null(), |b| b.as_ptr() as *const c_int);
        let st = unsafe { rtce_gather_accum(self.ctx, dst, p) };
        self.check(st)
    }

    /// The frame-by-frame form: the bands of frame slots `0 .. frames` of the latest batch, in place into `dst`'s slots.
    pub fn gather_frames(&mut self, dst: i32, bands: Option<&[[i32; 2]]>, frames: i32) -> Result<(), Error> {
        let p = bands.map_or(std::ptr::null(), |b| b.as_ptr() as *const c_int);
        let st = unsafe { rtce_gather_frames(self.ctx, dst, p, frames) };
        self.check(st)
    }

    /// Rank `dst`: the gathered image (waits for the exchange).
    pub fn gathered(&mut self) -> Result<(usize, usize, Vec<f32>), Error> {
        let (mut w, mut h) = (0, 0);
        let mut p: *mut f32 = std::ptr::null_mut();
        let st = unsafe { rtce_level_info(self.ctx, -1, 0, &mut w, &mut h, &mut p) };
        self.check(st)?;
        let mut out = vec![0f32; (w * h) as usize];
        let st = unsafe { rtce_readback_gathered(self.ctx, out.as_mut_ptr(), 0) };
        self.check(st)?;
        Ok((w as usize, h as usize, out))
    }
}

impl Drop for ConeMarcher {
    fn drop(&mut self) {
        unsafe { rtce_destroy(self.ctx) }
    }
}

/// Pinned host memory for `render_to_host` targets (cudaHostAlloc behind the ABI, so the Rust side needs no CUDA crate).
pub struct PinnedF32 {
    ptr: *mut f32,
    len: usize,
}

impl PinnedF32 {
    pub fn new(len: usize) -> Result<Self, Error> {
        let mut p: *mut c_void = std::ptr::null_mut();
        let st = unsafe { rtce_host_alloc(&mut p, len * std::mem::size_of::<f32>()) };
        if st != 0 {
            let msg = unsafe { CStr::from_ptr(rtce_status_string(st)) }.to_string_lossy().into_owned();
            return Err(Error { status: st, message: msg });
        }
        Ok(PinnedF32 { ptr: p as *mut f32, len })
    }
    pub fn as_slice(&self) -> &[f32] {
        unsafe { std::slice::from_raw_parts(self.ptr, self.len) }
    }
    pub fn as_mut_ptr(&mut self) -> *mut f32 {
        self.ptr
    }
}

impl Drop for PinnedF32 {
    fn drop(&mut self) {
        unsafe {
            rtce_host_free(self.ptr as *mut c_void);
        }
    }
}

/// Headless replacement of the reference's `main` loop body (src/main.rs:778-927 without window, swapchain and
/// input): trace + shade one frame of the scene `mutable_data` (the bytes src/main.rs:524-605 builds) and return
/// the RGBA8 image.  The camera is the reference's own state (`Data::rotation`: yaw, pitch; the movement of this
/// frame in camera axes), turned into push constants by `camera_pose` exactly as main.rs:770-773 does.
pub fn render_headless(mutable_data: &[u8], width: f32, height: f32, fov: f32, render_dist: f32,
                       yaw: f32, pitch: f32, start_pos: [f32; 3], local_delta: [f32; 3])
                       -> Result<(usize, usize, Vec<u8>), Error> {
    let mut m = ConeMarcher::new(0, render_dist)?;
    m.set_scene(mutable_data)?;
    let cb = ConstantBuffer { view: [width, height], ratio: [fov, fov * height / width] };   // main.rs:608-611
    m.set_view(&cb, 0)?;
    let mut pc = PushConstantData { rot: [0.0, 0.0, 0.0, 1.0], pos: start_pos, iter: 0, imageSize: [0.0; 2] };
    camera_pose(yaw, pitch, local_delta, &mut pc);   // main.rs:770-773
    let mut out = vec![0u8; (width as usize) * (height as usize) * 4];
    m.render_rgba8(std::slice::from_ref(&pc), &mut out)?;   // main.rs:886-914: compute half, render-pass half, "present"
    Ok((width as usize, height as usize, out))
}
