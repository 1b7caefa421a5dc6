//! Raw FFI declarations for `librm_b200.so` (`include/rm_b200.h`), one `pub fn` per exported symbol.
//!
//! This is the file a maintainer of the reference would drop in as `src/ray_marching/rm_sys.rs` (INTEGRATION.md shows
//! the two call sites in `renderer.rs` that change).  There is no Rust toolchain in the build image, so it is NOT
//! compiled here; `tests/test_abi_cpu.py` keeps its function list and argument counts in step with the C header.
//! Byte blocks (`uniforms144`, `limits12`) are exactly the reference's `as_shader_bytes()` output
//! (`renderer.rs:21-27`), `words` is `CSGCommandBufferBuilder::buffer` (`builder.rs:27-30`).
#![allow(non_camel_case_types, dead_code)]

use core::ffi::{c_char, c_void};

#[repr(C)] pub struct RmCtx { _private: [u8; 0] }
#[repr(C)] pub struct RmGroup { _private: [u8; 0] }

pub const RM_OK: i32 = 0;
pub const RM_E_INVALID_ARG: i32 = -1;
pub const RM_E_JIT: i32 = -9;
pub const RM_E_UNSUPPORTED: i32 = -10;
// command opcodes beyond the reference's four (builder.rs:1-24): the values its #[repr(u32)] enum reserves in comments
pub const RM_OP_PLANE: u32 = 2;
pub const RM_OP_INTERSECTION: u32 = 102;
pub const RM_OP_SMOOTH_UNION: u32 = 103;
pub const RM_OP_TRANSLATION_PUSH: u32 = 200;
pub const RM_OP_TRANSLATION_POP: u32 = 201;
pub const RM_OP_ROTATION_PUSH: u32 = 202;
pub const RM_OP_ROTATION_POP: u32 = 203;
pub const RM_OP_SCALE_PUSH: u32 = 204;
pub const RM_OP_SCALE_POP: u32 = 205;
pub const RM_MODE_EXACT: i32 = 0;
pub const RM_MODE_FAST: i32 = 1;
pub const RM_KERNEL_WAVEFRONT: i32 = 0;
pub const RM_KERNEL_SIMPLE: i32 = 1;
pub const RM_KERNEL_WAVEFRONT2: i32 = 2;
pub const RM_KERNEL_DIRECT: i32 = 3;
pub const RM_UNIFORMS_BYTES: usize = 144;
pub const RM_LIMITS_BYTES: usize = 12;
pub const RM_IPC_HANDLE_BYTES: usize = 64;

#[link(name = "rm_b200")]
extern "C" {
    // lifetime — RayMarchingResources::new / Drop (renderer.rs:51-175, main.rs:30-34)
    pub fn rm_create(cuda_device: i32, out: *mut *mut RmCtx) -> i32;
    pub fn rm_create_ex(cuda_device: i32, max_words: u32, out: *mut *mut RmCtx) -> i32;
    pub fn rm_destroy(ctx: *mut RmCtx);
    pub fn rm_set_limits(ctx: *mut RmCtx, limits12: *const u8) -> i32;
    pub fn rm_set_mode(ctx: *mut RmCtx, mode: i32) -> i32;
    pub fn rm_get_mode(ctx: *mut RmCtx) -> i32;
    pub fn rm_set_kernel(ctx: *mut RmCtx, kernel: i32) -> i32;
    pub fn rm_set_early_out(ctx: *mut RmCtx, on: i32) -> i32;

    // per frame — RayMarchingCallback::prepare (renderer.rs:196-242)
    pub fn rm_prepare(ctx: *mut RmCtx, uniforms144: *const u8, cmd_count: u32, words: *const u32, n_words: u32) -> i32;
    pub fn rm_prepare_uniforms(ctx: *mut RmCtx, uniforms144: *const u8) -> i32;
    pub fn rm_program_uploads(ctx: *mut RmCtx) -> u64;
    pub fn rm_validate_program(cmd_count: u32, words: *const u32, n_words: u32, max_words: u32) -> i32;

    // program specialisation: compile the scene into the kernel (optional)
    pub fn rm_specialize(ctx: *mut RmCtx) -> i32;
    pub fn rm_set_specialize(ctx: *mut RmCtx, on: i32) -> i32;   // 0 interpreter, 1 at first paint, 2 in the background
    pub fn rm_spec_info(ctx: *mut RmCtx, active: *mut i32, rays_per_thread: *mut i32, min_blocks: *mut i32,
                        compile_ms: *mut f64, cubin_bytes: *mut u64, compiles: *mut u64, launches: *mut u64) -> i32;
    pub fn rm_spec_status(ctx: *mut RmCtx) -> *const c_char;
    pub fn rm_spec_source(cmd_count: u32, words: *const u32, n_words: u32, mode: i32, rays_per_thread: i32,
                          buf: *mut c_char, cap: usize, needed: *mut usize) -> i32;
    pub fn rm_spec_compile_check(cmd_count: u32, words: *const u32, n_words: u32, mode: i32, rays_per_thread: i32,
                                 cubin_bytes: *mut u64, compile_ms: *mut f64, log: *mut c_char, log_cap: usize) -> i32;
    pub fn rm_spec_params(cmd_count: u32, words: *const u32, n_words: u32, mode: i32, buf: *mut f32, cap_floats: usize,
                          needed: *mut usize) -> i32;
    pub fn rm_set_cull(ctx: *mut RmCtx, on: i32) -> i32;
    pub fn rm_get_cull_stats(ctx: *mut RmCtx, executed_map_flops: *mut u64, culled_paints: *mut u64, reset: i32) -> i32;
    pub fn rm_paint_culled_static(ctx: *mut RmCtx, w: u32, h: u32, row0: u32, nrows: u32, dst_rgba8_dev: *mut c_void,
                                  pitch_bytes: usize, cuda_stream: *mut c_void) -> i32;
    pub fn rm_set_spec_literal_after(ctx: *mut RmCtx, n_paints: u32) -> i32;
    pub fn rm_spec_tier(ctx: *mut RmCtx) -> i32;
    pub fn rm_spec_cache(new_capacity: u64, entries: *mut u64, capacity: *mut u64, evictions: *mut u64) -> i32;

    // per frame — RayMarchingCallback::paint (renderer.rs:244-255)
    pub fn rm_paint(ctx: *mut RmCtx, w: u32, h: u32, row0: u32, nrows: u32, dst_rgba8_dev: *mut c_void,
                    pitch_bytes: usize, cuda_stream: *mut c_void) -> i32;
    pub fn rm_paint_strips(ctx: *mut RmCtx, w: u32, h: u32, strip_rows: u32, first_strip: u32, strip_step: u32,
                           dst_rgba8_dev: *mut c_void, pitch_bytes: usize, cuda_stream: *mut c_void,
                           rows_painted: *mut u32) -> i32;
    pub fn rm_paint_strips_into(ctx: *mut RmCtx, w: u32, h: u32, strip_rows: u32, first_strip: u32, strip_step: u32,
                                frame_rgba8_dev: *mut c_void, pitch_bytes: usize, cuda_stream: *mut c_void,
                                rows_painted: *mut u32) -> i32;
    pub fn rm_shared_frame_create(ctx: *mut RmCtx, bytes: usize, dev_ptr: *mut *mut c_void, handle64: *mut u8) -> i32;
    pub fn rm_shared_frame_open(ctx: *mut RmCtx, handle64: *const u8, dev_ptr: *mut *mut c_void) -> i32;
    pub fn rm_shared_frame_close(ctx: *mut RmCtx, dev_ptr: *mut c_void) -> i32;
    pub fn rm_copy_to_host(ctx: *mut RmCtx, dev_ptr: *const c_void, host: *mut c_void, bytes: usize,
                           cuda_stream: *mut c_void) -> i32;
    pub fn rm_copy_strips_to_host(ctx: *mut RmCtx, band_dev: *const c_void, w: u32, h: u32, strip_rows: u32, first_strip: u32,
                                  strip_step: u32, host_frame: *mut c_void, cuda_stream: *mut c_void) -> i32;
    pub fn rm_paint_poses(ctx: *mut RmCtx, uniforms144_array: *const u8, n_poses: u32, w: u32, h: u32,
                          dst_rgba8_dev: *mut c_void, pitch_bytes: usize, frame_stride_bytes: usize,
                          cuda_stream: *mut c_void) -> i32;
    pub fn rm_paint_poses_host(ctx: *mut RmCtx, uniforms144_array: *const u8, n_poses: u32, w: u32, h: u32,
                               host_rgba8: *mut c_void, host_frame_stride_bytes: usize, host_ring_frames: u32) -> i32;
    pub fn rm_download(ctx: *mut RmCtx, host_rgba8: *mut c_void, bytes: usize) -> i32;
    pub fn rm_sync(ctx: *mut RmCtx) -> i32;

    // several GPUs from one host thread
    pub fn rm_group_create(cuda_devices: *const i32, n: i32, out: *mut *mut RmGroup) -> i32;
    pub fn rm_group_destroy(g: *mut RmGroup);
    pub fn rm_group_size(g: *mut RmGroup) -> i32;
    pub fn rm_group_ctx(g: *mut RmGroup, i: i32) -> *mut RmCtx;
    pub fn rm_group_set_mode(g: *mut RmGroup, mode: i32) -> i32;
    pub fn rm_group_set_limits(g: *mut RmGroup, limits12: *const u8) -> i32;
    pub fn rm_group_specialize(g: *mut RmGroup) -> i32;
    pub fn rm_group_prepare(g: *mut RmGroup, uniforms144: *const u8, cmd_count: u32, words: *const u32, n_words: u32) -> i32;
    pub fn rm_group_paint_bands(g: *mut RmGroup, w: u32, h: u32, strip_rows: u32, host_rgba8: *mut c_void) -> i32;
    pub fn rm_group_frame(g: *mut RmGroup, dev_ptr: *mut *mut c_void, pitch_bytes: *mut usize) -> i32;
    pub fn rm_group_paint_poses(g: *mut RmGroup, uniforms144_array: *const u8, n_poses: u32, w: u32, h: u32,
                                host_rgba8: *mut c_void, host_frame_stride_bytes: usize) -> i32;
    pub fn rm_group_last_error(g: *mut RmGroup) -> *const c_char;

    // verification surface (untimed)
    pub fn rm_paint_debug(ctx: *mut RmCtx, w: u32, h: u32, row0: u32, nrows: u32, cls_host: *mut u8, t_host: *mut f32,
                          counters4: *mut u64, rgba_host: *mut u8) -> i32;
    pub fn rm_paint_aov(ctx: *mut RmCtx, w: u32, h: u32, row0: u32, nrows: u32, cls_host: *mut u8, t_host: *mut f32,
                        rgba_host: *mut u8, specialised: *mut i32) -> i32;
    pub fn rm_map(ctx: *mut RmCtx, pts_host: *const f32, n: u64, out_host: *mut f32) -> i32;
    pub fn rm_set_counters(ctx: *mut RmCtx, enable: i32) -> i32;
    pub fn rm_get_counters(ctx: *mut RmCtx, counters4: *mut u64, reset: i32) -> i32;

    // measurement helpers
    pub fn rm_mark(ctx: *mut RmCtx, slot: i32) -> i32;
    pub fn rm_elapsed_ms(ctx: *mut RmCtx, slot_a: i32, slot_b: i32, ms: *mut f32) -> i32;
    pub fn rm_flush_l2(ctx: *mut RmCtx) -> i32;
    pub fn rm_fp32_peak(ctx: *mut RmCtx, tflops: *mut f64, sm_count: *mut i32, sm_clock_khz: *mut i32) -> i32;
    pub fn rm_launch_count(ctx: *mut RmCtx) -> u64;
    pub fn rm_host_alloc(bytes: usize) -> *mut c_void;
    pub fn rm_host_free(p: *mut c_void);
    pub fn rm_host_register(p: *mut c_void, bytes: usize) -> i32;
    pub fn rm_host_unregister(p: *mut c_void) -> i32;

    pub fn rm_last_error(ctx: *mut RmCtx) -> *const c_char;
    pub fn rm_strerror(code: i32) -> *const c_char;
    pub fn rm_version() -> *const c_char;
}
