//! `extern "C"` declarations of include/rtw.h (ABI version 3). SOURCE ONLY: never compiled in this
//! repository's build image (no Rust toolchain). Field order and types mirror the C structs;
//! tests/test_rust_shim.py holds every struct's field list and every function name below against
//! raytracing-in-a-weekend_b200/_abi.py (which tests/test_abi.py holds against the built library).
#![allow(non_camel_case_types, dead_code)]
use std::os::raw::{c_char, c_double, c_int, c_void};

pub const RTW_ABI_VERSION: u32 = 3;
pub const RTW_OK: c_int = 0;
pub const RTW_E_INVALID: c_int = -1;
pub const RTW_E_CUDA: c_int = -2;
pub const RTW_E_NOMEM: c_int = -3;
pub const RTW_E_UNSUPPORTED: c_int = -4;

#[repr(C)] #[derive(Clone, Copy)]
pub struct rtw_material { pub metallicness: f32, pub opacity: f32, pub ir: f32, pub emitted: [f32; 3] }
#[repr(C)]
pub struct rtw_noise { pub ranvec: [[f32; 3]; 256], pub perm_x: [u32; 256], pub perm_y: [u32; 256], pub perm_z: [u32; 256] }
#[repr(C)] #[derive(Clone, Copy)]
pub struct rtw_texture { pub width: u32, pub height: u32, pub rgb: *const f32, pub noise_id: i32, pub noise_scale: f32 }
#[repr(C)] #[derive(Clone, Copy)]
pub struct rtw_sphere { pub origin: [f32; 3], pub radius: f32, pub col_mod: [f32; 3], pub velocity: [f32; 3], pub mat_id: u32, pub tex_id: u32 }
#[repr(C)] #[derive(Clone, Copy)]
pub struct rtw_quad { pub origin: [f32; 3], pub u: [f32; 3], pub v: [f32; 3], pub velocity: [f32; 3], pub mat_id: u32, pub tex_id: u32, pub shape: u32 }
pub const RTW_SHAPE_PARALLELOGRAM: u32 = 0;
pub const RTW_SHAPE_TRIANGLE: u32 = 1;
pub const RTW_DENSITY_SURFACE: u32 = 0;
pub const RTW_DENSITY_CONST: u32 = 1;
#[repr(C)] #[derive(Clone, Copy)]
pub struct rtw_instance {
    pub first_sphere: u32, pub n_spheres: u32, pub first_quad: u32, pub n_quads: u32,
    pub translation: [f32; 3], pub rotation: [f32; 3], pub density_kind: u32, pub density: f32,
}
#[repr(C)] #[derive(Clone, Copy)]
pub struct rtw_light { pub kind: u32, pub index: u32 }
pub const RTW_LIGHT_SPHERE: u32 = 0;
pub const RTW_LIGHT_QUAD: u32 = 1;
#[repr(C)]
pub struct rtw_scene_desc {
    pub abi_version: u32,
    pub n_spheres: u32, pub n_top_spheres: u32, pub n_quads: u32, pub n_top_quads: u32, pub n_instances: u32,
    pub n_materials: u32, pub n_textures: u32, pub n_noises: u32,
    pub spheres: *const rtw_sphere, pub quads: *const rtw_quad, pub instances: *const rtw_instance,
    pub materials: *const rtw_material, pub textures: *const rtw_texture, pub noises: *const rtw_noise,
    pub background: [f32; 3], pub n_lights: u32, pub lights: *const rtw_light,
}
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct rtw_camera {
    pub width: u32, pub height: u32, pub origin: [f32; 3], pub pixel00: [f32; 3],
    pub delta_u: [f32; 3], pub delta_v: [f32; 3], pub u: [f32; 3], pub v: [f32; 3], pub lens_radius: f32,
}
#[repr(C)] #[derive(Clone, Copy)]
pub struct rtw_render_opts {
    pub samples: u32, pub depth: u32, pub gamma: f32, pub integrator: u32, pub mode: u32,
    pub shutter_speed: f32, pub fps: f32, pub frame: u32, pub seed: u64, pub mint: f32, pub maxt: f32,
    pub sample_begin: u32, pub sample_count: u32, pub row_begin: u32, pub row_count: u32,
    pub output: u32, pub flags: u32, pub light_weight: f32, pub row_band: u32, pub row_stride: u32,
}
pub const RTW_OUT_RESOLVED: u32 = 0;
pub const RTW_OUT_SUM: u32 = 1;
pub const RTW_FLAG_STRICT_FP: u32 = 1;
pub const RTW_FLAG_WAVEFRONT: u32 = 2;
pub const RTW_FLAG_COUNT_WORK: u32 = 4;
pub const RTW_FLAG_MEGAKERNEL: u32 = 8;
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct rtw_stats {
    pub paths: u64, pub rays: u64, pub nan_paths: u64, pub gpu_launches: u64, pub kernel_ms: f32, pub total_ms: f32,
    pub nodes: u64, pub sphere_tests: u64, pub quad_tests: u64, pub instance_tests: u64, pub hits: u64,
    pub iterations: u64, pub extend_ms: f32, pub shade_ms: f32,
}
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct rtw_scene_info {
    pub n_tlas_nodes: u32, pub n_blas_nodes: u32, pub n_tlas_leaves: u32, pub n_flattened_instances: u32,
    pub bvh_max_depth: u32, pub n_baked_instances: u32, pub device_bytes: u64, pub h2d_bytes: u64,
}
pub enum rtw_scene {}
pub enum rtw_scene_multi {}
pub const RTW_SPLIT_SAMPLES: u32 = 0;
pub const RTW_SPLIT_TILES: u32 = 1;
pub const RTW_SPLIT_FRAMES: u32 = 2;
pub const RTW_PROX_SQUARE: c_int = 0;
pub const RTW_PROX_EDGES: c_int = 1;
pub const RTW_NOALLOC_MAX_INTERIOR_PIXELS: u32 = 1_046_528;
pub const RTW_NOALLOC_MAX_WINDOW: u32 = 16;
pub const RTW_BENCH_FP32: c_int = 0;
pub const RTW_BENCH_L2: c_int = 1;
pub const RTW_BENCH_HBM: c_int = 2;

extern "C" {
    pub fn rtw_abi_version() -> c_int;
    pub fn rtw_last_error() -> *const c_char;
    pub fn rtw_device_count() -> c_int;
    /// Viewport::new / new_from_res (viewport.rs:308-428) restated; a Rust caller normally copies the Viewport's
    /// own derived fields instead (gpu.rs `gpu_camera`).
    pub fn rtw_camera_init(cam: *mut rtw_camera, width: u32, aspect_ratio: f32, vfov_deg: f32, origin: *const f32,
                           direction: *const f32, vup: *const f32, lens_radius: f32) -> c_int;
    pub fn rtw_camera_init_res(cam: *mut rtw_camera, width: u32, height: u32, vfov_deg: f32, origin: *const f32,
                               direction: *const f32, vup: *const f32, lens_radius: f32) -> c_int;
    pub fn rtw_render_opts_default(o: *mut rtw_render_opts);
    pub fn rtw_scene_create(desc: *const rtw_scene_desc, device: c_int, out: *mut *mut rtw_scene) -> c_int;
    pub fn rtw_scene_destroy(scene: *mut rtw_scene);
    /// Releases the device / pinned blocks kept from destroyed scenes (`device` < 0: every device).
    pub fn rtw_trim_memory(device: c_int) -> c_int;
    /// Optional on-device LBVH rebuild of every BVH level / back to the host SAH tree. The scene must be idle.
    pub fn rtw_scene_rebuild_lbvh(scene: *mut rtw_scene, out_build_ms: *mut f32) -> c_int;
    pub fn rtw_scene_use_sah(scene: *mut rtw_scene) -> c_int;
    pub fn rtw_render(scene: *mut rtw_scene, cam: *const rtw_camera, opts: *const rtw_render_opts, out_rgb: *mut f32) -> c_int;
    pub fn rtw_render_device(scene: *mut rtw_scene, cam: *const rtw_camera, opts: *const rtw_render_opts, d_out: *mut f32, stream: *mut c_void) -> c_int;
    pub fn rtw_render_frames(scene: *mut rtw_scene, cam: *const rtw_camera, opts: *const rtw_render_opts, first: u32, n: u32, out_rgb: *mut f32) -> c_int;
    pub fn rtw_resolve_device(d_rgb: *mut f32, n_pixels: u64, total_samples: u32, gamma: f32, device: c_int, stream: *mut c_void) -> c_int;
    pub fn rtw_quantize_device(d_rgb: *const f32, d_rgb8: *mut u8, n_values: u64, device: c_int, stream: *mut c_void) -> c_int;
    /// write_img_f32's quantiser (write_img.rs:12-14) for a host image, run on the device.
    pub fn rtw_quantize(rgb: *const f32, rgb8: *mut u8, n_values: u64, device: c_int) -> c_int;
    /// rtw_render + that quantiser fused on the device; `out_rgb8` is host memory (H*W*3 bytes).
    pub fn rtw_render_rgb8(scene: *mut rtw_scene, cam: *const rtw_camera, opts: *const rtw_render_opts, out_rgb8: *mut u8) -> c_int;
    pub fn rtw_bilateral_filter_device(d_rgb8: *const u8, width: u32, height: u32, size: u32, proximity: c_int, d_out_rgb8: *mut u8,
                                       d_params: *mut f32, device: c_int, stream: *mut c_void) -> c_int;
    /// The device filter with its scratch in a static device array (no allocation inside the call); limits
    /// RTW_NOALLOC_MAX_INTERIOR_PIXELS / RTW_NOALLOC_MAX_WINDOW. No counterpart in the reference (include/rtw.h).
    pub fn rtw_bilateral_filter_device_noalloc(d_rgb8: *const u8, width: u32, height: u32, size: u32, proximity: c_int, d_out_rgb8: *mut u8,
                                               d_params: *mut f32, device: c_int, stream: *mut c_void) -> c_int;
    /// Rust2's `postprocessing::bilateral_filter(&img, Proximity::new(size, type))` (postprocessing.rs:70-135) on the
    /// device: `rgb8` / `out_rgb8` are the raw buffers of `ImageBuffer<Rgb<u8>, Vec<u8>>` (`img.as_raw()`),
    /// proximity 0 = Square, 1 = Edges; `out_params` / `out_kernel_ms` may be null.
    pub fn rtw_bilateral_filter(rgb8: *const u8, width: u32, height: u32, size: u32, proximity: c_int, out_rgb8: *mut u8,
                                out_params: *mut f32, out_kernel_ms: *mut f32, device: c_int) -> c_int;
    /// Several devices of this process behind one call (SURVEY 8b device list + split; include/rtw.h).
    pub fn rtw_scene_multi_create(desc: *const rtw_scene_desc, devices: *const c_int, n_devices: c_int, out: *mut *mut rtw_scene_multi) -> c_int;
    pub fn rtw_scene_multi_destroy(ms: *mut rtw_scene_multi);
    pub fn rtw_render_multi(ms: *mut rtw_scene_multi, cam: *const rtw_camera, opts: *const rtw_render_opts, split: u32, band_rows: u32,
                            out_rgb: *mut f32) -> c_int;
    pub fn rtw_render_frames_multi(ms: *mut rtw_scene_multi, cam: *const rtw_camera, opts: *const rtw_render_opts, first: u32, n: u32,
                                   out_rgb: *mut f32) -> c_int;
    pub fn rtw_multi_get_stats(ms: *mut rtw_scene_multi, total: *mut rtw_stats, per_device: *mut rtw_stats, combine_ms: *mut f32) -> c_int;
    pub fn rtw_scene_multi_device_count(ms: *const rtw_scene_multi) -> c_int;
    pub fn rtw_get_stats(scene: *mut rtw_scene, out: *mut rtw_stats) -> c_int;
    pub fn rtw_get_sched_stats(scene: *mut rtw_scene, out: *mut u64) -> c_int;
    pub fn rtw_microbench(device: c_int, which: c_int, reps: c_int, out_value: *mut c_double) -> c_int;
    pub fn rtw_scene_get_info(scene: *mut rtw_scene, out: *mut rtw_scene_info) -> c_int;
}
