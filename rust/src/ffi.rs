//! extern "C" declarations of include/rtb200.h (librtb200.so, sm_100a CUDA behind a C ABI).
//! SOURCE ONLY: the build image has no Rust toolchain; the struct layouts are checked against the header by
//! tests/test_host.py::test_struct_layouts_match_header through the Python mirror of the same structs.
#![allow(dead_code)]
use std::os::raw::{c_char, c_void};

pub const RTB_OK: i32 = 0;
pub const RTB_FLAG_COUNTERS: u32 = 1;
pub const RTB_FLAG_PROFILE: u32 = 2;
pub const RTB_FLAG_FRAME_ROWS: u32 = 4;
pub const RTB_FLAG_GATHER_PEER: u32 = 8;
pub const RTB_OBJ_MTL: u32 = 1;
pub const RTB_OBJ_PINNED: u32 = 2;
pub const RTB_COLOR_STATIC_MAX: u64 = 1 << 20;
pub const RTB_TIE_REFERENCE: u32 = 0;
pub const RTB_TIE_CALLER: u32 = 1;
pub const RTB_TIE_INDEX: u32 = 2;

#[repr(C)]
pub struct RtbScene { _private: [u8; 0] }

#[repr(C)]
pub struct RtbSceneDesc {
    pub n_tris: u32,
    pub verts: *const f32,      // n_tris * 9
    pub colors: *const f32,     // n_tris * 3
    pub tie_rank: *const i32,   // optional
    pub light: [f32; 9],
    pub device: i32,
}

#[repr(C)]
#[derive(Default)]
pub struct RtbSceneOptions {
    pub struct_size: u32, pub shadow_mode: i32, pub primary_mode: i32, pub term_mode: i32,
    pub chunk_rays: u64, pub term_bytes: u64, pub grid_entry_cap: u64,
    pub grid_cell_factor: f32, pub cam_cell_px: f32, pub cdiv_off: i32, pub trace_chunks: i32,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct RtbCameraPose { pub eye: [f32; 3], pub look_at: [f32; 3], pub up: [f32; 3], pub distance: f32 }

#[repr(C)]
#[derive(Default)]
pub struct RtbImportStats { pub read_s: f64, pub parse_s: f64, pub alloc_s: f64, pub file_bytes: u64, pub threads: u32, pub n_vertices: u32 }

#[repr(C)]
#[derive(Default)]
pub struct RtbSequenceStats {
    pub seconds: f64, pub device_ms: f32, pub frames: u32, pub cam_grid_redo: u32,
    pub primary_rays: u64, pub shadow_rays: u64,
}

pub type RtbFrameFn = Option<unsafe extern "C" fn(user: *mut c_void, pose_index: u32, rgb8: *const u8, width: u32, rows: u32)>;

#[repr(C)]
#[derive(Clone, Copy)]
pub struct RtbRenderParams {
    pub width: u32, pub height: u32, pub nb_ray: u32, pub nb_light_sample: u32,
    pub eye: [f32; 3], pub look_at: [f32; 3], pub up: [f32; 3], pub distance: f32,
    pub band_rows: u32, pub rank: u32, pub world: u32, pub flags: u32,
}

#[repr(C)]
#[derive(Default)]
pub struct RtbStats {
    pub primary_rays: u64, pub primary_hits: u64, pub shadow_rays: u64, pub shadow_occluded: u64,
    pub box_tests: u64, pub tri_tests: u64, pub kernel_ms: f32, pub launches: u32,
    pub primary_ms: f32, pub shadow_ms: f32, pub shade_ms: f32, pub queued_hits: u32,
    pub queued_groups: u32, pub queued_walks: u32,
    pub grid_cells: u32, pub grid_entries: u32, pub grid_big: u32, pub grid_build_ms: f32,
    pub grid_rebuilds: u32, pub grid_redo: u32,
}

#[repr(C)]
#[derive(Default)]
pub struct RtbSceneInfo {
    pub n_tris: u32, pub n_nodes: u32, pub max_depth: u32, pub ploc_iterations: u32,
    pub sah_cost: f32, pub build_ms: f32, pub tie_mode: u32,
}

#[link(name = "rtb200")]
extern "C" {
    pub fn rtb_scene_create(desc: *const RtbSceneDesc, out: *mut *mut RtbScene) -> i32;
    pub fn rtb_scene_create_ex(desc: *const RtbSceneDesc, options: *const RtbSceneOptions, out: *mut *mut RtbScene) -> i32;
    pub fn rtb_scene_destroy(scene: *mut RtbScene);
    pub fn rtb_scene_get_info(scene: *const RtbScene, info: *mut RtbSceneInfo) -> i32;
    pub fn rtb_scene_set_materials(scene: *mut RtbScene, ks_ns: *const f32) -> i32;
    pub fn rtb_scene_set_samples(scene: *mut RtbScene, samples: *const f32, n_pairs: u32) -> i32;
    pub fn rtb_scene_set_samples_async(scene: *mut RtbScene, samples: *const f32, n_pairs: u32) -> i32;
    pub fn rtb_scenes_set_samples(scenes: *const *mut RtbScene, n_gpus: u32, samples: *const f32, n_pairs: u32) -> i32;
    pub fn rtb_render_multi(scenes: *const *mut RtbScene, n_gpus: u32, params: *const RtbRenderParams, rgb8: *mut u8, stats: *mut RtbStats) -> i32;
    pub fn rtb_render_sequence(scene: *mut RtbScene, params: *const RtbRenderParams, poses: *const RtbCameraPose, n_poses: u32,
                               frames: *mut u8, png_pattern: *const c_char, on_frame: RtbFrameFn, user: *mut c_void,
                               stats: *mut RtbSequenceStats) -> i32;
    pub fn rtb_render_sequence_multi(scenes: *const *mut RtbScene, n_gpus: u32, params: *const RtbRenderParams, poses: *const RtbCameraPose,
                                     n_poses: u32, frames: *mut u8, png_pattern: *const c_char, stats: *mut RtbSequenceStats) -> i32;
    pub fn rtb_measure_peaks(device: i32, fp32_tflops: *mut f32, l2_read_gbs: *mut f32, hbm_read_gbs: *mut f32) -> i32;
    pub fn rtb_render(scene: *mut RtbScene, params: *const RtbRenderParams, rgb8: *mut u8, stats: *mut RtbStats) -> i32;
    pub fn rtb_render_begin(scene: *mut RtbScene, params: *const RtbRenderParams, rgb8: *mut u8) -> i32;
    pub fn rtb_render_end(scene: *mut RtbScene, stats: *mut RtbStats) -> i32;
    pub fn rtb_render_device(scene: *mut RtbScene, params: *const RtbRenderParams, d_rgb8: *mut c_void, cuda_stream: *mut c_void) -> i32;
    pub fn rtb_render_pixels(scene: *mut RtbScene, params: *const RtbRenderParams, pixels: *const u32, n_pix: u64,
                             rgb_lin: *mut f32, rgb8: *mut u8, hit_count: *mut u8, prim0: *mut i32, t0: *mut f32,
                             stats: *mut RtbStats) -> i32;
    pub fn rtb_color_to_u8(scene: *mut RtbScene, linear: *const f32, n: u64, out: *mut u8) -> i32;
    pub fn rtb_color_to_u8_static(scene: *mut RtbScene, linear: *const f32, n: u64, out: *mut u8) -> i32;
    pub fn rtb_scene_get_counters(scene: *mut RtbScene, out16: *mut u64) -> i32;
    pub fn rtb_selftest_division(scene: *mut RtbScene, seed: u64, n: u64, mismatches: *mut u64) -> i32;
    pub fn rtb_scene_get_tie_ranks(scene: *mut RtbScene, out: *mut i32) -> i32;
    pub fn rtb_scene_get_nodes(scene: *mut RtbScene, out: *mut f32) -> i32;
    pub fn rtb_owned_rows(height: u32, band_rows: u32, rank: u32, world: u32) -> u32;
    pub fn rtb_deinterleave_bands(device: i32, d_staged: *const c_void, stride_bytes: u64, d_frame: *mut c_void,
                                  width: u32, height: u32, band_rows: u32, world: u32, cuda_stream: *mut c_void) -> i32;
    pub fn rtb_peer_alloc(device: i32, bytes: u64, d_ptr_out: *mut *mut c_void, handle_out: *mut u8) -> i32;
    pub fn rtb_peer_open(device: i32, handle: *const u8, d_ptr_out: *mut *mut c_void) -> i32;
    pub fn rtb_peer_close(device: i32, d_ptr: *mut c_void, owner: i32) -> i32;
    pub fn rtb_import_obj(path: *const c_char, verts_out: *mut *mut f32, n_out: *mut u32) -> i32;
    pub fn rtb_import_obj_ex(path: *const c_char, flags: u32, verts_out: *mut *mut f32, kd_out: *mut *mut f32,
                             ks_ns_out: *mut *mut f32, n_out: *mut u32) -> i32;
    pub fn rtb_import_obj_parallel(path: *const c_char, n_threads: u32, flags: u32, verts_out: *mut *mut f32, n_out: *mut u32,
                                   stats: *mut RtbImportStats) -> i32;
    pub fn rtb_import_obj_parallel_ex(path: *const c_char, n_threads: u32, flags: u32, verts_out: *mut *mut f32, kd_out: *mut *mut f32,
                                      ks_ns_out: *mut *mut f32, n_out: *mut u32, stats: *mut RtbImportStats) -> i32;
    pub fn rtb_free_pinned(p: *mut c_void);
    pub fn rtb_spheres_closest_hit(device: i32, spheres_xyzr: *const f32, n_spheres: u32, origins: *const f32, dirs: *const f32,
                                   n_rays: u64, t_out: *mut f32, idx_out: *mut i32, normal_out: *mut f32) -> i32;
    pub fn rtb_spheres_bounds(spheres_xyzr: *const f32, n: u32, boxes6_out: *mut f32);
    pub fn rtb_gen_random_spheres(seed: u64, spheres_xyzr_out: *mut f32, colors_out: *mut f32) -> u32;
    pub fn rtb_gen_random_triangles(seed: u64, verts_out: *mut f32, colors_out: *mut f32) -> u32;
    pub fn rtb_free(p: *mut c_void);
    pub fn rtb_fill_samples(seed: u64, n_pairs: u32, out: *mut f32);
    pub fn rtb_write_png(path: *const c_char, rgb8: *const u8, width: u32, height: u32) -> i32;
    pub fn rtb_last_error() -> *const c_char;
    pub fn rtb_version() -> *const c_char;
}
