//! FFI binding of librrt_b200.so (include/rrt.h) for conor722/rust-ray-tracer.
//! NOT COMPILED IN THIS REPOSITORY'S CI: the build image has no cargo/rustc. Drop this file in
//! `src/rrt_ffi.rs`, add `mod rrt_ffi;` to `src/main.rs`, and use `build.rs` from this directory.
#![allow(non_camel_case_types, dead_code)]
use std::os::raw::{c_char, c_int, c_void};

pub const RRT_MAX_LIGHTS: usize = 16;

#[repr(C)]
pub struct rrt_texture { pub width: i32, pub height: i32, pub rgb: *const u8 }

#[repr(C)]
#[derive(Clone, Copy)]
pub struct rrt_material {
    pub ka: [f64; 3], pub kd: [f64; 3], pub ks: [f64; 3],
    pub ns: f64, pub kr: f64, pub tex_id: i32, pub bump_id: i32,
}

#[repr(C)]
pub struct rrt_scene_desc {
    pub n_tris: u64,
    pub v: *const f64, pub vt: *const f64, pub vn: *const f64,
    pub material_id: *const i32,
    pub n_materials: u32, pub materials: *const rrt_material,
    pub n_textures: u32, pub textures: *const rrt_texture,
    pub root_min: [f64; 3], pub root_max: [f64; 3],
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct rrt_light { pub kind: i32, pub _pad: i32, pub intensity: f64, pub vec: [f64; 3] }

#[repr(C)]
pub struct rrt_render_params {
    pub width: i32, pub height: i32,
    pub origin: [f64; 3], pub viewport: [f64; 3],
    pub n_lights: i32, pub max_depth: i32,
    pub lights: [rrt_light; RRT_MAX_LIGHTS],
    pub surface_offset: f64,
    pub use_camera_rot: i32, pub _pad: i32,
    pub camera_rot: [f64; 9],
}

#[repr(C)] pub struct rrt_scene { _private: [u8; 0] }

#[link(name = "rrt_b200")]
extern "C" {
    pub fn rrt_last_error() -> *const c_char;
    pub fn rrt_render_params_default(p: *mut rrt_render_params);
    pub fn rrt_scene_create(desc: *const rrt_scene_desc, device: c_int, out: *mut *mut rrt_scene) -> c_int;
    pub fn rrt_scene_destroy(scene: *mut rrt_scene);
    pub fn rrt_scene_order_hint(scene: *mut rrt_scene, enable: c_int) -> c_int;
    /// optional, once per frame size: allocate at set-up what the first rrt_render of that size would allocate inside the call
    pub fn rrt_scene_reserve(scene: *mut rrt_scene, params: *const rrt_render_params, shard: *const rrt_shard,
                             with_device_frame: c_int) -> c_int;
    pub fn rrt_render(scene: *mut rrt_scene, params: *const rrt_render_params, fb_0rgb: *mut u32,
                      hit_ids: *mut i32, device_ms: *mut f32) -> c_int;
    pub fn rrt_render_frames(scene: *mut rrt_scene, params: *const rrt_render_params, n_frames: i32,
                             fbs: *const *mut u32, total_ms: *mut f32) -> c_int;
    pub fn rrt_render_device(scene: *mut rrt_scene, params: *const rrt_render_params, shard: *const rrt_shard,
                             d_fb: *mut u32, d_hit_ids: *mut i32, cuda_stream: *mut c_void) -> c_int;
    // multi-GPU: fused render + gather into a (possibly peer-mapped) full frame, include/rrt.h
    pub fn rrt_render_device_into_frame(scene: *mut rrt_scene, params: *const rrt_render_params,
                                        shard: *const rrt_shard, d_frame: *mut u32, cuda_stream: *mut c_void) -> c_int;
    pub fn rrt_frame_alloc(device: c_int, width: i32, height: i32, d_frame: *mut *mut u32) -> c_int;
    pub fn rrt_frame_free(device: c_int, d_frame: *mut u32) -> c_int;
    pub fn rrt_ipc_get_handle(device: c_int, d_ptr: *const c_void, handle: *mut u8 /* [64] */) -> c_int;
    pub fn rrt_ipc_open_handle(device: c_int, handle: *const u8 /* [64] */, d_ptr: *mut *mut c_void) -> c_int;
    pub fn rrt_ipc_close_handle(device: c_int, d_ptr: *mut c_void) -> c_int;
    pub fn rrt_copy_frame_to_host(device: c_int, d_frame: *const u32, width: i32, height: i32, fb_0rgb: *mut u32,
                                  cuda_stream: *mut c_void) -> c_int;
}

/// rrt_shard (include/rrt.h): tile row j of `tile_rows` scan-lines belongs to shard j % n_shards.
#[repr(C)] pub struct rrt_shard { pub tile_rows: i32, pub n_shards: i32, pub shard_index: i32 }

pub fn last_error() -> String {
    unsafe { std::ffi::CStr::from_ptr(rrt_last_error()).to_string_lossy().into_owned() }
}
