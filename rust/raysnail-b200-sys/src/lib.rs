//! `extern "C"` declarations of `include/raysnail_b200.h` (ABI version 1): the drop-in boundary
//! for `TakePhotoSettings::shot_to_target` (raysnail `src/camera.rs:261-287`).
//!
//! Untested source (no Rust toolchain where this repository is built).  Field names, order and
//! types are compared with the C header by `tests/test_rust_binding.py`; call [`check_abi`] once
//! at start-up to compare the struct sizes with the ones the library was compiled with.
#![allow(non_camel_case_types)]

use std::os::raw::{c_char, c_int, c_void};

pub const RS_ABI_VERSION: c_int = 3;

// rs_status
pub const RS_OK: c_int = 0;
pub const RS_ERR_INVALID_ARGUMENT: c_int = 1;
pub const RS_ERR_CUDA: c_int = 2;
pub const RS_ERR_NO_DEVICE: c_int = 3;
pub const RS_ERR_NO_SCENE: c_int = 4;
pub const RS_ERR_UNSUPPORTED: c_int = 5;
pub const RS_ERR_OUT_OF_MEMORY: c_int = 6;
pub const RS_ERR_PARSE: c_int = 7;
pub const RS_ERR_IO: c_int = 8;

// rs_node_type (one per `Hittable` implementor, src/hittable/)
pub const RS_NODE_SPHERE: i32 = 0;
pub const RS_NODE_BOX: i32 = 1;
pub const RS_NODE_RECT: i32 = 2;
pub const RS_NODE_QUADRIC: i32 = 3;
pub const RS_NODE_DIFFERENCE: i32 = 4;
pub const RS_NODE_INTERSECTION: i32 = 5;
pub const RS_NODE_TRANSFORM: i32 = 6;
pub const RS_NODE_MEDIUM: i32 = 7;
pub const RS_NODE_RAYMARCHER: i32 = 8;
pub const RS_NODE_BVH: i32 = 9;

pub const RS_RECT_XY: i32 = 0;
pub const RS_RECT_XZ: i32 = 1;
pub const RS_RECT_YZ: i32 = 2;

// rs_material_type (src/material/)
pub const RS_MAT_LAMBERTIAN: i32 = 0;
pub const RS_MAT_METAL: i32 = 1;
pub const RS_MAT_DIFFUSE_METAL: i32 = 2;
pub const RS_MAT_DIELECTRIC: i32 = 3;
pub const RS_MAT_DIFFUSE_LIGHT: i32 = 4;
pub const RS_MAT_ISOTROPIC: i32 = 5;
pub const RS_MAT_BLINN_PHONG: i32 = 6;
pub const RS_MAT_MIXED: i32 = 7;

// rs_texture_type (src/texture/)
pub const RS_TEX_SOLID: i32 = 0;
pub const RS_TEX_CHECKER: i32 = 1;
pub const RS_TEX_IMAGE: i32 = 2;
pub const RS_TEX_PERLIN: i32 = 3;

pub const RS_FLAG_PIXEL_CENTRE: u32 = 1;
pub const RS_FLAG_KERNEL_TIMING: u32 = 2;
pub const RS_FLAG_MIS: u32 = 4;
pub const RS_REDO_REFERENCE: i32 = 0;
pub const RS_REDO_NOISE: i32 = 1;
pub const RS_REDO_NOISE_FIXED: i32 = 2;
pub const RS_RNG_PIXEL: u32 = 0;
pub const RS_RNG_LENS: u32 = 1;
pub const RS_RNG_SCATTER: u32 = 2;
pub const RS_RNG_MIXED: u32 = 3;
pub const RS_RNG_LIGHT: u32 = 4;
pub const RS_RNG_BRDF: u32 = 5;
pub const RS_RNG_MEDIUM: u32 = 6;
pub const RS_OPT_BVH_WIDTH: c_int = 1;
pub const RS_OPT_WAVE_STREAMS: c_int = 2;
pub const RS_OPT_LIST_MAX: c_int = 3;
pub const RS_OPT_WAIT_BUDGET_MS: c_int = 4;
pub const RS_OPT_INJECT_FAULT: c_int = 100;
pub const RS_IPC_HANDLE_BYTES: usize = 64;
pub const RS_MAX_PEERS: usize = 16;
pub const RS_STATIC_MAX_PIXELS: usize = 4194304;

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct rs_node {
    pub type_: i32,
    pub material: i32,
    pub child0: i32,
    pub child1: i32,
    pub first_xform: i32,
    pub n_xform: i32,
    pub rect_axis: i32,
    pub reserved: i32,
    pub p: [f64; 10],
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct rs_xform {
    pub m: [f64; 16],
    pub inv: [f64; 16],
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct rs_material {
    pub type_: i32,
    pub texture: i32,
    pub mat1: i32,
    pub mat2: i32,
    pub has_curve: i32,
    pub phong_exponent: i32,
    pub phong_factor: f64,
    pub a: f64,
    pub b: f64,
    pub color: [f32; 4],
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct rs_texture {
    pub type_: i32,
    pub odd: i32,
    pub even: i32,
    pub width: i32,
    pub height: i32,
    pub texel_offset: i32,
    pub perlin_kind: i32,
    pub perlin_depth: i32,
    pub perlin_smooth: i32,
    pub perlin_vector: i32,
    pub point_count: i32,
    pub perm_offset: i32,
    pub value_offset: i32,
    pub reserved: i32,
    pub scale: f64,
    pub color: [f32; 4],
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct rs_scene_desc {
    pub nodes: *const rs_node,
    pub n_nodes: i32,
    pub roots: *const i32,
    pub root_bbox: *const f64,
    pub n_roots: i32,
    pub lights: *const i32,
    pub n_lights: i32,
    pub xforms: *const rs_xform,
    pub n_xforms: i32,
    pub tri_verts: *const f64,
    pub tri_normals: *const f64,
    pub tri_material: *const i32,
    pub n_tris: i32,
    pub materials: *const rs_material,
    pub n_materials: i32,
    pub textures: *const rs_texture,
    pub n_textures: i32,
    pub texels: *const u8,
    pub n_texel_bytes: i64,
    pub perlin_perm: *const i32,
    pub n_perlin_perm: i32,
    pub perlin_values: *const f64,
    pub n_perlin_values: i32,
    pub default_material: i32,
    pub bg0: [f32; 4],
    pub bg1: [f32; 4],
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct rs_camera {
    pub origin: [f64; 3],
    pub lb: [f64; 3],
    pub horizontal_full: [f64; 3],
    pub vertical_full: [f64; 3],
    pub horizontal_unit: [f64; 3],
    pub vertical_unit: [f64; 3],
    pub aperture: f64,
    pub shutter_speed: f64,
    pub width: i32,
    pub height: i32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct rs_render_params {
    pub sqrt_spp: i32,
    pub depth: i32,
    pub gamma: i32,
    pub stratum_row_begin: i32,
    pub stratum_row_end: i32,
    pub max_wave_paths: i32,
    pub seed: u64,
    pub flags: u32,
    pub reserved: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct rs_stats {
    pub n_samples: u64,
    pub n_segments: u64,
    pub n_launches: u64,
    pub ms_raygen: f64,
    pub ms_extend: f64,
    pub ms_shade: f64,
    pub ms_accumulate: f64,
    pub ms_total: f64,
}

/// Opaque handle of `rs_create_multi` (several devices behind one call).
#[repr(C)]
pub struct rs_multi {
    _private: [u8; 0],
}

/// Opaque context: one CUDA device, one host thread at a time.
#[repr(C)]
pub struct rs_ctx {
    _private: [u8; 0],
}

extern "C" {
    pub fn rs_abi_version() -> c_int;
    pub fn rs_abi_struct_size(which: c_int) -> c_int;
    pub fn rs_create(out: *mut *mut rs_ctx, device: c_int) -> c_int;
    pub fn rs_destroy(ctx: *mut rs_ctx);
    pub fn rs_last_error(ctx: *const rs_ctx) -> *const c_char;
    pub fn rs_set_option(ctx: *mut rs_ctx, option: c_int, value: c_int) -> c_int;
    pub fn rs_upload_scene(ctx: *mut rs_ctx, scene: *const rs_scene_desc) -> c_int;
    pub fn rs_render(ctx: *mut rs_ctx, cam: *const rs_camera, params: *const rs_render_params, pixel_mask: *const u8, out_rgba: *mut f32) -> c_int;
    pub fn rs_render_accumulate(ctx: *mut rs_ctx, cam: *const rs_camera, params: *const rs_render_params, d_pixel_mask: *const u8, d_accum: *mut f32, stream: *mut c_void) -> c_int;
    pub fn rs_finalize(ctx: *mut rs_ctx, d_accum: *const f32, d_pixel_mask: *const u8, d_out_rgba: *mut f32, width: i32, height: i32, n_samples: i32, gamma: i32, stream: *mut c_void) -> c_int;
    pub fn rs_device_alloc(ctx: *mut rs_ctx, bytes: i64, d_ptr: *mut *mut c_void) -> c_int;
    pub fn rs_device_free(ctx: *mut rs_ctx, d_ptr: *mut c_void) -> c_int;
    pub fn rs_ipc_export(ctx: *mut rs_ctx, d_ptr: *const c_void, handle: *mut u8) -> c_int;
    pub fn rs_ipc_open(ctx: *mut rs_ctx, handle: *const u8, d_ptr: *mut *mut c_void) -> c_int;
    pub fn rs_ipc_close(ctx: *mut rs_ctx, d_ptr: *mut c_void) -> c_int;
    pub fn rs_finalize_peers(ctx: *mut rs_ctx, d_accums: *const *const f32, n_peers: i32, d_pixel_mask: *const u8, d_out_rgba: *mut f32, pixel_begin: i64, pixel_end: i64, n_samples: i32, gamma: i32, stream: *mut c_void) -> c_int;
    pub fn rs_combine_passes(ctx: *mut rs_ctx, d_old: *const f32, d_new: *const f32, d_out: *mut f32, n_pixels: i64, pass: f32, stream: *mut c_void) -> c_int;
    pub fn rs_noise_map(ctx: *mut rs_ctx, d_pixels: *const f32, d_noise: *mut f32, width: i32, height: i32, quirk: i32, stream: *mut c_void) -> c_int;
    /// No counterpart in the reference: the redo map computed without any device allocation (fixed 16 MiB scratch
    /// inside the library, at most RS_STATIC_MAX_PIXELS pixels; calls that overlap on one GPU race on it).
    pub fn rs_redo_map_static(ctx: *mut rs_ctx, d_pixels: *const f32, d_redo_mask: *mut u8, d_redo_count: *mut u64, width: i32, height: i32, quirk: i32, noise_threshold: f32, stream: *mut c_void) -> c_int;
    pub fn rs_quantize_rgb8(ctx: *mut rs_ctx, d_pixels: *const f32, d_rgb: *mut u8, n_pixels: i64, stream: *mut c_void) -> c_int;
    pub fn rs_debug_primary(ctx: *mut rs_ctx, cam: *const rs_camera, params: *const rs_render_params, sample_index: i32, obj_id: *mut i32, t: *mut f64, normal3: *mut f64, point3: *mut f64, front: *mut i32, material: *mut i32, ray_od6: *mut f64) -> c_int;
    pub fn rs_debug_scatter(ctx: *mut rs_ctx, cam: *const rs_camera, params: *const rs_render_params, sample_index: i32, kind: *mut i32, next_od6: *mut f64, weight3: *mut f64, emitted3: *mut f64) -> c_int;
    pub fn rs_get_stats(ctx: *const rs_ctx, out: *mut rs_stats) -> c_int;
    pub fn rs_render_passes(ctx: *mut rs_ctx, cam: *const rs_camera, params: *const rs_render_params, n_passes: i32, redo_mode: i32, noise_threshold: f32, out_rgba: *mut f32, out_rgb8: *mut u8, redo_counts: *mut i64) -> c_int;
    pub fn rs_pass_seed(seed: u64, pass: i32) -> u64;
    pub fn rs_philox4x32_10(counter: *const u32, key: *const u32, out: *mut u32) -> c_int;
    pub fn rs_philox_uniform(seed: u64, pixel: u32, sample: u32, depth: u32, purpose: u32, call: u32, out4: *mut f64) -> c_int;
    pub fn rs_create_multi(out: *mut *mut rs_multi, device_ids: *const i32, n: i32) -> c_int;
    pub fn rs_destroy_multi(m: *mut rs_multi);
    pub fn rs_multi_last_error(m: *const rs_multi) -> *const c_char;
    pub fn rs_multi_device_count(m: *const rs_multi) -> c_int;
    pub fn rs_multi_set_option(m: *mut rs_multi, option: c_int, value: c_int) -> c_int;
    pub fn rs_multi_upload_scene(m: *mut rs_multi, scene: *const rs_scene_desc) -> c_int;
    pub fn rs_multi_render(m: *mut rs_multi, cam: *const rs_camera, params: *const rs_render_params, pixel_mask: *const u8, out_rgba: *mut f32, stats_sum: *mut rs_stats) -> c_int;
    pub fn rs_signal(ctx: *mut rs_ctx, d_counter: *mut u32, stream: *mut c_void) -> c_int;
    pub fn rs_wait_counter(ctx: *mut rs_ctx, d_counter: *const u32, expect: u32, stream: *mut c_void) -> c_int;
    pub fn rs_wait_check(ctx: *mut rs_ctx, stream: *mut c_void) -> c_int;
}

/// Compares `size_of` of every `#[repr(C)]` struct with what the loaded library was compiled
/// with (`rs_abi_struct_size`) and the ABI version.  Call once before anything else.
pub fn check_abi() -> Result<(), String> {
    use std::mem::size_of;
    let v = unsafe { rs_abi_version() };
    if v != RS_ABI_VERSION {
        return Err(format!("libraysnail_b200 ABI version {} != {}", v, RS_ABI_VERSION));
    }
    let sizes: [(c_int, usize, &str); 8] = [
        (0, size_of::<rs_node>(), "rs_node"),
        (1, size_of::<rs_xform>(), "rs_xform"),
        (2, size_of::<rs_material>(), "rs_material"),
        (3, size_of::<rs_texture>(), "rs_texture"),
        (4, size_of::<rs_scene_desc>(), "rs_scene_desc"),
        (5, size_of::<rs_camera>(), "rs_camera"),
        (6, size_of::<rs_render_params>(), "rs_render_params"),
        (7, size_of::<rs_stats>(), "rs_stats"),
    ];
    for (which, mine, name) in sizes.iter() {
        let theirs = unsafe { rs_abi_struct_size(*which) };
        if theirs as usize != *mine {
            return Err(format!("{}: {} bytes here, {} in the library", name, mine, theirs));
        }
    }
    Ok(())
}
